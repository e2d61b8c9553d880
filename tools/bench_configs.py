"""Measures the BASELINE.json configs other than bench.py's headline one (development / profiles
helper; bench.py is the driver contract).  One JSON line per case:
  config 1 shape (Adirondack 1436x992, D=145, block 17, SSD, smooth 1.0 and 0.9)
  config 3 (1920x1080, D=256, block 5/9/15 x SSD/SAD/NCC)
  config 5 (1280x720 pairs, D=128, 7x7 SAD + sub-pixel, batch API with pinned buffers -> pairs/s)
  reconstruction chain (disparity -> depth -> vertices -> faces) on a 4K map
Device-resident timings use CUDA events on the context stream (median of --iters after 3 warm-ups)."""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import stereo_reconstruction_b200 as s
from stereo_reconstruction_b200 import capi, synth

ap = argparse.ArgumentParser()
ap.add_argument("--iters", type=int, default=10)
ap.add_argument("--batch", type=int, default=48, help="pairs of the config-5 sample (the config names 1024)")
ap.add_argument("--skip", default="", help="comma list of: c1,c3,c5,recon")
a = ap.parse_args()
skip = set(a.skip.split(","))
ctx = capi.default_context(0)
COST = {"ssd": capi.COST_SSD_L2, "sad": capi.COST_SAD, "ncc": capi.COST_NCC}


def timed(fn):
    for _ in range(3):
        fn()
    ctx.sync()
    ts = []
    for _ in range(a.iters):
        ctx.l2_flush()
        ctx.timer_start(); fn(); ts.append(ctx.timer_stop())
    return float(np.median(ts))


def search_case(tag, w, h, D, block, cost, smooth=1.0, subpixel=False, lr_check=False):
    L, R = synth.synth_pair(0xB200, 0, w, h, 0, D)
    pair = capi.Pair(ctx, w, h, w, h)
    pair.upload(L, R)
    p = capi.search_params(block=block, min_disp=0, max_disp=D, smooth=smooth, cost=COST[cost], subpixel=subpixel, lr_check=lr_check)
    ms = timed(lambda: pair.run(p))
    ev = synth.effective_evals(w, w, h, block, 0, D)
    print(json.dumps({"case": tag, "w": w, "h": h, "D": D, "block": block, "cost": cost, "smooth": smooth,
                      "subpixel": subpixel, "lr_check": lr_check, "kernel_family": ctx.last_kernel, "ms": round(ms, 4),
                      "effective_MDE_per_s": round(ev / ms / 1e3, 1), "nominal_MDE_per_s": round(w * h * D / ms / 1e3, 1)}), flush=True)
    pair.close()


if "c1" not in skip:
    for sm in (1.0, 0.9):
        search_case("config1-shape", 1436, 992, 145, 17, "ssd", smooth=sm)
    search_case("config2-shape-lr", 1436, 992, 145, 17, "ssd", lr_check=True)
    search_case("config2-shape-lr-subpix", 1436, 992, 145, 17, "sad", lr_check=True, subpixel=True)
    # main.cpp:40-65 as one device pipeline on the config-1/2 shape: search (block 17, smooth 0.9) -> 8-bit hand-off ->
    # outlier removal (k = 500) -> depth -> vertices -> faces; the disparity map never leaves HBM (outputs NULL = timing
    # of the kernels only, n_faces is the only D2H)
    import ctypes as C
    w, h, D = 1436, 992, 145
    L, R = synth.synth_pair(0xB200, 0, w, h, 0, D)
    pair = capi.Pair(ctx, w, h, w, h)
    pair.upload(L, R)
    p = capi.search_params(block=17, min_disp=0, max_disp=D, smooth=0.9)
    nf = C.c_uint32(0)
    null = C.c_void_p(None)

    def pipeline():
        pair.run(p)
        ctx.check(ctx._l.sr_pair_reconstruct(ctx.h, pair.h, 1, 500, C.c_float(1.5), C.c_float(0.8), C.c_float(2000.0), C.c_float(1.0),
                                             C.c_float(2000.0), C.c_float(2000.0), C.c_float(w / 2), C.c_float(h / 2), C.c_float(1.0),
                                             null, null, null, null, C.byref(nf)))
    for _ in range(2):
        pipeline()
    ctx.sync()
    t0 = time.perf_counter()
    for _ in range(a.iters):
        pipeline()
    ctx.sync()
    dt = (time.perf_counter() - t0) / a.iters
    print(json.dumps({"case": "main.cpp-pipeline-resident", "w": w, "h": h, "D": D, "block": 17, "smooth": 0.9,
                      "ms": round(dt * 1e3, 4), "faces": int(nf.value),
                      "note": "search + u8 hand-off + removeDisparityOutliers(500) + depth + back-projection + face list, host-timed incl. the n_faces read"}), flush=True)
    pair.close()
if "c3" not in skip:
    for block in (5, 9, 15):
        for cost in ("ssd", "sad", "ncc"):
            search_case("config3", 1920, 1080, 256, block, cost)
if "c5" not in skip:
    w, h, D, n = 1280, 720, 128, a.batch
    search_case("config5-one-pair-resident", w, h, D, 7, "sad", subpixel=True)
    # distinct pinned buffers for a ring of 8 pairs, reused round-robin over the batch (content differs per slot)
    ring = 8
    Ls, Rs, outs, hs = [], [], [], []
    for i in range(ring):
        L, R = synth.synth_pair(0xB200, i, w, h, 0, D)
        Lp, lh = capi.pinned_empty(L.shape, np.uint8); Lp[...] = L
        Rp, rh = capi.pinned_empty(R.shape, np.uint8); Rp[...] = R
        o, oh = capi.pinned_empty((h, w), np.float64)
        Ls.append(Lp); Rs.append(Rp); outs.append(o); hs += [lh, rh, oh]
    p = capi.search_params(block=7, min_disp=0, max_disp=D, smooth=1.0, cost=capi.COST_SAD, subpixel=True)
    lb = [Ls[i % ring] for i in range(n)]; rb = [Rs[i % ring] for i in range(n)]; ob = [outs[i % ring] for i in range(n)]
    capi.block_search_batch(ctx, lb, rb, p, ob)            # the first call builds the ring of resident pairs
    t0 = time.perf_counter()
    capi.block_search_batch(ctx, lb, rb, p, ob)
    dt = time.perf_counter() - t0
    ev = synth.effective_evals(w, w, h, 7, 0, D)
    print(json.dumps({"case": "config5-batch-e2e", "pairs": n, "w": w, "h": h, "D": D, "block": 7, "cost": "sad", "subpixel": True,
                      "seconds": round(dt, 4), "pairs_per_s": round(n / dt, 1), "effective_MDE_per_s": round(n * ev / dt / 1e6, 1),
                      "h2d_bytes_per_pair": 2 * w * h * 3, "d2h_bytes_per_pair": w * h * 8}), flush=True)
    for x in hs:
        capi.free_pinned(x)
if "secondary" not in skip:
    # the reference-exact secondary paths (round-1 review: unmeasured): Left variant, varBlock, block > 17 (direct kernels,
    # O(window) per candidate like the reference), LinearSearch; and the reference's default call end to end
    def variant_case(tag, L, R, params, note):
        pair = capi.Pair(ctx, L.shape[1], L.shape[0], R.shape[1], R.shape[0])
        pair.upload(L, R)
        ms = timed(lambda: pair.run(params))
        w2, h2 = R.shape[1], R.shape[0]
        D = params.max_disp - params.min_disp
        print(json.dumps({"case": tag, "w": w2, "h": h2, "D": D, "block": params.block, "kernel_family": ctx.last_kernel, "ms": round(ms, 4),
                          "nominal_MDE_per_s": round(w2 * h2 * D / ms / 1e3, 1), "note": note}), flush=True)
        pair.close()
    L, R = synth.synth_pair(0xB200, 0, 900, 750, 0, 128)
    variant_case("left-variant", L, R, capi.search_params(variant=capi.VARIANT_LEFT, block=17, min_disp=0, max_disp=128, smooth=0.9),
                 "computeDisparityMapLeft(0.9), Teddy-sized: unbiased winners from the tile kernel + multi-CTA row wavefront")
    variant_case("varblock", L, R, capi.search_params(block=17, min_disp=0, max_disp=128, smooth=0.9, var_block=True, thres=19.0),
                 "computeDisparityMapRight(0.9, varBlock = true, 19.0): per-pixel window growth, direct kernel")
    variant_case("block-21-direct", L, R, capi.search_params(block=21, min_disp=0, max_disp=128, smooth=1.0),
                 "windows beyond 17: direct kernel, O(window) per candidate")
    variant_case("block-17-tile", L, R, capi.search_params(block=17, min_disp=0, max_disp=128, smooth=1.0), "the same shape through the tile kernel, for scale")
    variant_case("linear-search", L, R, capi.search_params(variant=capi.VARIANT_LINEAR, block=1, min_disp=0, max_disp=200, smooth=1.0),
                 "LinearSearch::computeDisparityMap(1.0): 200 candidates, one-pixel window")
    # main.cpp:40 end to end from and to pinned host buffers: BlockSearch(17, 0, 145).computeDisparityMapRight(0.9)
    w, h, D = 1436, 992, 145
    L, R = synth.synth_pair(0xB200, 0, w, h, 0, D)
    Lp, a1 = capi.pinned_empty(L.shape, np.uint8); Lp[...] = L
    Rp, a2 = capi.pinned_empty(R.shape, np.uint8); Rp[...] = R
    o, a3 = capi.pinned_empty((h, w), np.float64)
    for smooth in (0.9, 1.0):
        p = capi.search_params(block=17, min_disp=0, max_disp=D, smooth=smooth)
        for _ in range(3):
            capi.block_search(ctx, Lp, Rp, p, out=o)
        t0 = time.perf_counter()
        for _ in range(a.iters):
            capi.block_search(ctx, Lp, Rp, p, out=o)
        dt = (time.perf_counter() - t0) / a.iters
        print(json.dumps({"case": "config1-shape-e2e", "w": w, "h": h, "D": D, "block": 17, "smooth": smooth, "ms": round(dt * 1e3, 4),
                          "note": "sr_block_search from and to pinned host buffers (H2D 8.5 MB + D2H 11.4 MB inside)"}), flush=True)
    for x in (a1, a2, a3):
        capi.free_pinned(x)
if "recon" not in skip:
    from stereo_reconstruction_b200 import reconstruction as rec
    w, h = 3840, 2160
    rng = np.random.default_rng(1)
    disp = rng.integers(1, 255, (h, w)).astype(np.float32)
    bgr = rng.integers(0, 255, (h, w, 3), dtype=np.uint8)
    K = np.array([[3000.0, 0, w / 2], [0, 3000.0, h / 2], [0, 0, 1]], np.float32)
    t0 = time.perf_counter()
    for _ in range(3):
        rec.reconstructFused(bgr, disp.copy(), K, 3000.0, 1.0, kernelSize=500)
    dt = (time.perf_counter() - t0) / 3
    print(json.dumps({"case": "reconstruction-chain-host-call", "w": w, "h": h, "seconds_per_call": round(dt, 4),
                      "note": "outliers(k=500) + depth + vertices + faces incl. H2D/D2H of pageable numpy buffers"}), flush=True)
