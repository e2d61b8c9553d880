#!/usr/bin/env python
"""bench.py -- headline benchmark of the stereo hot path (BASELINE.json metric: million disparity
evaluations per second, MDE/s).

A step = one pass of the hot path (block-matching search, cost + WTA + resolve to the CV_64F map)
over one synthetic rectified pair.  Default workload = BASELINE.json configs[3]:
3840x2160, 512 disparities, 9x9 block, SAD; at N>1 the pair is row-band sharded and the bands are
gathered to rank 0 over NCCL ("scaling": "strong").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload 4k512|1080p256|720p128|adirondack]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...      # the reference CPU algorithm (oracle port) on host cores

Prints ONE JSON line on rank 0.  `value` = effective evaluations (reference semantics: candidates
before the break rule) of the whole pair / device time per step with inputs resident in HBM;
`e2e` = the same through the public host-pointer API (pinned host buffers, H2D + D2H inside).
At N=1 the line also carries `roofline_recon`: the HBM-bound Reconstruction kernels (main.cpp:41-65) on the same
pair, timed after every other timed region (secondary; SR_BENCH_NO_RECON=1 skips it).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (w, h, max_disp, block, cost)   -- BASELINE.json configs
    "4k512": (3840, 2160, 512, 9, "sad"),       # configs[3]
    "1080p256": (1920, 1080, 256, 9, "ssd"),    # configs[2] (one point of the sweep)
    "720p128": (1280, 720, 128, 7, "sad"),      # configs[4] (one pair)
    "adirondack": (1436, 992, 145, 17, "ssd"),  # configs[0] shape (synthetic content)
}
# algorithmic integer ops per evaluation: pointwise cost + 2 sliding sums + WTA
OPS_PER_EVAL = {"ssd": 8, "sad": 7}
# dram__bytes_read.sum + dram__bytes_write.sum of the search kernel from the committed ncu capture
# (profiles/), per launch, keyed by workload; None = not captured for this workload
NCU_TRAFFIC = {"4k512": 105734144 + 12579584}   # profiles/r02_tile_sad_4k512_ncu_full.csv (read + write bytes, N=1 launch)


def host_threads():
    """All host threads this process may use.  torch.distributed.run exports OMP_NUM_THREADS=1 when it starts more than
    one rank: the CPU arms pass an explicit thread count to the oracle instead of inheriting that."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


def config_for(args, world):
    """`config` of the JSON line -- identical in both arms (--impl b200 / reference) for the same flags."""
    if args.workload == "batch720":
        return {"workload": f"batch720: {args.batch_pairs} synthetic 1280x720 pairs per rank and step, 128 disparities, 7x7 SAD + "
                            "sub-pixel (BASELINE.json configs[4]; the config names 1024 pairs in total)",
                "sharding": "by pair, p mod N, no collective", "evals": "effective (reference break rule)"}
    w, h, D, block, cost = WORKLOADS[args.workload]
    cost = args.cost or cost
    block = args.block or block
    return {"workload": f"{args.workload}: synthetic {w}x{h} rectified pair, {D} disparities, {block}x{block} {cost.upper()}, "
                        "smooth 1.0 (BASELINE.json configs)",
            "sharding": "single GPU" if world == 1 else "row bands with halo rows, bands gathered to rank 0",
            "l2": "384 MB written between timed iterations", "evals": "effective (reference break rule)",
            "nominal_MDE": w * h * D / 1e6}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="4k512", choices=sorted(WORKLOADS) + ["batch720"])
    ap.add_argument("--batch-pairs", type=int, default=32, help="pairs per rank and step of --workload batch720")
    ap.add_argument("--cost", default=None, choices=["ssd", "sad"])
    ap.add_argument("--block", type=int, default=None)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    return ap.parse_args()


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.lines = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "20"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def mark(self, timeout=3.0):
        """Waits until nvidia-smi delivers its first sample (its start-up takes longer than a short timed region),
        then marks the start of the region: only samples from here on are reported."""
        t0 = time.time()
        while self.proc and not self.lines and time.time() - t0 < timeout:
            time.sleep(0.005)
        self.start = len(self.lines)

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        start = getattr(self, "start", 0)
        for ln in (self.lines[start:] or self.lines[-1:]):   # a region shorter than the 20 ms period: the sample just before it
            f = [t.strip() for t in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def _oracle_sample(orc, L, R, block, D, ocost, mid, rows, nthreads):
    """Runs the CPU port on `rows` output rows starting at `mid`, on a crop that carries the
    window halo (so the port's per-call setup is proportional to the sample, not the image)."""
    pad = block
    y0 = max(mid - pad, 0)
    y1 = min(mid + rows + pad, L.shape[0])
    t0 = time.perf_counter()
    orc.block_search_right_fast(L[y0:y1], R[y0:y1], block, 0, D, 1.0, cost=ocost, rows=(mid - y0, mid - y0 + rows), nthreads=nthreads)
    return time.perf_counter() - t0


def _ref_sample(ref, L, R, block, D, mid, rows, nthreads):
    """The reference's own BlockSearch.cpp (oracle/_ref, compiled unchanged) on `rows` output rows: the rows are cut into
    one band per thread, each band is a BlockSearch of its own on a crop with halo rows.  -> (seconds, threads used)"""
    pad = block
    y0 = max(mid - pad, 0)
    y1 = min(mid + rows + pad, L.shape[0])
    t0 = time.perf_counter()
    _, used = ref.block_search_right_rows(L[y0:y1], R[y0:y1], block, 0, D, (mid - y0, mid - y0 + rows), nthreads)
    return time.perf_counter() - t0, used


def _ref_available():
    try:
        from oracle import ref
        if ref.available():
            ref.lib()
            return ref
    except Exception as e:   # a missing .so on a box without the reference tree: fall back to the port, say so
        print(f"[bench] oracle/_ref unavailable ({e}); timing the oracle port", file=sys.stderr, flush=True)
    return None


def port_baseline(orc, L, R, w, h, D, block, cost, seconds, nthreads):
    """Optimised CPU port (box sums + OpenMP, bit-identical to the reference) on a bounded row sample."""
    from stereo_reconstruction_b200 import synth
    ocost = orc.COST_SAD if cost == "sad" else orc.COST_SSD_L2
    mid = block
    unit = 16 * nthreads                                          # the port parallelises over 16-row chunks
    cal = min(unit, h - 2 * block)
    _oracle_sample(orc, L, R, block, D, ocost, mid, 2, nthreads)
    per_row = _oracle_sample(orc, L, R, block, D, ocost, mid, cal, nthreads) / cal
    rows = int(seconds / max(per_row, 1e-6))
    rows = max(min(cal, h - 2 * block), min(h - 2 * block, (rows // unit) * unit))
    sec = _oracle_sample(orc, L, R, block, D, ocost, mid, rows, nthreads)
    ev = synth.effective_evals(w, w, rows, block, 0, D)
    return {"value": ev / sec / 1e6, "unit": "MDE/s", "cores": nthreads, "kind": "port",
            "sample": f"rows {mid}..{mid + rows} of the same pair ({rows}/{h} rows, {sec:.2f} s), {cost.upper()} {block}x{block}: box-sum + "
                      f"OpenMP port of BlockSearch::computeDisparityMapRight (bit-identical to the reference), {nthreads} threads"}


def run_reference(args, rank, world):
    """The reference's own CPU implementation of the path on the box's host cores: oracle/_ref = the reference's
    BlockSearch.cpp compiled UNCHANGED (header stand-ins for OpenCV/Eigen, oracle/Makefile), one BlockSearch per row band,
    bands on all host threads (the reference itself has no threads).  Its only cost is SSD-L2 (BlockSearch.cpp:156-158),
    so that is what is timed, on the workload's shape, window and disparity range.  Falls back to the oracle port when the
    prebuilt library is missing.  Rank 0 only; the other ranks exit without work."""
    if rank != 0:
        return
    nthreads = host_threads()
    os.environ["OMP_NUM_THREADS"] = str(nthreads)                 # before the OpenMP runtimes load (torchrun exports 1)
    from oracle import orc
    from stereo_reconstruction_b200 import synth
    w, h, D, block, cost = WORKLOADS["720p128" if args.workload == "batch720" else args.workload]
    cost = args.cost or cost
    block = args.block or block
    L, R = orc.synth_pair(0xB200, 0, w, h, 0, D)
    mid = block
    budget = min(3.0, 150.0 / max(args.steps + args.warmup, 1))   # seconds per step: the whole run ends within a few minutes
    ref = _ref_available()
    if ref is not None:
        sec1, used = _ref_sample(ref, L, R, block, D, mid, nthreads, nthreads)          # one row per thread
        rows = max(nthreads, min(h - 2 * block, int(budget / max(sec1, 1e-6)) * nthreads))
        for _ in range(args.warmup):
            _ref_sample(ref, L, R, block, D, mid, rows, nthreads)
        ts = [_ref_sample(ref, L, R, block, D, mid, rows, nthreads)[0] for _ in range(args.steps)]
        kind, what = "reference", (f"oracle/_ref (the reference's BlockSearch.cpp compiled unchanged), SSD-L2 {block}x{block} "
                                   f"(the reference's only cost), one BlockSearch per row band on {used} threads")
    else:
        ocost = orc.COST_SAD if cost == "sad" else orc.COST_SSD_L2
        unit = 16 * nthreads
        cal = min(unit, h - 2 * block)
        _oracle_sample(orc, L, R, block, D, ocost, mid, 2, nthreads)
        per_row = _oracle_sample(orc, L, R, block, D, ocost, mid, cal, nthreads) / cal
        rows = int(budget / max(per_row, 1e-6))
        rows = max(min(cal, h - 2 * block), min(h - 2 * block, (rows // unit) * unit))
        for _ in range(args.warmup):
            _oracle_sample(orc, L, R, block, D, ocost, mid, rows, nthreads)
        ts = [_oracle_sample(orc, L, R, block, D, ocost, mid, rows, nthreads) for _ in range(args.steps)]
        used, kind = nthreads, "port"
        what = f"box-sum + OpenMP port of BlockSearch::computeDisparityMapRight, {cost.upper()} {block}x{block}, {nthreads} threads"
    ev = synth.effective_evals(w, w, rows, block, 0, D)
    sec = float(np.mean(ts))
    val = ev / sec / 1e6
    sample = f"{rows} of {h} rows of the {w}x{h} pair per step (rows {mid}..{mid + rows}), D={D}; {what}"
    cpu = {"value": val, "unit": "MDE/s", "cores": used, "kind": kind, "sample": sample}
    if kind == "reference":   # the optimised port beside it (the number round 1 reported as the reference arm)
        cpu["port"] = port_baseline(orc, L, R, w, h, D, block, cost, min(4.0, budget * 2), nthreads)
    line = {"impl": "reference", "metric": "disparity_evals_per_s", "value": val, "unit": "MDE/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
            "scaling": "weak" if args.workload == "batch720" else "strong", "vs_baseline": None, "dtype": "u8/int32", "data": "synthetic",
            "config": config_for(args, world), "cpu_baseline": cpu,
            "e2e": {"value": val, "unit": "MDE/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


def recon_roofline(ctx, pair, w, h, iters=5):
    """HBM-bound part of the path: main.cpp:41-65 on the search result resident in HBM --
    8-bit hand-off -> removeDisparityOutliers(500) -> depth -> vertices -> face count (sr_pair_reconstruct, outputs
    NULL: nothing leaves HBM but n_faces).  CUDA events on the context stream, L2 flushed between iterations.
    Algorithmic bytes per pixel after SURVEY.md 8d: hand-off 8+4, a6 4+4, a8 4+4, a9 3+4+32, a10 validity 32."""
    import ctypes as C
    f, null, nf = C.c_float, C.c_void_p(None), C.c_uint32(0)
    focal = 3000.0

    def pipeline():
        ctx.check(ctx._l.sr_pair_reconstruct(ctx.h, pair.h, 1, 500, f(1.5), f(0.8), f(focal), f(1.0), f(focal), f(focal),
                                             f(w / 2), f(h / 2), f(1.0), null, null, null, null, C.byref(nf)))
    for _ in range(2):
        pipeline()
    ts = []
    for _ in range(iters):
        ctx.l2_flush()
        ctx.timer_start()
        pipeline()
        ts.append(ctx.timer_stop())
    ms = float(np.median(ts))
    peak, src = 6551.7, "fallback of B200_PROFILING.md"
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
        src = "MEASURED_PEAKS.json hbm_gbs"
    except Exception:
        pass
    alg = w * h * 99
    return {"bound": "hbm", "achieved": alg / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
            "frac": alg / (ms * 1e-3) / 1e9 / peak, "ms": ms, "algorithmic_bytes": alg, "algorithmic_bytes_per_pixel": 99,
            "implementation_bytes_per_pixel": 75, "faces": int(nf.value), "peak_source": src,
            "what": "sr_pair_reconstruct on the resident search result: u8 hand-off, outliers k=500, depth, vertices, "
                    "face count (main.cpp:41-65); per-kernel figures in DESIGN.md 4.2c / profiles/r02_recon_4k_hbm.jsonl"}


def cpu_baseline(args, L, R, w, h, D, block, cost):
    """The CPU reference beside the GPU number, on rank 0 at N=1, bounded to about --cpu-seconds of CPU work in total:
    (reference) oracle/_ref = the reference's own code on all host threads, SSD-L2; (literal_1_thread) the same on one
    thread, as the reference actually runs; (port) the optimised box-sum port with the workload's cost."""
    from oracle import orc
    from stereo_reconstruction_b200 import synth
    nthreads = host_threads()
    mid = block
    out = None
    ref = _ref_available()
    if ref is not None:
        sec1, used = _ref_sample(ref, L, R, block, D, mid, nthreads, nthreads)
        rows = max(nthreads, min(h - 2 * block, int(0.4 * args.cpu_seconds / max(sec1, 1e-6)) * nthreads))
        sec, used = _ref_sample(ref, L, R, block, D, mid, rows, nthreads)
        ev = synth.effective_evals(w, w, rows, block, 0, D)
        out = {"value": ev / sec / 1e6, "unit": "MDE/s", "cores": used, "kind": "reference",
               "sample": f"rows {mid}..{mid + rows} of the same pair ({rows}/{h} rows, {sec:.2f} s): oracle/_ref = the reference's "
                         f"BlockSearch.cpp compiled unchanged, SSD-L2 {block}x{block} (its only cost), one BlockSearch per row band on {used} threads"}
        rows1 = max(1, rows // max(used, 1) // 2)
        sec_1, _ = _ref_sample(ref, L, R, block, D, mid, rows1, 1)
        out["literal_1_thread"] = {"value": synth.effective_evals(w, w, rows1, block, 0, D) / sec_1 / 1e6, "unit": "MDE/s", "cores": 1,
                                   "sample": f"{rows1} rows, {sec_1:.2f} s, extrapolates to {sec_1 / rows1 * h:.0f} s for the pair"}
    port = port_baseline(orc, L, R, w, h, D, block, cost, 0.4 * args.cpu_seconds, nthreads)
    if out is None:
        return port
    out["port"] = port
    return out


def emit(line):
    """The ONE JSON line of this run goes to the real stdout; everything else that writes to fd 1 (NCCL's version banner,
    library chatter) was sent to stderr by quiet_stdout()."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = 1


def quiet_stdout():
    global _REAL_STDOUT
    if _REAL_STDOUT == 1:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def main():
    args = parse()
    quiet_stdout()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if args.workload == "batch720":
        run_batch(args, rank, world, local_rank)
        return
    import torch
    import torch.distributed as dist
    import stereo_reconstruction_b200 as s
    from stereo_reconstruction_b200 import sharding, synth
    capi = s.capi
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        # stdout carries exactly one JSON line: NCCL's "NCCL version ..." banner (NCCL_DEBUG=VERSION) goes to stderr
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=dev)
    w, h, D, block, cost = WORKLOADS[args.workload]
    cost = args.cost or cost
    block = args.block or block
    ctx = capi.Context(local_rank)
    ext = torch.cuda.ExternalStream(ctx.stream, device=dev)

    L, R = synth.synth_pair(0xB200, 0, w, h, 0, D)
    # pinned host copies for the end-to-end leg
    Lp, lp_h = capi.pinned_empty(L.shape, np.uint8); Lp[...] = L
    Rp, rp_h = capi.pinned_empty(R.shape, np.uint8); Rp[...] = R
    outp, out_h = capi.pinned_empty((h, w), np.float64)
    params = capi.search_params(block=block, min_disp=0, max_disp=D, smooth=1.0,
                                cost=capi.COST_SAD if cost == "sad" else capi.COST_SSD_L2)
    pair = capi.Pair(ctx, w, h, w, h)
    pair.upload(Lp, Rp)
    r0, r1 = sharding.band_rows(h, rank, world)
    full = sharding.result_tensor(pair, dev)
    full16 = sharding.result16_tensor(pair, dev) if world > 1 else None   # int16 wire format of the band gather
    total_evals = synth.effective_evals(w, w, h, block, 0, D)
    # band gather: stores into rank 0's HBM over NVLink (capi_peer.cu); NCCL point-to-point where peer mapping fails
    peer = None
    if world > 1 and os.environ.get("SR_GATHER", "peer") != "nccl":   # "peer" (default), "peer2" (two-kernel form), "nccl"
        try:
            peer = sharding.PeerGather(pair, h, rank, world)
        except RuntimeError as e:
            if rank == 0:
                print(f"[bench] {e}; gathering over NCCL", file=sys.stderr, flush=True)

    def gather():
        if peer is not None:
            peer.gather()
        else:
            with torch.cuda.stream(ext):
                sharding.gather_bands16(pair, full16, h, rank, world)

    half = (block - 1) // 2
    h2d_rows = h if world == 1 else sum(min(b1 + half + 1, h) - max(b0 - half, 0)
                                        for b0, b1 in (sharding.band_rows(h, r, world) for r in range(world)))

    def barrier():
        ctx.sync()
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()

    two_kernel = os.environ.get("SR_GATHER") == "peer2"   # development: search + separate push kernel (round-1 form)

    def step_resident():
        if peer is not None and two_kernel:
            pair.run(params, r0, r1)
            peer.gather()
        elif peer is not None:
            peer.run_and_gather(params)      # ranks > 0: the resolve pass stores the int16 wire format into rank 0's window
        else:
            pair.run(params, r0, r1)
            if world > 1:
                gather()

    # end to end at N > 1: the CV_64F result lives in ONE host buffer shared by the ranks (POSIX shared memory, pinned in
    # every process): each rank uploads its band's rows (+ halo), searches and downloads its band straight into its rows
    # of that buffer -- N PCIe links instead of rank 0's one, no device-side gather on this route
    shm = out_shm = None
    if world > 1:
        from multiprocessing import shared_memory
        name = f"sr_bench_out_{os.environ.get('MASTER_PORT', '0')}"
        if rank == 0:
            try:
                old = shared_memory.SharedMemory(name=name); old.close(); old.unlink()
            except FileNotFoundError:
                pass
            shm = shared_memory.SharedMemory(name=name, create=True, size=h * w * 8)
        dist.barrier()
        if rank != 0:
            shm = shared_memory.SharedMemory(name=name)
        out_shm = np.ndarray((h, w), np.float64, buffer=shm.buf)
        capi.host_register(out_shm)

    def step_e2e_shared():
        pair.search_band(Lp, Rp, params, r0, r1, out=out_shm)         # sr_pair_search_band: H2D (band + halo) + kernels + D2H of the band

    def step_e2e():
        if world == 1:
            capi.block_search(ctx, Lp, Rp, params, out=outp)          # sr_block_search: H2D + kernels + D2H
        else:
            # every rank: H2D of its band's image rows (+ halo), band search, by row chunks (sr_pair_search_band);
            # rank 0 streams its own band back to the host inside that pipeline, the others keep theirs in HBM
            pair.search_band(Lp, Rp, params, r0, r1, out=outp if rank == 0 else None)
            gather()
            if rank == 0:
                ext.synchronize()
                pair.download(params, out=outp, row0=r1, row1=h)   # the gathered bands (rank 0 owns rows [0, r1))
            ctx.sync()

    sampler = ClockSampler(local_rank) if rank == 0 else None
    peak_dual, peak_single = ctx.int_peak()
    if sampler:
        sampler.mark()
    for _ in range(max(args.warmup, 3)):
        ctx.l2_flush()                       # warm-up steps look exactly like timed steps (first flush allocates)
        step_resident()
    barrier()
    if world > 1:
        # The ranks leave the host-side barrier up to ~1 ms apart; the first step after it would measure that skew (rank 0
        # waiting for the late ranks' bands: 1.73 ms against 0.92 ms for every later step at N=2; at N=8 the pushers' NEXT
        # step waiting for the late root's acknowledgement: 0.81 against 0.30 ms).  Two untimed steps let the gather's flag
        # and acknowledgement words pull the ranks into step; the K timed steps follow.
        for _ in range(2):
            ctx.l2_flush()
            step_resident()
    launches0 = ctx.launches
    step_ms, kern_ms = [], []
    for _ in range(args.steps):
        ctx.l2_flush()                       # between timed iterations: write 384 MB > 126 MB L2
        ctx.timer_start()
        step_resident()
        step_ms.append(ctx.timer_stop())
        if ctx.last_kernel == "tile":
            kern_ms.append(ctx.search_kernel_ms())
    barrier()
    launches = ctx.launches - launches0 - args.steps * 0   # memset of the flush is not a kernel of ours
    if os.environ.get("SR_BENCH_DEBUG"):
        print(f"[rank {rank}] step ms {[round(x, 3) for x in step_ms]} kernel ms {[round(x, 3) for x in kern_ms]}", file=sys.stderr, flush=True)
    t_total = torch.tensor([sum(step_ms)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_total, op=dist.ReduceOp.MAX)
    ms_per_step = float(t_total.item()) / args.steps

    # end to end through the host-pointer API
    for _ in range(2):
        step_e2e()
    barrier()
    e2e_t = []
    for _ in range(max(3, min(args.steps, 10))):
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        step_e2e()
        e2e_t.append(time.perf_counter() - t0)
    barrier()
    e2e_shared_s = None
    if world > 1:
        for _ in range(2):
            step_e2e_shared()
        barrier()
        ts = []
        for _ in range(max(3, min(args.steps, 10))):
            dist.barrier()
            t0 = time.perf_counter()
            step_e2e_shared()
            ts.append(time.perf_counter() - t0)
        barrier()
        tt = torch.tensor([float(np.mean(ts))], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e_shared_s = float(tt.item())
    clocks = sampler.stop() if sampler else None
    # the same call with pageable buffers (what a cv::Mat holds): staged through pinned memory by host threads
    pageable_ms = None
    if world == 1:
        out_pg = np.zeros((h, w), np.float64)
        for _ in range(2):
            capi.block_search(ctx, L, R, params, out=out_pg)
        t0 = time.perf_counter()
        for _ in range(3):
            capi.block_search(ctx, L, R, params, out=out_pg)
        pageable_ms = (time.perf_counter() - t0) / 3 * 1e3
        del out_pg
    t_e2e = torch.tensor([float(np.mean(e2e_t))], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_e2e, op=dist.ReduceOp.MAX)
    e2e_s = float(t_e2e.item())

    # parity spot check against nothing here (tests do that); report result checksum for reproducibility
    if rank == 0:
        res = pair.download(params)
        checksum = float(res.sum())
        gather_ok = shared_ok = None
        if world > 1:
            pair.run(params)                 # the whole pair on this GPU: the gathered bands must reproduce it
            single = pair.download(params)
            gather_ok = bool(np.array_equal(res, single))
            shared_ok = bool(np.array_equal(out_shm, single))
        band_evals = synth.effective_evals(w, w, r1 - r0, block, 0, D)
        k_ms = float(np.mean(kern_ms)) if kern_ms else None
        ops = OPS_PER_EVAL[cost]
        roofline = None
        if k_ms:
            achieved = band_evals * ops / (k_ms * 1e-3) / 1e12
            in_bytes = 2 * w * (r1 - r0) * 3 + w * (r1 - r0) * 8   # algorithmic: BGR in (both images) + CV_64F out
            roofline = {"bound": "alu", "achieved": achieved, "peak": peak_dual / 1e12, "unit": "Tint-op/s",
                        "frac": achieved / (peak_dual / 1e12),
                        "traffic": NCU_TRAFFIC.get(args.workload) if world == 1 else None,   # captured for the whole-pair launch only
                        "kernel": "search_tile_kernel", "kernel_ms": k_ms, "ops_per_eval": ops,
                        "evals_per_launch": band_evals, "kernel_MDE_per_s": band_evals / (k_ms * 1e-3) / 1e6,
                        "peak_source": "measured live: VABSDIFF4+IDP.4A dual-pipe issue (sr_int_peak)",
                        "peak_single_pipe": peak_single / 1e12, "frac_of_single_pipe": achieved / (peak_single / 1e12),
                        "hbm": {"algorithmic_GBps": in_bytes / (k_ms * 1e-3) / 1e9, "peak_GBps": 6547.2,
                                "note": "14 B/pixel algorithmic; the kernel is ALU-bound (SURVEY.md 8d)"}}
        line = {
            "metric": "disparity_evals_per_s", "value": total_evals / (ms_per_step * 1e-3) / 1e6, "unit": "MDE/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u8/int32", "data": "synthetic",
            "config": config_for(args, world),
            "gather": (None if world == 1 else
                       "int16 stores into rank 0's HBM over NVLink (peer memory), widened by rank 0 as the flags arrive" if peer is not None else
                       "NCCL point-to-point to rank 0 (int16 wire format, widened on rank 0)"),
            "effective_MDE": total_evals / 1e6,
            "pairs_per_s": 1e3 / ms_per_step,
            "nominal_value": w * h * D / (ms_per_step * 1e-3) / 1e6,
            "e2e": ({"value": total_evals / e2e_s / 1e6, "unit": "MDE/s", "ms_per_step": e2e_s * 1e3,
                     "h2d_bytes_per_step": int(h2d_rows * w * 3 * 2), "d2h_bytes_per_step": int(w * h * 8),
                     "pageable_ms_per_step": pageable_ms,
                     "api": "sr_block_search (pinned host buffers; pageable_ms_per_step = same call on pageable buffers)"} if world == 1 else
                    {"value": total_evals / e2e_shared_s / 1e6, "unit": "MDE/s", "ms_per_step": e2e_shared_s * 1e3,
                     "h2d_bytes_per_step": int(h2d_rows * w * 3 * 2), "d2h_bytes_per_step": int(w * h * 8),
                     "api": "sr_pair_search_band on every rank: H2D of the band's image rows + halo, band search, D2H of the band into its rows "
                            "of ONE result buffer in POSIX shared memory pinned by every process (N PCIe links, no device-side gather)",
                     "result_equals_single_gpu": shared_ok,
                     "via_rank0": {"value": total_evals / e2e_s / 1e6, "ms_per_step": e2e_s * 1e3,
                                   "api": "sr_pair_search_band + band gather into rank 0's HBM + sr_search_download on rank 0 (one PCIe link carries the 66 MB)"}}),
            "gpu_launches": int(launches),
            "kernel_family": ctx.last_kernel,
            "clocks": clocks,
            "roofline": roofline,
            "result_checksum": checksum,
        }
        if world > 1:
            line["gathered_map_equals_single_gpu"] = gather_ok
        if world == 1 and not os.environ.get("SR_BENCH_NO_RECON"):
            # secondary, after every timed region: the HBM-bound Reconstruction kernels on the same pair
            try:
                line["roofline_recon"] = recon_roofline(ctx, pair, w, h)
            except Exception as e:   # never at the expense of the headline line
                line["roofline_recon"] = {"error": str(e)}
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(args, L, R, w, h, D, block, cost)
        emit(line)
    barrier()
    capi.free_pinned(lp_h); capi.free_pinned(rp_h); capi.free_pinned(out_h)
    if world > 1:
        pair.peer_status()                   # a failed step of the gather protocol is an error on every rank
        capi.host_unregister(out_shm)
        del out_shm
        dist.barrier()
        shm.close()
        if rank == 0:
            shm.unlink()
        dist.destroy_process_group()


def run_batch(args, rank, world, local_rank):
    """BASELINE.json configs[4]: batch of 1280x720 pairs, 128 disparities, 7x7 SAD + sub-pixel, sharded by
    pair (pair p -> rank p mod N, no collective).  A step = `--batch-pairs` pairs per rank through
    sr_block_search_batch (pinned host buffers, H2D / kernels / D2H pipelined); weak scaling."""
    import torch
    import torch.distributed as dist
    import stereo_reconstruction_b200 as s
    from stereo_reconstruction_b200 import synth
    capi = s.capi
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=dev)
    w, h, D, block = 1280, 720, 128, 7
    n = args.batch_pairs
    ctx = capi.Context(local_rank)
    ring = 8   # distinct pinned pairs, reused round-robin over the batch (global pair index p = rank + i * world)
    Ls, Rs, outs, hs = [], [], [], []
    for i in range(ring):
        L, R = synth.synth_pair(0xB200, rank + i * world, w, h, 0, D)
        Lp, a = capi.pinned_empty(L.shape, np.uint8); Lp[...] = L
        Rp, b = capi.pinned_empty(R.shape, np.uint8); Rp[...] = R
        o, c = capi.pinned_empty((h, w), np.float64)
        Ls.append(Lp); Rs.append(Rp); outs.append(o); hs += [a, b, c]
    p = capi.search_params(block=block, min_disp=0, max_disp=D, smooth=1.0, cost=capi.COST_SAD, subpixel=True)
    lb = [Ls[i % ring] for i in range(n)]; rb = [Rs[i % ring] for i in range(n)]; ob = [outs[i % ring] for i in range(n)]
    evals = synth.effective_evals(w, w, h, block, 0, D)

    def barrier():
        ctx.sync(); torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()

    # device-resident leg: the n pairs' kernels on resident pairs, 8 per tile-kernel launch (sr_search_run_batch)
    res = [capi.Pair(ctx, w, h, w, h) for _ in range(ring)]
    for i, q in enumerate(res):
        q.upload(Ls[i], Rs[i])
    sampler = ClockSampler(local_rank) if rank == 0 else None
    for _ in range(max(args.warmup, 3)):
        capi.search_run_batch(ctx, res, p)
    l0 = ctx.launches
    barrier()
    ctx.timer_start()
    for _ in range(args.steps):
        for i0 in range(0, n, ring):
            capi.search_run_batch(ctx, res[:min(ring, n - i0)], p)
    dev_ms = ctx.timer_stop()
    launches = ctx.launches - l0
    barrier()
    for _ in range(2):
        capi.block_search_batch(ctx, lb, rb, p, ob)       # same path as the timed calls (the first one builds the ring of pairs)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        capi.block_search_batch(ctx, lb, rb, p, ob)
    e2e_s = time.perf_counter() - t0
    barrier()
    clocks = sampler.stop() if sampler else None
    t = torch.tensor([dev_ms * 1e-3, e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_s, e2e_s = float(t[0]), float(t[1])
    if rank == 0:
        total_pairs = n * args.steps * world
        line = {"metric": "disparity_evals_per_s", "value": total_pairs * evals / dev_s / 1e6, "unit": "MDE/s", "n_gpus": world,
                "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": dev_s * 1e3 / args.steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/int32", "data": "synthetic",
                "config": config_for(args, world),
                "pairs_per_s": total_pairs / dev_s,
                "e2e": {"value": total_pairs * evals / e2e_s / 1e6, "unit": "MDE/s", "pairs_per_s": total_pairs / e2e_s,
                        "h2d_bytes_per_step": int(n * 2 * w * h * 3), "d2h_bytes_per_step": int(n * w * h * 8),
                        "api": "sr_block_search_batch (pinned host buffers)"},
                "gpu_launches": int(launches), "kernel_family": ctx.last_kernel, "clocks": clocks, "roofline": None}
        emit(line)
    barrier()
    for x in hs:
        capi.free_pinned(x)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
