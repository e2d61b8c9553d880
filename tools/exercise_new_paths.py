"""Small runs of the entry points added late in round 2 (Left variant through the tile kernel, chunked smoothing, batched
sub-pixel pass, bilinear warp, warped upload, fused reconstruction): a quick exercise on the GPU box, e.g. under a memory
checker where one is available (compute-sanitizer is closed on this pool; bounds were checked by hand)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import stereo_reconstruction_b200 as sr
from stereo_reconstruction_b200 import capi, synth
ctx = capi.default_context()
R_ = sr.reconstruction
for (w, h, D, B, s) in ((300, 60, 40, 9, 0.9), (530, 40, 70, 17, 0.9), (120, 30, 200, 3, 1.0), (200, 50, 20, 5, 1.3)):
    L, R = synth.synth_pair(7, 0, w, h, 0, min(D, 60))
    out = sr.BlockSearch(L, R, B, 0, D, ctx=ctx).computeDisparityMapLeft(s)
    print("left", w, h, D, B, s, int(out.sum()))
L, R = synth.synth_pair(9, 0, 640, 200, 0, 48)
for chunks in ("1", "3", "5"):
    os.environ["SR_PIPE_CHUNKS"] = chunks
    for s in (0.9, 1.0):
        o = capi.block_search(ctx, L, R, capi.search_params(block=9, min_disp=0, max_disp=48, smooth=s))
        print("right chunks", chunks, s, int(o.sum()))
os.environ.pop("SR_PIPE_CHUNKS")
p = capi.search_params(block=7, min_disp=0, max_disp=48, smooth=1.0, cost=capi.COST_SAD, subpixel=True)
pairs = [capi.Pair(ctx, 640, 200, 640, 200) for _ in range(3)]
for q in pairs: q.upload(L, R)
capi.search_run_batch(ctx, pairs, p); print("batch subpix", float(pairs[2].download(p).sum()))
M = np.array([[0.98, -0.03, 9.0], [0.02, 1.01, -4.0], [-4e-5, 1.5e-5, 1.0]])
img = np.ascontiguousarray(L)
print("warp", int(R_.warpPerspectiveLinear(img, M, (701, 233), ctx=ctx).sum()))
pw = capi.Pair(ctx, 650, 210, 630, 190); pw.upload_warped(L, M, R, np.linalg.inv(M)); pw.run(capi.search_params(block=9, min_disp=0, max_disp=48, smooth=0.9))
print("upload_warped", float(pw.download(capi.search_params(block=9, min_disp=0, max_disp=48, smooth=0.9)).sum()))
bgr = img; disp = np.abs(np.random.default_rng(1).normal(30, 10, (200, 640))).astype(np.float32)
K = np.array([[500.0, 0, 320], [0, 500, 100], [0, 0, 1]], np.float32)
got = R_.reconstructFused(bgr, disp, K, 500.0, 0.1, kernelSize=9, thrMesh=0.5, ctx=ctx)
print("recon faces", len(got["faces"]))
print("done")
