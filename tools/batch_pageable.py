"""sr_block_search_batch on pageable (cv::Mat-like) buffers next to pinned ones: pairs/s for the config-5 shape."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from stereo_reconstruction_b200 import capi, synth
ctx = capi.default_context(0)
w, h, D, n = 1280, 720, 128, 32
pairs = [synth.synth_pair(0xB200, i, w, h, 0, D) for i in range(n)]
p = capi.search_params(block=7, min_disp=0, max_disp=D, smooth=1.0, cost=capi.COST_SAD, subpixel=True)
outs = [np.zeros((h, w), np.float64) for _ in range(n)]
Ls, Rs = [a for a, _ in pairs], [b for _, b in pairs]
def rate(L, R, O):
    capi.block_search_batch(ctx, L, R, p, O)
    t0 = time.perf_counter()
    for _ in range(3): capi.block_search_batch(ctx, L, R, p, O)
    return 3 * n / (time.perf_counter() - t0)
print("pageable: %.0f pairs/s" % rate(Ls, Rs, outs))
ref = [o.copy() for o in outs]
hs = []
def pin(a):
    b, hnd = capi.pinned_empty(a.shape, a.dtype); b[...] = a; hs.append(hnd); return b
Lp, Rp, Op = [pin(a) for a in Ls], [pin(a) for a in Rs], [pin(o) for o in outs]
for o in Op: o[...] = 0
print("pinned:   %.0f pairs/s" % rate(Lp, Rp, Op))
print("identical:", all(np.array_equal(np.asarray(a), b) for a, b in zip(Op, ref)))
for hnd in hs: capi.free_pinned(hnd)
