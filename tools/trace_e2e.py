"""One-shot sr_block_search on the 4K pair from and to pinned host buffers: median time; with SR_PIPE_TRACE=1 the library
prints the per-chunk timeline (upload done / search start / search end / download done) on stderr."""
import sys, os, time
sys.path.insert(0, ".")
import numpy as np
from stereo_reconstruction_b200 import capi, synth
ctx = capi.default_context()
w, h, D = 3840, 2160, 512
L, R = synth.synth_pair(0xB200, 0, w, h, 0, D)
Lp, a1 = capi.pinned_empty(L.shape, np.uint8); Lp[...] = L
Rp, a2 = capi.pinned_empty(R.shape, np.uint8); Rp[...] = R
o, a3 = capi.pinned_empty((h, w), np.float64)
p = capi.search_params(block=9, min_disp=0, max_disp=D, smooth=1.0, cost=capi.COST_SAD)
for _ in range(4):
    capi.block_search(ctx, Lp, Rp, p, out=o)
ts = []
for _ in range(10):
    t0 = time.perf_counter(); capi.block_search(ctx, Lp, Rp, p, out=o); ts.append(time.perf_counter() - t0)
print("e2e ms median", round(1e3 * float(np.median(ts)), 4), "min", round(1e3 * min(ts), 4), flush=True)
os.environ["SR_PIPE_TRACE"] = "1"
