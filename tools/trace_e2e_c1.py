"""Config-1 shape through the one-shot call (pinned host buffers): median time per SR_PIPE_CHUNKS setting."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from stereo_reconstruction_b200 import capi, synth
ctx = capi.default_context()
w, h, D = (int(v) for v in (sys.argv[1:4] if len(sys.argv) >= 4 else (1436, 992, 145)))
B = int(sys.argv[4]) if len(sys.argv) >= 5 else 17
L, R = synth.synth_pair(0xB200, 0, w, h, 0, D)
Lp, a1 = capi.pinned_empty(L.shape, np.uint8); Lp[...] = L
Rp, a2 = capi.pinned_empty(R.shape, np.uint8); Rp[...] = R
o, a3 = capi.pinned_empty((h, w), np.float64)
p = capi.search_params(block=B, min_disp=0, max_disp=D, smooth=float(os.environ.get("SR_TRACE_SMOOTH", "1.0")))
for _ in range(4):
    capi.block_search(ctx, Lp, Rp, p, out=o)
ts = []
for _ in range(20):
    t0 = time.perf_counter(); capi.block_search(ctx, Lp, Rp, p, out=o); ts.append(time.perf_counter() - t0)
print("chunks", os.environ.get("SR_PIPE_CHUNKS", "default"), w, h, D, B, "e2e ms median", round(1e3 * float(np.median(ts)), 4), "min", round(1e3 * min(ts), 4), flush=True)
