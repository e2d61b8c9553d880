"""Config-1 shape with the raster rule (smooth 0.9): device-resident time; run under ncu for the per-kernel list."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from stereo_reconstruction_b200 import capi, synth
ctx = capi.default_context()
w, h, D, B = 1436, 992, 145, 17
L, R = synth.synth_pair(0xB200, 0, w, h, 0, D)
pair = capi.Pair(ctx, w, h, w, h); pair.upload(L, R)
for smooth in (0.9, 1.0):
    p = capi.search_params(block=B, min_disp=0, max_disp=D, smooth=smooth)
    for _ in range(3): pair.run(p)
    ctx.sync(); ts = []
    for _ in range(9):
        ctx.timer_start(); pair.run(p); ts.append(ctx.timer_stop())
    print(json.dumps({"w": w, "h": h, "D": D, "block": B, "smooth": smooth, "ms": round(float(np.median(ts)), 4)}), flush=True)
