"""A/B timing of the tile kernel on a few shapes (device-resident, CUDA events, L2 not flushed, median of 9).
Usage (GPU box): SR_TILE_NW=8|16 python tools/ab_tile.py"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from stereo_reconstruction_b200 import capi, synth
ctx = capi.default_context()
SHAPES = [(3840, 2160, 512, 9, "sad"), (3840, 2160, 512, 9, "ssd"), (1920, 1080, 256, 5, "sad"), (1920, 1080, 256, 9, "sad"),
          (1280, 720, 128, 7, "sad"), (1436, 992, 145, 11, "ssd"), (3840, 270, 512, 9, "sad")]
for w, h, D, B, cost in SHAPES:
    L, R = synth.synth_pair(0xB200, 0, w, h, 0, D)
    pair = capi.Pair(ctx, w, h, w, h); pair.upload(L, R)
    p = capi.search_params(block=B, min_disp=0, max_disp=D, smooth=1.0, cost=capi.COST_SAD if cost == "sad" else capi.COST_SSD_L2)
    for _ in range(3): pair.run(p)
    ctx.sync()
    ts = []
    for _ in range(9):
        ctx.timer_start(); pair.run(p); ts.append(ctx.timer_stop())
    out = pair.download(p)
    print(json.dumps({"nw": os.environ.get("SR_TILE_NW", "default"), "zmask": os.environ.get("SR_TILE_ZMASK", "default"), "w": w, "h": h, "D": D, "block": B, "cost": cost,
                      "ms": round(float(np.median(ts)), 4), "kernel_ms": round(ctx.search_kernel_ms(), 4) if hasattr(ctx, "search_kernel_ms") else None,
                      "checksum": int(np.asarray(out, np.int64).sum())}), flush=True)
    pair.close()
