"""Left variant (computeDisparityMapLeft) timing: tile-kernel path against the direct top-3 path, per chunk size.
Usage (GPU box): python tools/left_bench.py [w h D block smooth]   -- env SR_LEFT_DIRECT / SR_LEFT_CHUNK are read by the library."""
import json, os, sys, time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from stereo_reconstruction_b200 import capi, synth

w, h, D, block = (int(v) for v in (sys.argv[1:5] if len(sys.argv) >= 5 else (900, 750, 128, 17)))
smooth = float(sys.argv[5]) if len(sys.argv) >= 6 else 0.9
ctx = capi.default_context()
L, R = synth.synth_pair(0xB200, 0, w, h, 0, D)
pair = capi.Pair(ctx, w, h, w, h)
pair.upload(L, R)
p = capi.search_params(variant=capi.VARIANT_LEFT, block=block, min_disp=0, max_disp=D, smooth=smooth)
for _ in range(2):
    pair.run(p)
ctx.sync()
ts = []
for _ in range(5):
    ctx.timer_start()
    pair.run(p)
    ts.append(ctx.timer_stop())
out = pair.download(p)
print(json.dumps({"case": "left-variant", "w": w, "h": h, "D": D, "block": block, "smooth": smooth,
                  "direct": os.environ.get("SR_LEFT_DIRECT", "0"), "chunk": os.environ.get("SR_LEFT_CHUNK", "default"),
                  "ms_median": round(float(np.median(ts)), 4), "ms_min": round(min(ts), 4),
                  "checksum": int(np.asarray(out, np.int64).sum()), "nonzero": int((out != 0).sum())}), flush=True)
