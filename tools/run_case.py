"""Runs one search configuration a few times (for ncu launch lists / captures of the non-headline paths)."""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import stereo_reconstruction_b200 as s
from stereo_reconstruction_b200 import capi, synth
ap = argparse.ArgumentParser()
ap.add_argument("--w", type=int, default=1436); ap.add_argument("--h", type=int, default=992)
ap.add_argument("--d", type=int, default=145); ap.add_argument("--block", type=int, default=17)
ap.add_argument("--cost", type=int, default=0); ap.add_argument("--smooth", type=float, default=0.9)
ap.add_argument("--subpixel", action="store_true"); ap.add_argument("--lr", action="store_true")
ap.add_argument("--iters", type=int, default=2)
a = ap.parse_args()
ctx = capi.default_context(0)
L, R = synth.synth_pair(0xB200, 0, a.w, a.h, 0, a.d)
pair = capi.Pair(ctx, a.w, a.h, a.w, a.h)
pair.upload(L, R)
p = capi.search_params(block=a.block, min_disp=0, max_disp=a.d, smooth=a.smooth, cost=a.cost, subpixel=a.subpixel, lr_check=a.lr)
for _ in range(a.iters):
    pair.run(p)
ctx.sync()
print(ctx.last_kernel, ctx.launches)
