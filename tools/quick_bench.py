"""Kernel-only timing of one synthetic pair (development helper; bench.py is the contract)."""
import argparse, sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import stereo_reconstruction_b200 as s
from oracle import orc

ap = argparse.ArgumentParser()
ap.add_argument("--w", type=int, default=3840); ap.add_argument("--h", type=int, default=2160)
ap.add_argument("--d", type=int, default=512); ap.add_argument("--block", type=int, default=9)
ap.add_argument("--cost", type=int, default=0); ap.add_argument("--smooth", type=float, default=1.0)
ap.add_argument("--chunks", type=int, default=1)
ap.add_argument("--iters", type=int, default=10); ap.add_argument("--check-rows", type=int, default=24)
a = ap.parse_args()
ctx = s.capi.default_context(0)
L, R = orc.synth_pair(0xB200, 0, a.w, a.h, 0, a.d)
pair = s.capi.Pair(ctx, a.w, a.h, a.w, a.h)
pair.upload(L, R)
p = s.capi.search_params(block=a.block, min_disp=0, max_disp=a.d, smooth=a.smooth, cost=a.cost)
def run_all():
    if a.chunks == 1:
        pair.run(p)
    else:
        for c in range(a.chunks):
            pair.run(p, a.h * c // a.chunks, a.h * (c + 1) // a.chunks)
for _ in range(3): run_all()
ctx.sync()
ts = []
for _ in range(a.iters):
    ctx.timer_start(); run_all(); ts.append(ctx.timer_stop())
ev = pair.effective_evals()
ms = float(np.median(ts))
print(f"kernel={ctx.last_kernel} {a.w}x{a.h} D={a.d} B={a.block} cost={a.cost} smooth={a.smooth}: median {ms:.3f} ms (min {min(ts):.3f}) "
      f"nominal {a.w*a.h*a.d/ms/1e3:.0f} MDE/s effective {ev/ms/1e3:.0f} MDE/s evals={ev}")
if a.check_rows:
    out = pair.download(p)
    rows = (a.h // 2, a.h // 2 + a.check_rows)
    if a.smooth == 1.0:
        want = orc.block_search_right_fast(L, R, a.block, 0, a.d, a.smooth, cost=a.cost, rows=rows)
        bad = int((out[rows[0]:rows[1]] != want[rows[0]:rows[1]]).sum())
        print(f"parity on rows {rows}: {bad} mismatching pixels")
