"""Stress of the Left-variant wavefront and the chunked smoothing pipeline against the oracle on random shapes
(development aid: races in the hand-over protocols would show up as sporadic mismatches).  GPU box only."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import stereo_reconstruction_b200 as sr
from oracle import orc
ctx = sr.capi.default_context()
rng = np.random.default_rng(int(sys.argv[1]) if len(sys.argv) > 1 else 1)
bad = 0
for it in range(int(sys.argv[2]) if len(sys.argv) > 2 else 40):
    w = int(rng.integers(40, 700)); h = int(rng.integers(20, 90)); B = int(rng.choice([3, 5, 7, 9, 11, 13, 15, 17])); D = int(rng.integers(2, 90))
    s = float(rng.choice([1.0, 0.9, 0.5, 0.97, 1.2]))
    if h <= B + 2 or w <= B + 2: continue
    L, R = orc.synth_pair(int(rng.integers(1, 1 << 30)), it, w, h, 0, min(D, 60), bool(rng.integers(0, 2)))
    want = orc.block_search_left(L, R, B, D, s)
    for rep in range(3):
        got = sr.BlockSearch(L, R, B, 0, D, ctx=ctx).computeDisparityMapLeft(s)
        if not np.array_equal(got, want):
            bad += 1; print("LEFT MISMATCH", w, h, B, D, s, int((got != want).sum()), flush=True)
    # chunked smoothing through the one-shot call
    os.environ["SR_PIPE_CHUNKS"] = str(int(rng.integers(1, 7)))
    Br = int(rng.choice([3, 5, 9, 17]))
    wantr = orc.block_search_right_fast(L, R, Br, 0, D, s)
    p = sr.capi.search_params(block=Br, min_disp=0, max_disp=D, smooth=s)
    for rep in range(2):
        gotr = sr.capi.block_search(ctx, L, R, p)
        if not np.array_equal(gotr, wantr):
            bad += 1; print("RIGHT MISMATCH", w, h, Br, D, s, os.environ["SR_PIPE_CHUNKS"], int((gotr != wantr).sum()), flush=True)
print("stress done, mismatches:", bad)
