"""End-to-end time of sr_block_search with pageable host buffers (what a cv::Mat holds) next to pinned ones."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import stereo_reconstruction_b200 as s
from stereo_reconstruction_b200 import capi, synth
ctx = capi.default_context(0)
w, h, D = 3840, 2160, 512
L, R = synth.synth_pair(0xB200, 0, w, h, 0, D)
out = np.zeros((h, w), np.float64)
p = capi.search_params(block=9, min_disp=0, max_disp=D, smooth=1.0, cost=capi.COST_SAD)
def timed(fn, n=5):
    fn(); fn()
    t0 = time.perf_counter()
    for _ in range(n): fn()
    return (time.perf_counter() - t0) / n * 1e3
print("pageable buffers: %.2f ms" % timed(lambda: capi.block_search(ctx, L, R, p, out=out)))
t0 = time.perf_counter()
for a in (L, R, out): capi.host_register(a)
print("sr_host_register of the three buffers: %.2f ms (once)" % ((time.perf_counter() - t0) * 1e3))
print("registered buffers: %.2f ms" % timed(lambda: capi.block_search(ctx, L, R, p, out=out)))
for a in (L, R, out): capi.host_unregister(a)
