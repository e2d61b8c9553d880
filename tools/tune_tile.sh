#!/bin/bash
# Development sweep of the tile kernel's per-warp instruction mix (SR_TILE_ZMASK: bit w = warp w slides its window sums on
# the FMA pipe), 4K / 512 / 9x9 SAD.   usage: tools/tune_tile.sh > gpurun_out/tune.log
for z in 0x0000 0xFFF0 0xFF00 0x0F0F 0xF0F0 0xFFFF 0x7BDE; do
  echo "SAD zmask=$z"; SR_TILE_ZMASK=$z python tools/quick_bench.py --cost 1 --iters 15 2>&1 | tail -2
done
echo "SSD"; python tools/quick_bench.py --cost 0 --iters 15 2>&1 | tail -2
