#!/bin/bash
# Development sweep of the tile kernel's per-warp instruction mixes (SR_TILE_ZMASK / SR_TILE_FMASK), 4K / 512 / 9x9.
# usage: tools/tune_tile.sh > gpurun_out/tune.log
for z in 0x0000 0xFFF0 0xFF00 0xFFFF 0x7BDE; do
  echo "SAD zmask=$z"; SR_TILE_ZMASK=$z python tools/quick_bench.py --cost 1 --iters 15 2>&1 | tail -2
done
for f in 0x0000 0xFFFF; do for z in 0x0000 0xFFFF; do
  echo "SSD fmask=$f zmask=$z"; SR_TILE_FMASK=$f SR_TILE_ZMASK=$z python tools/quick_bench.py --cost 0 --iters 15 2>&1 | tail -2
done; done
