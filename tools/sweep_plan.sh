#!/bin/bash
# Development: the tile kernel's time over forced schedules (SR_TILE_NBANDS, SR_TILE_TAIL = "percent,divisor")
echo "== 4K"; for nb in 8 12 16 20 24 28 32 40; do for tl in 0,2 10,4 20,3; do
  echo -n "nb=$nb tail=$tl: "; SR_TILE_NBANDS=$nb SR_TILE_TAIL=$tl python tools/quick_bench.py --cost 1 --iters 9 --check-rows 0 2>&1 | tail -1 | sed 's/.*median \([0-9.]*\) ms.*/\1/'
done; done
echo "== 3840x270 (one of 8 bands)"; for nb in 1 2 3 4 5 6 8; do for tl in 0,2 15,2 20,3 30,3 40,2; do
  echo -n "nb=$nb tail=$tl: "; SR_TILE_NBANDS=$nb SR_TILE_TAIL=$tl python tools/quick_bench.py --cost 1 --h 270 --iters 15 --check-rows 0 2>&1 | tail -1 | sed 's/.*median \([0-9.]*\) ms.*/\1/'
done; done
echo "== auto"; SR_TILE_PLAN_DEBUG=1 python tools/quick_bench.py --cost 1 --h 270 --iters 15 --check-rows 0 2>&1 | tail -2
