set -e
python -m pytest tests/test_gpu_search.py -x -q -m gpu 2>&1 | tail -3
for t in 4 8 12 16; do echo "threads $t"; SR_STAGE_THREADS=$t python tools/e2e_pageable.py 2>&1 | grep pageable; done
echo default; python tools/e2e_pageable.py
