"""The oracle against the reference ITSELF at full size: tests/golden/ref_full/*.npz are whole disparity maps computed
by oracle/_ref/libsr_ref.so (the reference's BlockSearch.cpp compiled unchanged, tests/golden/make_ref_full.py) on every
shipped MiddEval3 trainingH pair (block 17, minD 0, maxD = calib ndisp, smoothFactor 0.9 and 1.0) and on the
reference's own rectified pairs (maxD 200, smoothFactor 0.9 = main.cpp:40 verbatim).  The box-sum oracle port that the
GPU tests and bench.py's cpu_baseline use must reproduce each of them pixel for pixel."""
import os

import numpy as np
import pytest

from refdata import REF_FULL, load_pair, ref_full_cases


@pytest.mark.parametrize("name", ref_full_cases(), ids=str)
def test_oracle_port_equals_reference_full_map(orc, name):
    z = np.load(os.path.join(REF_FULL, name + ".npz"))
    L, R = load_pair(str(z["kind"]), str(z["scene"]))
    assert L.shape[:2] == tuple(z["left_shape"]) and R.shape[:2] == tuple(z["right_shape"])
    got = orc.block_search_right_fast(L, R, int(z["block"]), int(z["min_d"]), int(z["max_d"]), float(z["smooth"]))
    bad = int((got != z["out"]).sum())
    assert bad == 0, f"{name}: {bad} pixels differ from the compiled reference"


def test_all_reference_maps_present():
    names = ref_full_cases()
    assert sum(n.startswith("middeval_") for n in names) == 30      # 15 pairs x smooth {0.9, 1.0}
    assert sum(n.startswith("rectified_") for n in names) >= 10     # 11 shipped; MotorcycleE (h1 < h2) throws in the reference
