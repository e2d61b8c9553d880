"""Pins the CPU oracle (oracle/sr_oracle.c) against the golden vectors produced by executing the
reference's arithmetic through cv2 (tests/golden/make_golden.py).  CPU only."""
import numpy as np
import pytest

from conftest import golden_cases


@pytest.mark.parametrize("name,z", golden_cases("right"), ids=lambda v: v if isinstance(v, str) else "")
def test_right_literal(orc, name, z):
    out, mb = orc.block_search_right(z["left"], z["right"], int(z["block"]), int(z["min_d"]), int(z["max_d"]),
                                     float(z["smooth"]), bool(z["var_block"]), float(z["thres"]), return_max_block=True)
    np.testing.assert_array_equal(out, z["out"])
    assert mb == int(z["max_block"])


@pytest.mark.parametrize("name,z", [c for c in golden_cases("right") if not int(c[1]["var_block"])],
                         ids=lambda v: v if isinstance(v, str) else "")
def test_right_fast_equals_golden(orc, name, z):
    out = orc.block_search_right_fast(z["left"], z["right"], int(z["block"]), int(z["min_d"]), int(z["max_d"]), float(z["smooth"]))
    np.testing.assert_array_equal(out, z["out"])


@pytest.mark.parametrize("name,z", golden_cases("left"), ids=lambda v: v if isinstance(v, str) else "")
def test_left(orc, name, z):
    out = orc.block_search_left(z["left"], z["right"], int(z["block"]), int(z["max_d"]), float(z["smooth"]))
    np.testing.assert_array_equal(out, z["out"])


@pytest.mark.parametrize("name,z", golden_cases("linear"), ids=lambda v: v if isinstance(v, str) else "")
def test_linear(orc, name, z):
    out = orc.linear_search(z["left"], z["right"], float(z["smooth"]))
    np.testing.assert_array_equal(out, z["out"])


@pytest.mark.parametrize("name,z", golden_cases("outliers"), ids=lambda v: v if isinstance(v, str) else "")
def test_outliers(orc, name, z):
    b = orc.box_blur(z["disp"], int(z["k"]))
    np.testing.assert_array_equal(b, z["blur"])
    d = orc.remove_outliers(z["disp"], int(z["k"]), float(z["front"]), float(z["back"]))
    np.testing.assert_array_equal(d, z["out"])


def test_outliers_full(orc):
    (name, z), = golden_cases("outliers_full")
    disp = z["disp_u8"].astype(np.float32)
    b = orc.box_blur(disp, int(z["k"]))
    np.testing.assert_array_equal(b[::25, ::25], z["blur_sample"])
    d = orc.remove_outliers(disp, int(z["k"]), float(z["front"]), float(z["back"]))
    np.testing.assert_array_equal(d[::25, ::25], z["out_sample"])
    assert float(d.astype(np.float64).sum()) == float(z["out_sum"])
    assert int((d != disp).sum()) == int(z["changed"])


def test_warp_nearest(orc):
    (name, z), = golden_cases("warp")
    out = orc.warp_nearest_f64(z["src"], z["M"], (int(z["dw"]), int(z["dh"])))
    np.testing.assert_array_equal(out, z["out"])


def test_quantize(orc):
    (name, z), = golden_cases("quantize")
    np.testing.assert_array_equal(orc.quantize_u8(z["vals"]), z["out"])
