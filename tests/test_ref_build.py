"""oracle/_ref/libsr_ref.so = the reference's own BlockSearch.cpp / LinearSearch.cpp / reconstruction.cpp compiled
UNCHANGED against the header stand-ins of oracle/refstub (recipe: oracle/Makefile).  These CPU tests
  (1) hold every arithmetic function of the stand-in against the cv2 function of the same name,
  (2) pin the C oracle (oracle/sr_oracle.c) to the compiled reference on random pairs, real-image crops and the
      committed cv2 golden vectors -- so the GPU parity tests (CUDA vs oracle) are parity with the reference's code.
The .so is built in the build container (where /root/reference exists) and travels to the GPU box prebuilt."""
import os

import numpy as np
import pytest

from conftest import golden_cases

ref = pytest.importorskip("oracle.ref")
if not ref.available():
    pytest.skip("oracle/_ref not built and no reference tree", allow_module_level=True)

try:
    import cv2
except ImportError:   # pragma: no cover
    cv2 = None
needs_cv2 = pytest.mark.skipif(cv2 is None, reason="cv2 not installed")


# ---------------- (1) the stand-in's OpenCV functions against cv2 ----------------
@needs_cv2
def test_standin_absdiff_norm_vs_cv2():
    rng = np.random.default_rng(0)
    for _ in range(50):
        h, w = int(rng.integers(1, 20)), int(rng.integers(1, 20))
        a = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        b = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        d, n = ref.cv_absdiff_norm(a, b)
        cd = cv2.absdiff(a, b)
        np.testing.assert_array_equal(d, cd)
        assert n == cv2.norm(cd, cv2.NORM_L2)        # bitwise: sqrt of the exact integer sum


@needs_cv2
def test_standin_mean_subtract_vs_cv2():
    """cv::mean (sum * 1/N in double) and cv::subtract(u8 Mat, Scalar) (scalar rounded half-to-even to the source depth,
    saturating subtract), incl. means that land exactly on k + 0.5 (BlockSearch.cpp:126-127)."""
    rng = np.random.default_rng(1)
    for t in range(1500):
        h, w = int(rng.integers(1, 9)), int(rng.integers(1, 9))
        a = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        if t % 3 == 0:
            a = (a // 64 * 64).astype(np.uint8)
        if t % 7 == 0 and h * w % 2 == 0:   # exact half means
            a[...] = 10
            a.reshape(-1, 3)[: h * w // 2] = 11
        m, s, n = ref.cv_mean_subtract(a)
        cm = cv2.mean(a)
        cs = cv2.subtract(a, cm)
        assert tuple(m) == tuple(cm[:3])
        np.testing.assert_array_equal(s, cs)
        assert n == cv2.norm(cs, cv2.NORM_L2)


@needs_cv2
@pytest.mark.parametrize("k", [1, 2, 3, 5, 8, 21, 50, 500])
def test_standin_blur_vs_cv2(k):
    rng = np.random.default_rng(k)
    x = (rng.random((47, 61)) * 200).astype(np.float32)
    np.testing.assert_array_equal(ref.cv_blur(x, k), cv2.blur(x, (k, k)))
    y = (rng.standard_normal((33, 29)) * 10.0 ** rng.integers(-3, 4, (33, 29))).astype(np.float32)   # order of additions visible
    np.testing.assert_array_equal(ref.cv_blur(y, k), cv2.blur(y, (k, k)))


def test_standin_roi_throws_like_cv_mat():
    assert not ref.roi_throws(10, 8, 0, 0, 10, 8)
    assert not ref.roi_throws(10, 8, 3, 2, 0, 0)          # empty ROI is legal
    assert ref.roi_throws(10, 8, -1, 0, 3, 3)
    assert ref.roi_throws(10, 8, 8, 0, 3, 3)
    assert ref.roi_throws(10, 8, 0, 6, 3, 3)


# ---------------- (2) the oracle against the compiled reference ----------------
def _rand_pair(rng, h2, w2, dh=0, dw=0, levels=256, black=0.05):
    h1, w1 = h2 + dh, w2 + dw
    R = (rng.integers(0, levels, (h2, w2, 3)) * (255 // (levels - 1))).astype(np.uint8)
    L = (rng.integers(0, levels, (h1, w1, 3)) * (255 // (levels - 1))).astype(np.uint8)
    R[rng.random((h2, w2)) < black] = 0
    L[rng.random((h1, w1)) < black] = 0
    return L, R


@pytest.mark.parametrize("seed", range(12))
def test_right_oracle_equals_reference(orc, seed):
    """computeDisparityMapRight (BlockSearch.cpp:88-179): ties, black pixels, unequal sizes, even / odd blocks,
    minDisparity, smoothFactor below / at / above 1 -- literal oracle and the box-sum port both equal the reference."""
    rng = np.random.default_rng(2000 + seed)
    h2, w2 = int(rng.integers(9, 28)), int(rng.integers(20, 56))
    L, R = _rand_pair(rng, h2, w2, int(rng.integers(0, 4)), int(rng.integers(-6, 10)), int(rng.choice([2, 4, 256])))
    block = int(rng.choice([1, 3, 4, 7, 9, 17, 21]))
    min_d = int(rng.choice([0, 0, 2]))
    max_d = min_d + int(rng.integers(1, 24))
    smooth = float(rng.choice([1.0, 0.9, 1.3, 0.5]))
    want, mb = ref.block_search_right(L, R, block, min_d, max_d, smooth, return_max_block=True)
    got, omb = orc.block_search_right(L, R, block, min_d, max_d, smooth, return_max_block=True)
    np.testing.assert_array_equal(got, want)
    assert omb == mb == block
    if block >= 2:
        np.testing.assert_array_equal(orc.block_search_right_fast(L, R, block, min_d, max_d, smooth), want)


@pytest.mark.parametrize("seed", range(10))
def test_right_varblock_oracle_equals_reference(orc, seed):
    """varBlock (BlockSearch.cpp:125-145): smooth content (windows grow), thresholds from "never grows" to "grows to
    the image"; the printed max block size is part of the result."""
    rng = np.random.default_rng(3000 + seed)
    h, w = int(rng.integers(10, 22)), int(rng.integers(24, 48))
    base = rng.integers(0, 256, (h // 4 + 2, w // 4 + 2, 3)).astype(np.float32)
    R = np.kron(base, np.ones((4, 4, 1), np.float32))[:h, :w] + rng.integers(0, [2, 6, 30][seed % 3], (h, w, 3))
    R = np.clip(R, 1, 255).astype(np.uint8)              # never black, never uniform: growth terminates
    L = np.roll(R, 3, axis=1)
    L = np.clip(L.astype(np.int32) + rng.integers(-2, 3, L.shape), 1, 255).astype(np.uint8)
    thres = [5.0, 19.0, 60.0, 150.0, 400.0][seed % 5]
    block = int(rng.choice([3, 5, 9]))
    smooth = float(rng.choice([1.0, 0.9]))
    want, mb = ref.block_search_right(L, R, block, 0, 12, smooth, True, thres, return_max_block=True)
    got, omb = orc.block_search_right(L, R, block, 0, 12, smooth, True, thres, return_max_block=True)
    np.testing.assert_array_equal(got, want)
    assert omb == mb


def test_right_h1_smaller_throws_in_both(orc):
    rng = np.random.default_rng(7)
    L, R = _rand_pair(rng, 20, 30, dh=-3, black=0)
    with pytest.raises(ref.RefError) as e:
        ref.block_search_right(L, R, 9, 0, 8, 1.0)
    assert e.value.code == ref.E_SHAPE                      # cv::Exception from the left ROI
    with pytest.raises(orc.OracleError) as e2:
        orc.block_search_right(L, R, 9, 0, 8, 1.0)
    assert e2.value.code == orc.E_SHAPE


@pytest.mark.parametrize("seed", range(8))
def test_left_oracle_equals_reference(orc, seed):
    """computeDisparityMapLeft (BlockSearch.cpp:24-86): the genuine raster dependency, largest d wins ties."""
    rng = np.random.default_rng(4000 + seed)
    h1, w1 = int(rng.integers(12, 26)), int(rng.integers(24, 50))
    L, R = _rand_pair(rng, h1 + int(rng.integers(0, 3)), w1 + int(rng.integers(-4, 6)), levels=int(rng.choice([2, 4, 256])))
    L = L[:h1, :w1] if L.shape[0] >= h1 and L.shape[1] >= w1 else np.ascontiguousarray(np.resize(L, (h1, w1, 3)))
    block = int(rng.choice([3, 5, 7, 9]))
    max_d = int(rng.integers(2, 20))
    smooth = float(rng.choice([1.0, 0.9, 0.5, 1.3]))
    want = ref.block_search_left(L, R, block, max_d, smooth)
    np.testing.assert_array_equal(orc.block_search_left(L, R, block, max_d, smooth), want)


@pytest.mark.parametrize("seed", range(4))
def test_linear_oracle_equals_reference(orc, seed):
    """LinearSearch (LinearSearch.cpp:10-59) where the reference stays in bounds (w1 >= w2 + 199)."""
    rng = np.random.default_rng(5000 + seed)
    h, w2 = int(rng.integers(6, 14)), int(rng.integers(10, 40))
    L, R = _rand_pair(rng, h, w2, dw=199 + int(rng.integers(0, 8)), levels=int(rng.choice([3, 256])))
    L = np.ascontiguousarray(L)
    smooth = float(rng.choice([1.0, 0.9, 1.2]))
    want, printed = ref.linear_search(L, R, smooth, return_printed=True)
    np.testing.assert_array_equal(orc.linear_search(L, R, smooth), want)
    assert printed.startswith(f"h1:{h}\nh2:{h}\n")          # LinearSearch.cpp:16-17 (the shim prints the same)


@pytest.mark.parametrize("name,z", golden_cases("right") + golden_cases("left"), ids=lambda v: v if isinstance(v, str) else "")
def test_reference_reproduces_cv2_goldens(name, z):
    """The committed golden vectors came from a cv2 transliteration of the reference (tests/golden/make_golden.py);
    the compiled reference reproduces them -- transliteration, stand-in and oracle all agree."""
    if str(z["kind"]) == "right":
        out, mb = ref.block_search_right(z["left"], z["right"], int(z["block"]), int(z["min_d"]), int(z["max_d"]), float(z["smooth"]),
                                         bool(z["var_block"]), float(z["thres"]), return_max_block=True)
        assert mb == int(z["max_block"])
    else:
        out = ref.block_search_left(z["left"], z["right"], int(z["block"]), int(z["max_d"]), float(z["smooth"]))
    np.testing.assert_array_equal(out, z["out"])


@pytest.mark.parametrize("seed", range(4))
def test_reconstruction_chain_oracle_equals_reference(orc, seed, tmp_path):
    """removeDisparityOutliers, scaleDisparityMap, convertDisparityToDepth, reconstruction() and WriteMesh
    (reconstruction.cpp:5-208): maps bit-identical, the COFF file byte-identical."""
    rng = np.random.default_rng(6000 + seed)
    h, w = int(rng.integers(8, 20)), int(rng.integers(10, 30))
    disp = rng.integers(0, 60, (h, w)).astype(np.float32)
    disp[rng.random((h, w)) < 0.1] = 0
    if seed % 2:
        disp += rng.random((h, w)).astype(np.float32)
    k = int(rng.choice([3, 7, 20, 500]))
    d1 = ref.remove_outliers(disp, k, 1.5, 0.8)
    np.testing.assert_array_equal(orc.remove_outliers(disp, k, 1.5, 0.8), d1)
    np.testing.assert_array_equal(orc.scale(d1, 0.37), ref.scale(d1, 0.37))
    d1[rng.random((h, w)) < 0.1] = 0
    depth = ref.disparity_to_depth(d1, 1733.74, 1.0)
    np.testing.assert_array_equal(orc.disparity_to_depth(d1, 1733.74, 1.0), depth)
    bgr = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    fx, fy, cx, cy = 1733.74, 1733.68, 792.27, 541.89
    thr = float(rng.choice([0.5, 1.0, 5.0]))
    path = ref.reconstruction(bgr, depth, fx, fy, cx, cy, thr, str(tmp_path))
    verts = orc.backproject(bgr, depth, fx, fy, cx, cy)
    mine = str(tmp_path / "oracle.off")
    orc.write_off(mine, verts, w, h, thr)
    assert open(path, "rb").read() == open(mine, "rb").read()
    # WriteMesh / CheckTriangularValidity on the oracle's vertex array (same 32-byte layout as struct Vertex)
    theirs = str(tmp_path / "ref_from_oracle_vertices.off")
    ref.write_mesh(theirs, verts, w, h, thr)
    assert open(theirs, "rb").read() == open(mine, "rb").read()
    faces = orc.mesh_faces(verts, w, h, thr)
    lines = open(theirs).read().split("\n")
    assert lines[1] == f"{w * h} {len(faces)} 0"
    for a, b, c in [(0, w, 1), (w, w + 1, 1), (w * (h - 2), w * (h - 1), w * (h - 2) + 1)]:
        assert ref.triangle_valid(verts, a, b, c, thr) == bool(
            orc.lib().orc_triangle_valid(verts.ctypes.data_as(__import__("ctypes").c_void_p), a, b, c, __import__("ctypes").c_float(thr)))


def test_rows_harness_equals_whole_image(orc):
    """bench.py's reference arm cuts the output rows into bands, one BlockSearch per band on a crop with halo rows:
    the same map as one BlockSearch over the whole pair (smoothFactor 1)."""
    L, R = orc.synth_pair(0xB200, 0, 96, 40, 0, 16)
    want = ref.block_search_right(L, R, 9, 0, 16, 1.0)
    for nt in (1, 3, 8):
        got, used = ref.block_search_right_rows(L, R, 9, 0, 16, (0, 40), nt)
        np.testing.assert_array_equal(got, want)
        assert 1 <= used <= nt
    got, _ = ref.block_search_right_rows(L, R, 9, 0, 16, (7, 29), 4)
    np.testing.assert_array_equal(got[7:29], want[7:29])
