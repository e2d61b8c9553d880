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


@pytest.mark.parametrize("name,z", golden_cases("warplin"), ids=lambda v: v if isinstance(v, str) else "")
def test_warp_linear(orc, name, z):
    """cv2.warpPerspective (INTER_LINEAR, BORDER_CONSTANT 0, CV_8UC3) goldens (rectification.cpp:486-493): all 1024
    weight-table entries, borders, strong perspective, vanishing denominator, singular matrices, saturated coordinates,
    a shipped-image crop."""
    out = orc.warp_linear_bgr(z["src"], z["M"], (int(z["dw"]), int(z["dh"])))
    np.testing.assert_array_equal(out, z["out"])


def test_warp_linear_goldens_cover_every_table_entry():
    """The `all_fractions` case lands on each of the 32 x 32 sub-pixel fractions at least once (recomputed here in the
    same arithmetic), so the closed-form weights stand for the whole BilinearTab_i."""
    z = dict(golden_cases("warplin"))["warplin_all_fractions"]
    m = np.linalg.inv(z["M"])
    seen = set()
    for y in range(int(z["dh"])):
        for x in range(int(z["dw"])):
            w = m[2, 0] * x + m[2, 1] * y + m[2, 2]
            X = int(np.rint((m[0, 0] * x + m[0, 1] * y + m[0, 2]) * 32 / w)); Y = int(np.rint((m[1, 0] * x + m[1, 1] * y + m[1, 2]) * 32 / w))
            if 0 <= X >> 5 < z["src"].shape[1] - 1 and 0 <= Y >> 5 < z["src"].shape[0] - 1:
                seen.add((Y & 31, X & 31))
    assert len(seen) == 1024


@pytest.mark.parametrize("name,z", golden_cases("warp"), ids=lambda v: v if isinstance(v, str) else "")
def test_warp_nearest(orc, name, z):
    """cv2.warpPerspective(INTER_NEAREST) goldens (rectification.cpp:82-87): half-pixel landings, strong perspective,
    (near-)singular matrices, maps narrower / lower than OpenCV's 64 x 16 evaluation block."""
    out = orc.warp_nearest_f64(np.asarray(z["src"], np.float64), z["M"], (int(z["dw"]), int(z["dh"])))
    np.testing.assert_array_equal(out, np.asarray(z["out"], np.float64))


@pytest.mark.parametrize("seed", range(10))
def test_search_ext_fast_equals_literal(orc, seed):
    """The box-sum version of the extension search (used as the oracle on whole MiddEval3 pairs) equals the literal
    O(window) definition: both directions, SSD / SAD, winner, winner cost and parabola, ties, black pixels, unequal
    widths, minDisparity, thread counts."""
    rng = np.random.default_rng(7000 + seed)
    h2, w2 = int(rng.integers(6, 22)), int(rng.integers(12, 50))
    h1, w1 = h2 + int(rng.integers(0, 3)), w2 + int(rng.integers(-6, 8))
    lv = int(rng.choice([2, 4, 256]))
    R = (rng.integers(0, lv, (h2, w2, 3)) * (255 // (lv - 1))).astype(np.uint8)
    L = (rng.integers(0, lv, (h1, w1, 3)) * (255 // (lv - 1))).astype(np.uint8)
    R[rng.random((h2, w2)) < 0.05] = 0
    L[rng.random((h1, w1)) < 0.05] = 0
    block = int(rng.choice([3, 4, 7, 9, 17]))
    mn = int(rng.choice([0, 0, 2]))
    mx = mn + int(rng.integers(1, 24))
    for cost in (orc.COST_SSD_L2, orc.COST_SAD):
        for lr in (False, True):
            try:
                want = orc.search_ext(L, R, block, mn, mx, cost, lr, True, True)
            except orc.OracleError:
                with pytest.raises(orc.OracleError):
                    orc.search_ext_fast(L, R, block, mn, mx, cost, lr, True, True)
                continue
            got = orc.search_ext_fast(L, R, block, mn, mx, cost, lr, True, True, nthreads=1 + seed % 3)
            for a, b in zip(got, want):
                np.testing.assert_array_equal(a, b)


def test_quantize(orc):
    (name, z), = golden_cases("quantize")
    np.testing.assert_array_equal(orc.quantize_u8(z["vals"]), z["out"])


@pytest.mark.parametrize("seed", range(8))
def test_right_fast_equals_literal_on_random_pairs(orc, seed):
    """The box-sum + OpenMP port that bench.py times as cpu_baseline / --impl reference equals the literal restatement
    of BlockSearch.cpp:88-179 (itself pinned by the cv2 goldens above) on random pairs: few grey levels (ties: the
    first minimum must win), black pixels (skipped), unequal image sizes, even / odd blocks, minDisparity, smoothFactor
    below / at / above 1, row sub-ranges and thread counts."""
    rng = np.random.default_rng(1000 + seed)
    h2, w2 = int(rng.integers(9, 30)), int(rng.integers(20, 60))
    h1, w1 = h2 + int(rng.integers(0, 4)), w2 + int(rng.integers(-6, 10))
    levels = int(rng.choice([2, 4, 256]))
    R = (rng.integers(0, levels, (h2, w2, 3)) * (255 // (levels - 1))).astype(np.uint8)
    L = (rng.integers(0, levels, (h1, w1, 3)) * (255 // (levels - 1))).astype(np.uint8)
    R[rng.random((h2, w2)) < 0.05] = 0
    block = int(rng.choice([3, 4, 7, 9, 17]))
    min_d = int(rng.choice([0, 0, 2]))
    max_d = min_d + int(rng.integers(1, 24))
    smooth = float(rng.choice([1.0, 0.9, 1.3]))
    want = orc.block_search_right(L, R, block, min_d, max_d, smooth)
    np.testing.assert_array_equal(orc.block_search_right_fast(L, R, block, min_d, max_d, smooth), want)
    np.testing.assert_array_equal(orc.block_search_right_fast(L, R, block, min_d, max_d, smooth, nthreads=3), want)
    if smooth == 1.0:                                   # a row range is independent of the rows above it only without the bias
        r0, r1 = h2 // 3, h2 - 2
        got = orc.block_search_right_fast(L, R, block, min_d, max_d, smooth, rows=(r0, r1))
        np.testing.assert_array_equal(got[r0:r1], want[r0:r1])


def test_box_blur_oracle_against_numpy_recurrence(orc):
    """A second, independent restatement of cv::blur's RowSum<float,double> / ColumnSum<double,float> recurrences
    (reconstruction.cpp:8) in plain Python on a small map with a wide dynamic range, where the order of the additions
    is visible in the result: the C oracle must follow the same order bit for bit."""
    rng = np.random.default_rng(5)
    h, w, k = 13, 17, 5
    src = (rng.standard_normal((h, w)) * 10.0 ** rng.integers(-6, 7, (h, w))).astype(np.float32)
    a = k // 2

    def r101(p, n):
        while p < 0 or p >= n:
            p = -p if p < 0 else 2 * (n - 1) - p
        return p
    rows = np.zeros((h, w), np.float64)
    for y in range(h):
        acc = 0.0
        for i in range(k):
            acc += float(src[y, r101(i - a, w)])
        rows[y, 0] = acc
        for x in range(1, w):
            acc += float(src[y, r101(x + k - 1 - a, w)]) - float(src[y, r101(x - 1 - a, w)])
            rows[y, x] = acc
    want = np.zeros((h, w), np.float32)
    scale = 1.0 / (float(k) * float(k))
    for x in range(w):
        s = 0.0
        for i in range(k - 1):
            s += rows[r101(i - a, h), x]
        for y in range(h):
            s0 = s + rows[r101(y - a + k - 1, h), x]
            want[y, x] = np.float32(s0 * scale)
            s = s0 - rows[r101(y - a, h), x]
    assert orc.box_blur(src, k).tobytes() == want.tobytes()


def _np_search_right(L, R, block, min_d, max_d, sad):
    """SURVEY.md 8c in plain numpy: the reference's window geometry, candidate order, break rule and first minimum
    (BlockSearch.cpp:105-172) on integer SAD / SSD costs, plus the sub-pixel parabola in fp32."""
    h2, w2, _ = R.shape
    h1, w1, _ = L.shape
    half = (block - 1) // 2
    disp = np.zeros((h2, w2), np.int32)
    sub = np.zeros((h2, w2), np.float32)
    f = np.float32
    for y in range(min(h1, h2)):
        for x in range(w2):
            if not R[y, x].any():
                continue
            l, r, u, dn = min(x, half), min(w2 - x - 1, half), min(y, half), min(h2 - y - 1, half)
            b = R[y - u:y + dn, x - l:x + r].astype(np.int64)
            costs = []
            for d in range(min_d, max_d):
                if x + d + r >= w1:
                    break
                a = L[y - u:y + dn, x + d - l:x + d + r].astype(np.int64)
                costs.append(int(np.abs(a - b).sum() if sad else ((a - b) ** 2).sum()))
            if not costs:
                continue
            bi = int(np.argmin(costs))                      # first minimum
            disp[y, x] = min_d + bi
            s = f(min_d + bi)
            if 1 <= bi < len(costs) - 1:
                cm, c0, cp = f(costs[bi - 1]), f(costs[bi]), f(costs[bi + 1])
                den = (cm - f(2.0) * c0) + cp
                if den > 0:
                    s = s + (cm - cp) / (f(2.0) * den)
            sub[y, x] = s
    return disp, sub


@pytest.mark.parametrize("seed", range(6))
def test_extension_oracle_against_numpy_and_reference_path(orc, seed):
    """The extension oracle (SAD / SSD integer costs, sub-pixel; no reference counterpart) against an
    independent numpy restatement, and its SSD disparities against the reference-pinned Right search with
    smoothFactor 1 (same geometry, break rule and tie-breaking: the extensions share the pinned machinery)."""
    rng = np.random.default_rng(2000 + seed)
    h, w2 = int(rng.integers(8, 14)), int(rng.integers(16, 30))
    w1 = w2 + int(rng.integers(-3, 6))
    levels = int(rng.choice([3, 256]))
    R = (rng.integers(0, levels, (h, w2, 3)) * (255 // (levels - 1))).astype(np.uint8)
    L = (rng.integers(0, levels, (h, w1, 3)) * (255 // (levels - 1))).astype(np.uint8)
    R[rng.random((h, w2)) < 0.05] = 0
    block = int(rng.choice([3, 5, 8, 9]))
    min_d = int(rng.choice([0, 1]))
    max_d = min_d + int(rng.integers(2, 12))
    for cost in (orc.COST_SSD_L2, orc.COST_SAD):
        d, _, sp = orc.search_ext(L, R, block, min_d, max_d, cost, want_subpix=True)
        want_d, want_sp = _np_search_right(L, R, block, min_d, max_d, sad=cost == orc.COST_SAD)
        np.testing.assert_array_equal(d, want_d)
        assert sp.tobytes() == want_sp.tobytes()
    ref = orc.block_search_right(L, R, block, min_d, max_d, 1.0)
    d, _, _ = orc.search_ext(L, R, block, min_d, max_d, orc.COST_SSD_L2)
    has = ref >= 0                                          # the reference stores -x where no candidate survives
    np.testing.assert_array_equal(d[has], ref[has].astype(np.int32))


@pytest.mark.parametrize("seed", range(4))
def test_extension_oracle_ncc_against_numpy(orc, seed):
    """NCC extension cost (exact integer moments, then one correctly rounded fp32 operation
    per step) restated independently in numpy float32; first candidate with the strictly smallest cost wins, flat
    windows cost +inf."""
    rng = np.random.default_rng(3000 + seed)
    h, w2 = int(rng.integers(8, 12)), int(rng.integers(16, 26))
    w1 = w2 + int(rng.integers(0, 5))
    levels = int(rng.choice([2, 256]))
    R = (rng.integers(0, levels, (h, w2, 3)) * (255 // (levels - 1))).astype(np.uint8)
    L = (rng.integers(0, levels, (h, w1, 3)) * (255 // (levels - 1))).astype(np.uint8)
    if levels == 2:
        R[2:6, 3:9] = 255                                   # flat windows: zero variance
    block, min_d, max_d = int(rng.choice([5, 7])), 0, int(rng.integers(3, 10))
    half = (block - 1) // 2
    f = np.float32
    want = np.zeros((h, w2), np.int32)
    for y in range(h):
        for x in range(w2):
            if not R[y, x].any():
                continue
            l, r, u, dn = min(x, half), min(w2 - x - 1, half), min(y, half), min(h - y - 1, half)
            b = R[y - u:y + dn, x - l:x + r].astype(np.int64)
            n = b.size
            best, bi = f(np.inf), -1
            for k, d in enumerate(range(min_d, max_d)):
                if x + d + r >= w1:
                    break
                a = L[y - u:y + dn, x + d - l:x + d + r].astype(np.int64)
                sl, sr, sll, srr, slr = int(a.sum()), int(b.sum()), int((a * a).sum()), int((b * b).sum()), int((a * b).sum())
                num, vl, vr = n * slr - sl * sr, n * sll - sl * sl, n * srr - sr * sr
                if n == 0 or vl <= 0 or vr <= 0:
                    c = f(np.inf)
                else:
                    rl, rr = f(1.0) / np.sqrt(f(vl)), f(1.0) / np.sqrt(f(vr))
                    c = -((f(num) * rl) * rr)
                if c < best:
                    best, bi = c, k
            want[y, x] = min_d + bi if bi >= 0 else 0
    got, _, _ = orc.search_ext(L, R, block, min_d, max_d, orc.COST_NCC)
    np.testing.assert_array_equal(got, want)


@pytest.mark.parametrize("seed", range(4))
def test_extension_oracle_left_referenced_and_lr_check_against_numpy(orc, seed):
    """Left-referenced search ("same geometry mirrored, d >= 0") and the left-right consistency rule
    (D_R valid iff x + D_R < w1 and |D_L(y, x + D_R) - D_R| <= 1, else 0) restated in numpy."""
    rng = np.random.default_rng(4000 + seed)
    h, w = int(rng.integers(8, 12)), int(rng.integers(18, 28))
    levels = int(rng.choice([3, 256]))
    R = (rng.integers(1, levels + 1, (h, w, 3)) * (255 // levels)).astype(np.uint8)      # no black pixels
    L = (rng.integers(1, levels + 1, (h, w, 3)) * (255 // levels)).astype(np.uint8)
    block, min_d, max_d = int(rng.choice([5, 7])), 0, int(rng.integers(3, 9))
    half = (block - 1) // 2
    want_l = np.zeros((h, w), np.int32)
    for y in range(h):
        for x in range(w):
            l, r, u, dn = min(x, half), min(w - x - 1, half), min(y, half), min(h - y - 1, half)
            a = L[y - u:y + dn, x - l:x + r].astype(np.int64)
            costs = []
            for d in range(min_d, max_d):
                cx = x - d                                   # partner column in the right image
                if cx - l < 0:
                    break
                b = R[y - u:y + dn, cx - l:cx + r].astype(np.int64)
                costs.append(int(np.abs(a - b).sum()))
            want_l[y, x] = min_d + int(np.argmin(costs)) if costs else 0
    dl, _, _ = orc.search_ext(L, R, block, min_d, max_d, orc.COST_SAD, left_ref=True)
    np.testing.assert_array_equal(dl, want_l)
    dr, _, _ = orc.search_ext(L, R, block, min_d, max_d, orc.COST_SAD)
    want = np.zeros_like(dr)
    for y in range(h):
        for x in range(w):
            d = int(dr[y, x])
            if x + d < w and abs(int(dl[y, x + d]) - d) <= 1:
                want[y, x] = d
    np.testing.assert_array_equal(orc.lr_check(dr, dl), want)
