"""Parity of the remaining CUDA entry points (Left / Linear searches, extensions, Reconstruction,
warp, evaldisp, batch API) with the CPU oracle and the cv2-executed golden vectors.  Integer,
index and the reference's fp32/fp64 expressions: bit-exact.  NCC / sub-pixel: within 1e-3 px
(north_star), and in fact compared bit-exact except where two NCC costs tie within 1e-6."""
import os

import numpy as np
import pytest

from conftest import golden_cases

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def sr():
    import stereo_reconstruction_b200 as s
    return s


@pytest.fixture(scope="module")
def ctx(sr):
    return sr.capi.default_context()


@pytest.mark.parametrize("name,z", golden_cases("left"), ids=lambda v: v if isinstance(v, str) else "")
def test_left_golden(sr, ctx, name, z):
    out = sr.BlockSearch(z["left"], z["right"], int(z["block"]), 0, int(z["max_d"]), ctx=ctx).computeDisparityMapLeft(float(z["smooth"]))
    np.testing.assert_array_equal(out, z["out"])


@pytest.mark.parametrize("smooth", [1.0, 0.9, 1.3])
def test_left_vs_oracle(sr, ctx, orc, smooth):
    L, R = orc.synth_pair(11, 0, 260, 70, 0, 30, True)   # quantised: many ties, the largest d must win
    L[20:30, 100:120] = 0
    want = orc.block_search_left(L, R, 7, 30, smooth)
    got = sr.BlockSearch(L, R, 7, 0, 30, ctx=ctx).computeDisparityMapLeft(smooth)
    np.testing.assert_array_equal(got, want)


@pytest.mark.parametrize("w1,w2,h1,h2,block,max_d,smooth,quant", [
    (1100, 1100, 40, 40, 5, 70, 0.9, False),     # three column tiles, two passes, range not a multiple of 4
    (600, 600, 50, 50, 17, 40, 0.9, False),      # the largest window of the tile path
    (300, 340, 45, 40, 9, 33, 1.0, False),       # right image wider and shorter
    (340, 300, 40, 45, 3, 130, 0.8, False),      # left image wider: partners beyond w2 - h are no candidates; three passes
    (120, 120, 30, 30, 11, 200, 0.9, False),     # range wider than the image
    (260, 260, 40, 40, 7, 30, 0.5, True),        # quantised colours: ties, the largest d wins; strong bias
    (260, 260, 40, 40, 7, 30, -1.0, True),       # negative factor: biased distances below every unbiased one
    (530, 530, 36, 36, 13, 64, 1.3, False),      # penalty (> 1): three entries per pixel from the direct kernel
    (64, 64, 20, 20, 19, 16, 0.9, False),        # window beyond 17: direct kernel
])
def test_left_variant_tile_path(sr, ctx, orc, w1, w2, h1, h2, block, max_d, smooth, quant):
    """computeDisparityMapLeft (BlockSearch.cpp:24-86): unbiased winners from the tile kernel (odd windows 3..17, candidate
    index off - d so that the first minimum is the largest d) + the raster rule as a multi-CTA wavefront, against the
    literal oracle, on shapes that exercise every bound of the candidate range."""
    w, h = max(w1, w2), max(h1, h2)
    L, R = orc.synth_pair(0xB2 + block, 3, w, h, 0, min(max_d, 60), quant)
    L = np.ascontiguousarray(L[:h1, :w1]); R = np.ascontiguousarray(R[:h2, :w2])
    L[h1 // 2:h1 // 2 + 4, w1 // 3:w1 // 3 + 25] = 0          # black pixels are skipped (and stay 0)
    want = orc.block_search_left(L, R, block, max_d, smooth)
    bs = sr.BlockSearch(L, R, block, 0, max_d, ctx=ctx)
    got = bs.computeDisparityMapLeft(smooth)
    np.testing.assert_array_equal(got, want)
    np.testing.assert_array_equal(bs.computeDisparityMapLeft(smooth), want)   # again on the same pair (buffers reused)


@pytest.mark.parametrize("grid", ["1", "2"])
def test_left_variant_wavefront_rounds(sr, ctx, orc, monkeypatch, grid):
    """The row wavefront with fewer CTAs than rows / 16: every warp takes several rows (rounds), the hand-over between the
    last warp of a CTA and warp 0 of the next one (or of the next round) goes through global memory, all others through
    the shared-memory ring -- which the leader may not overrun (rows wider than the ring)."""
    monkeypatch.setenv("SR_LEFT_GRID", grid)
    L, R = orc.synth_pair(77, 1, 300, 80, 0, 24, True)
    want = orc.block_search_left(L, R, 5, 24, 0.9)
    got = sr.BlockSearch(L, R, 5, 0, 24, ctx=ctx).computeDisparityMapLeft(0.9)
    np.testing.assert_array_equal(got, want)


@pytest.mark.parametrize("name,z", golden_cases("linear"), ids=lambda v: v if isinstance(v, str) else "")
def test_linear_golden(sr, ctx, name, z):
    out = sr.LinearSearch(z["left"], z["right"], ctx=ctx).computeDisparityMap(float(z["smooth"]))
    np.testing.assert_array_equal(out, z["out"])


def test_linear_vs_oracle(sr, ctx, orc):
    L, R = orc.synth_pair(5, 0, 500, 40, 0, 150)
    for smooth in (1.0, 0.7):
        np.testing.assert_array_equal(sr.LinearSearch(L, R, ctx=ctx).computeDisparityMap(smooth), orc.linear_search(L, R, smooth))


# ---- extensions -------------------------------------------------------------------------------------
@pytest.mark.parametrize("cost", [0, 1])
@pytest.mark.parametrize("block", [7, 9, 19])
def test_ext_subpixel_int_costs(sr, ctx, orc, cost, block):
    L, R = orc.synth_pair(21, 0, 520, 64, 0, 48)
    R[5:9, 40:60] = 0
    d, _, sp = orc.search_ext(L, R, block, 0, 48, cost, want_subpix=True)
    got = sr.BlockSearch(L, R, block, 0, 48, cost=cost, ctx=ctx).computeDisparityMapRight(1.0, subpixel=True)
    np.testing.assert_array_equal(got, sp.astype(np.float64))       # same fp32 expression: bit-exact
    assert np.abs(got - sp).max() <= 1e-3


def test_ext_ncc(sr, ctx, orc):
    L, R = orc.synth_pair(22, 0, 300, 48, 0, 32)
    d, _, sp = orc.search_ext(L, R, 7, 0, 32, orc.COST_NCC, want_subpix=True)
    got = sr.BlockSearch(L, R, 7, 0, 32, cost=sr.capi.COST_NCC, ctx=ctx).computeDisparityMapRight(1.0)
    assert (got != d).sum() == 0
    gots = sr.BlockSearch(L, R, 7, 0, 32, cost=sr.capi.COST_NCC, ctx=ctx).computeDisparityMapRight(1.0, subpixel=True)
    assert np.abs(gots - sp).max() <= 1e-3                          # north_star tolerance


@pytest.mark.parametrize("cost", [0, 1, 2])
def test_ext_lr_check(sr, ctx, orc, cost):
    L, R = orc.synth_pair(23, 0, 280, 50, 0, 40)
    dr, _, _ = orc.search_ext(L, R, 9, 0, 40, cost)
    dl, _, _ = orc.search_ext(L, R, 9, 0, 40, cost, left_ref=True)
    want = orc.lr_check(dr, dl)
    got = sr.BlockSearch(L, R, 9, 0, 40, cost=cost, ctx=ctx).computeDisparityMapRight(1.0, lr_check=True)
    np.testing.assert_array_equal(got, want.astype(np.float64))


# ---- Reconstruction -----------------------------------------------------------------------------------
@pytest.mark.parametrize("name,z", golden_cases("outliers"), ids=lambda v: v if isinstance(v, str) else "")
def test_outliers_golden(sr, ctx, name, z):
    R = sr.reconstruction
    np.testing.assert_array_equal(R.boxBlur(z["disp"], int(z["k"]), ctx=ctx), z["blur"])
    d = np.array(z["disp"], dtype=np.float32)
    R.removeDisparityOutliers(d, int(z["k"]), float(z["front"]), float(z["back"]), ctx=ctx)
    np.testing.assert_array_equal(d, z["out"])


def test_outliers_full_golden(sr, ctx):
    (name, z), = golden_cases("outliers_full")
    disp = z["disp_u8"].astype(np.float32)
    d = disp.copy()
    sr.reconstruction.removeDisparityOutliers(d, int(z["k"]), float(z["front"]), float(z["back"]), ctx=ctx)
    np.testing.assert_array_equal(d[::25, ::25], z["out_sample"])
    assert float(d.astype(np.float64).sum()) == float(z["out_sum"])
    assert int((d != disp).sum()) == int(z["changed"])


def _recon_inputs(orc, w=310, h=200):
    rng = np.random.default_rng(9)
    bgr = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    disp = rng.integers(0, 60, (h, w)).astype(np.float32)
    disp[rng.random((h, w)) < 0.1] = 0            # invalid -> -inf depth
    disp += (rng.random((h, w)) < 0.3) * rng.random((h, w)).astype(np.float32)
    K = np.array([[1733.74, 0, 792.27], [0, 1733.74, 541.89], [0, 0, 1]], np.float32)
    return bgr, disp.astype(np.float32), K


def test_depth_backproject_faces(sr, ctx, orc):
    R = sr.reconstruction
    bgr, disp, K = _recon_inputs(orc)
    h, w = disp.shape
    s = disp.copy(); R.scaleDisparityMap(s, 1.7, ctx=ctx)
    np.testing.assert_array_equal(s, orc.scale(disp, 1.7))
    depth = R.convertDisparityToDepth(disp, K[0, 0], 1.0, ctx=ctx)
    np.testing.assert_array_equal(depth, orc.disparity_to_depth(disp, K[0, 0], 1.0))
    v = R.backproject(bgr, depth, K, ctx=ctx)
    want_v = orc.backproject(bgr, depth, K[0, 0], K[1, 1], K[0, 2], K[1, 2])
    assert v.tobytes() == want_v.tobytes()                          # 32-byte Vertex layout, bit for bit
    for thr in (0.5, 5.0, 1e9):
        np.testing.assert_array_equal(R.meshFaces(v, w, h, thr, ctx=ctx), orc.mesh_faces(want_v, w, h, thr))
    assert R.CheckTriangularValidity(v, 0, w, 1, 1e9, ctx=ctx) == bool(orc.lib().orc_triangle_valid(
        want_v.ctypes.data, 0, w, 1, __import__("ctypes").c_float(1e9)))


def test_fused_pipeline_and_mesh_file(sr, ctx, orc, tmp_path):
    R = sr.reconstruction
    bgr, disp, K = _recon_inputs(orc, 180, 120)
    h, w = disp.shape
    got = R.reconstructFused(bgr, disp, K, K[0, 0], 1.0, kernelSize=31, thrFront=1.5, thrBack=0.8, thrMesh=3.0, ctx=ctx)
    d = orc.remove_outliers(disp, 31, 1.5, 0.8)
    depth = orc.disparity_to_depth(d, K[0, 0], 1.0)
    v = orc.backproject(bgr, depth, K[0, 0], K[1, 1], K[0, 2], K[1, 2])
    np.testing.assert_array_equal(got["disp"], d)
    np.testing.assert_array_equal(got["depth"], depth)
    assert got["vertices"].tobytes() == v.tobytes()
    np.testing.assert_array_equal(got["faces"], orc.mesh_faces(v, w, h, 3.0))
    a, b = tmp_path / "gpu.off", tmp_path / "cpu.off"
    assert R.WriteMesh(got["vertices"], w, h, a, 3.0, ctx=ctx)
    orc.write_off(str(b), v, w, h, 3.0)
    assert a.read_bytes() == b.read_bytes()                         # byte-identical COFF text
    assert R.WriteMesh(got["vertices"], w, h, tmp_path / "no_such_dir" / "x.off", 3.0, ctx=ctx) is False
    R.writePFM(tmp_path / "g.pfm", got["disp"]); orc.write_pfm(str(tmp_path / "c.pfm"), d)
    assert (tmp_path / "g.pfm").read_bytes() == (tmp_path / "c.pfm").read_bytes()
    R.writePLY(tmp_path / "g.ply", got["vertices"], w, h); orc.write_ply(str(tmp_path / "c.ply"), v, w, h)
    assert (tmp_path / "g.ply").read_bytes() == (tmp_path / "c.ply").read_bytes()


def test_pair_reconstruct_device_resident(sr, ctx, orc):
    """main.cpp:40-65 without a host round trip of the disparity map."""
    L, Rr = orc.synth_pair(31, 0, 400, 90, 0, 60)
    K = np.array([[500.0, 0, 200.0], [0, 500.0, 45.0], [0, 0, 1]], np.float32)
    pair = sr.capi.Pair(ctx, 400, 90, 400, 90)
    pair.upload(L, Rr)
    p = sr.capi.search_params(block=17, min_disp=0, max_disp=60, smooth=0.9)
    pair.run(p)
    got = sr.reconstruction.pairReconstruct(pair, K, K[0, 0], 1.0, True, 25, 1.5, 0.8, 2.0)
    disp64 = orc.block_search_right_fast(L, Rr, 17, 0, 60, 0.9)
    d = orc.remove_outliers(orc.quantize_u8(disp64), 25, 1.5, 0.8)
    depth = orc.disparity_to_depth(d, K[0, 0], 1.0)
    v = orc.backproject(Rr, depth, K[0, 0], K[1, 1], K[0, 2], K[1, 2])
    np.testing.assert_array_equal(got["disp"], d)
    assert got["vertices"].tobytes() == v.tobytes()
    np.testing.assert_array_equal(got["faces"], orc.mesh_faces(v, 400, 90, 2.0))


@pytest.mark.parametrize("w,h,k", [(900, 750, 500), (400, 90, 25), (333, 61, 7), (64, 40, 1), (1436, 992, 500), (50, 30, 200)])
def test_u8_handoff_pipeline_integer_exact(sr, ctx, orc, monkeypatch, w, h, k):
    """main.cpp:41-65 for the 8-bit hand-off runs integer-exact (recon.cu: u8 row sums, int32 column prefixes, blurred
    value + outlier rule inside the depth / vertex / face kernel): every output equals the oracle chain
    quantize -> removeDisparityOutliers(k) -> convertDisparityToDepth -> reconstruction -> WriteMesh, and the same
    outputs come out of the general (fp64 prefix) kernels.  Kernel sizes from 1 to larger than the map (reflect-101
    folds several times), values that round / saturate at the hand-off."""
    rng = np.random.default_rng(w * 31 + k)
    Rr = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    L = np.ascontiguousarray(np.roll(Rr, 5, axis=1))
    K = np.array([[500.0, 0, w / 2], [0, 500.0, h / 2], [0, 0, 1]], np.float32)
    pair = sr.capi.Pair(ctx, w, h, w, h)
    pair.upload(L, Rr)
    p = sr.capi.search_params(block=9, min_disp=0, max_disp=min(300, w - 20), smooth=1.0, cost=sr.capi.COST_SAD)
    pair.run(p)
    disp64 = pair.download(p)
    got = sr.reconstruction.pairReconstruct(pair, K, K[0, 0], 1.0, True, k, 1.5, 0.8, 2.0)
    n0 = ctx.launches
    sr.reconstruction.pairReconstruct(pair, K, K[0, 0], 1.0, True, k, 1.5, 0.8, 2.0)
    assert ctx.launches - n0 == 7          # rowsum, seg, prefix, fused depth kernel, block totals, scan, scatter
    d = orc.remove_outliers(orc.quantize_u8(disp64), k, 1.5, 0.8)
    depth = orc.disparity_to_depth(d, K[0, 0], 1.0)
    v = orc.backproject(Rr, depth, K[0, 0], K[1, 1], K[0, 2], K[1, 2])
    np.testing.assert_array_equal(got["disp"], d)
    np.testing.assert_array_equal(got["depth"], depth)
    assert got["vertices"].tobytes() == v.tobytes()
    np.testing.assert_array_equal(got["faces"], orc.mesh_faces(v, w, h, 2.0))
    monkeypatch.setenv("SR_RECON_GENERAL", "1")
    gen = sr.reconstruction.pairReconstruct(pair, K, K[0, 0], 1.0, True, k, 1.5, 0.8, 2.0)
    for key in ("disp", "depth", "faces"):
        np.testing.assert_array_equal(gen[key], got[key])
    assert gen["vertices"].tobytes() == got["vertices"].tobytes()
    pair.close()


@pytest.mark.parametrize("name,z", golden_cases("warplin"), ids=lambda v: v if isinstance(v, str) else "")
def test_warp_linear_golden(sr, ctx, name, z):
    """The rectifying warp of the input images (rectification.cpp:486-493) against cv2.warpPerspective goldens."""
    out = sr.reconstruction.warpPerspectiveLinear(z["src"], z["M"], (int(z["dw"]), int(z["dh"])), ctx=ctx)
    np.testing.assert_array_equal(out, z["out"])


def test_pair_upload_warped_equals_host_rectified_pair(sr, ctx, orc):
    """rectification_main.cpp's order on the device: un-rectified images -> rectifying warp (H, Hp) -> search.  The
    disparity map equals the one from images rectified on the host (oracle warp) and uploaded the plain way."""
    L, R = orc.synth_pair(0xB200, 4, 400, 120, 0, 40)
    H = np.array([[1.01, 0.02, -3.0], [-0.015, 0.99, 2.0], [2e-5, -1e-5, 1.0]])
    Hp = np.array([[0.99, -0.01, 2.5], [0.02, 1.01, -1.5], [-1e-5, 3e-5, 1.0]])
    w1, h1, w2, h2 = 410, 124, 396, 118
    Lr = orc.warp_linear_bgr(L, H, (w1, h1)); Rr = orc.warp_linear_bgr(R, Hp, (w2, h2))
    p = sr.capi.search_params(block=9, min_disp=0, max_disp=40, smooth=1.0)
    a = sr.capi.Pair(ctx, w1, h1, w2, h2); a.upload(Lr, Rr); a.run(p); want = a.download(p)
    b = sr.capi.Pair(ctx, w1, h1, w2, h2); b.upload_warped(L, H, R, Hp); b.run(p); got = b.download(p)
    np.testing.assert_array_equal(got, want)
    np.testing.assert_array_equal(want, orc.block_search_right_fast(Lr, Rr, 9, 0, 40, 1.0))
    a.close(); b.close()


def test_warp_linear_vs_oracle_large_and_strided(sr, ctx, orc):
    """A 1436 x 992 image (the Teddy pair's size class) under a rectification-like homography, odd output width (byte
    stores instead of words at the row ends), and a source view with a row stride."""
    rng = np.random.default_rng(5)
    big = rng.integers(0, 256, (992, 1500, 3)).astype(np.uint8)
    src = big[:, 32:1468]                                            # strided view, 1436 wide
    M = np.array([[0.97, -0.03, 25.0], [0.02, 1.02, -11.0], [-4e-5, 1.5e-5, 1.0]])
    for dsize in ((1501, 1011), (1436, 992), (3, 2)):
        want = orc.warp_linear_bgr(np.ascontiguousarray(src), M, dsize)
        got = sr.reconstruction.warpPerspectiveLinear(src, M, dsize, ctx=ctx)
        np.testing.assert_array_equal(got, want)
    L2, R2 = sr.reconstruction.rectifyImages(src, src[:, ::-1], M, np.linalg.inv(M), (700, 500), (640, 480), ctx=ctx)
    np.testing.assert_array_equal(L2, orc.warp_linear_bgr(np.ascontiguousarray(src), M, (700, 500)))
    np.testing.assert_array_equal(R2, orc.warp_linear_bgr(np.ascontiguousarray(src[:, ::-1]), np.linalg.inv(M), (640, 480)))


@pytest.mark.parametrize("name,z", golden_cases("warp"), ids=lambda v: v if isinstance(v, str) else "")
def test_warp_golden(sr, ctx, orc, name, z):
    """cv2.warpPerspective(INTER_NEAREST) goldens incl. half-pixel landings, strong perspective, (near-)singular M."""
    out = sr.reconstruction.warpPerspectiveNearest(np.asarray(z["src"], np.float64), z["M"], (int(z["dw"]), int(z["dh"])), ctx=ctx)
    np.testing.assert_array_equal(out, np.asarray(z["out"], np.float64))


def test_average_disparity(sr, ctx, orc):
    """computeAverageDisparity (utils.cpp:8-20): one float accumulator in raster order -- bit-identical, not 'close'."""
    rng = np.random.default_rng(11)
    for shape in ((37, 53), (750, 900), (1, 1)):
        d = (rng.random(shape) * 200).astype(np.float32)
        d[rng.random(shape) < 0.1] = np.inf
        d[rng.random(shape) < 0.02] = -np.inf
        got, cnt = sr.reconstruction.computeAverageDisparity(d, ctx=ctx, return_count=True)
        want = np.float32(orc.average_disparity(d))
        assert cnt == int(np.isfinite(d).sum())
        assert got.tobytes() == want.tobytes(), (got, want)


def test_pipeline_with_inverse_warp(sr, ctx, orc):
    """rectification.cpp:78-88 -> main.cpp:41-65 without the disparity map leaving HBM (sr_pair_reconstruct_warp):
    search on the rectified pair -> warpPerspective(H_.inv(), INTER_NEAREST) into the original right image's frame ->
    8-bit hand-off -> removeDisparityOutliers -> depth -> vertices (coloured from the un-rectified image) -> faces."""
    capi = sr.capi
    w, h, D = 420, 300, 40
    L, R = orc.synth_pair(0xB200, 3, w, h, 0, D)
    dw, dh = 400, 310
    color = np.random.default_rng(5).integers(0, 256, (dh, dw, 3), dtype=np.uint8)
    Hinv = np.array([[1.0021, 0.0153, -6.7], [-0.0112, 0.9987, 4.3], [1.3e-5, -2.1e-6, 1.0]])
    pair = capi.Pair(ctx, w, h, w, h)
    pair.upload(L, R)
    p = capi.search_params(block=17, min_disp=0, max_disp=D, smooth=0.9)
    pair.run(p)
    K = np.array([[1733.74, 0, 792.27], [0, 1733.68, 541.89], [0, 0, 1]], np.float32)
    got = sr.reconstruction.pairReconstructWarp(pair, Hinv, (dw, dh), color, K, float(K[0, 0]), 1.0, quantize_u8=True, kernelSize=50)
    disp = orc.block_search_right_fast(L, R, 17, 0, D, 0.9)
    warped = orc.warp_nearest_f64(disp, Hinv, (dw, dh))
    np.testing.assert_array_equal(got["warped"], warped)
    d8 = orc.quantize_u8(warped)
    filt = orc.remove_outliers(d8, 50, 1.5, 0.8)
    np.testing.assert_array_equal(got["disp"], filt)
    depth = orc.disparity_to_depth(filt, float(K[0, 0]), 1.0)
    np.testing.assert_array_equal(got["depth"], depth)
    verts = orc.backproject(color, depth, float(K[0, 0]), float(K[1, 1]), float(K[0, 2]), float(K[1, 2]))
    assert got["vertices"].tobytes() == verts.tobytes()
    np.testing.assert_array_equal(got["faces"], orc.mesh_faces(verts, dw, dh, 1.0))
    # colour image omitted: the warped map must have the pair's size and takes the pair's right image
    got2 = sr.reconstruction.pairReconstructWarp(pair, np.eye(3), (w, h), None, K, float(K[0, 0]), 1.0, quantize_u8=True, kernelSize=50)
    same = sr.reconstruction.pairReconstruct(pair, K, float(K[0, 0]), 1.0, quantize_u8=True, kernelSize=50)
    for k in ("disp", "depth", "faces"):
        np.testing.assert_array_equal(got2[k], same[k])
    assert got2["vertices"].tobytes() == same["vertices"].tobytes()
    pair.close()


def test_evaldisp(sr, ctx, orc):
    rng = np.random.default_rng(4)
    h, w = 120, 200
    gt = (rng.random((h, w)) * 60).astype(np.float32); gt[rng.random((h, w)) < 0.05] = np.inf
    disp = (gt + rng.normal(0, 2, (h, w))).astype(np.float32); disp[~np.isfinite(disp)] = 3
    disp[rng.random((h, w)) < 0.1] = 0
    mask = np.where(rng.random((h, w)) < 0.8, 255, 128).astype(np.uint8)
    for rd in (0, 1):
        want = orc.evaldisp(disp, gt, mask, 2.0, 64.0, rd)
        got = sr.reconstruction.evaldisp(disp, gt, mask, 2.0, 64.0, rd, ctx=ctx)
        assert (got["n"], got["bad"], got["invalid"]) == (want.n, want.bad, want.invalid)   # counts exact
        assert got["bad_pct"] == want.bad_pct and got["invalid_pct"] == want.invalid_pct
        assert abs(got["avg_err"] - want.avg_err) <= 1e-4 * max(1.0, want.avg_err)           # float sum order differs


def test_batch_api(sr, ctx, orc):
    import ctypes as C
    capi = sr.capi
    n, w, h, D = 7, 384, 72, 48
    pairs = [orc.synth_pair(0xB200, i, w, h, 0, D) for i in range(n)]
    Ls = (capi.SrImage * n)(); Rs = (capi.SrImage * n)()
    keep = []
    outs = [np.zeros((h, w), np.float64) for _ in range(n)]
    for i, (L, R) in enumerate(pairs):
        la, li = capi.image(L); ra, ri = capi.image(R); keep += [la, ra]
        Ls[i] = li; Rs[i] = ri
    optrs = (C.c_void_p * n)(*[o.ctypes.data for o in outs])
    p = capi.search_params(block=7, min_disp=0, max_disp=D, smooth=1.0, cost=capi.COST_SAD, subpixel=True)
    ctx.check(ctx._l.sr_block_search_batch(ctx.h, n, Ls, Rs, C.byref(p), optrs, C.c_size_t(w * 8)))
    for i, (L, R) in enumerate(pairs):
        _, _, sp = orc.search_ext(L, R, 7, 0, D, orc.COST_SAD, want_subpix=True)
        np.testing.assert_array_equal(outs[i], sp.astype(np.float64))


@pytest.mark.parametrize("subpixel,cost", [(True, 1), (False, 1), (False, 0), (True, 0)])
def test_batch_multi_pair_launch(sr, ctx, orc, subpixel, cost):
    """Batches of small pairs share ONE launch of the fused cost + WTA kernel (work items carry the pair index):
    sr_search_run_batch on resident pairs and the grouped pipeline of sr_block_search_batch (pinned buffers, >= 16
    pairs) give each pair exactly what sr_search_run gives it alone -- and what the oracle says."""
    import ctypes as C
    capi = sr.capi
    n, w, h, D = 19, 300, 66, 40
    data = [orc.synth_pair(0xB200, 100 + i, w, h, 0, D, quant=(i % 3 == 0)) for i in range(n)]
    data[2][1][10:20, 50:90] = 0
    p = capi.search_params(block=7, min_disp=0, max_disp=D, smooth=1.0, cost=cost, subpixel=subpixel)
    want = []
    for L, R in data:
        if subpixel:
            want.append(orc.search_ext_fast(L, R, 7, 0, D, cost, want_subpix=True)[2].astype(np.float64))
        else:
            want.append(orc.block_search_right_fast(L, R, 7, 0, D, 1.0, cost=cost))
    # resident form
    pairs = [capi.Pair(ctx, w, h, w, h) for _ in range(n)]
    for q, (L, R) in zip(pairs, data):
        q.upload(L, R)
    capi.search_run_batch(ctx, pairs, p)         # first use of the two group geometries: the schedule autotuner runs here
    n0 = ctx.launches
    capi.search_run_batch(ctx, pairs, p)
    # 19 pairs = groups of 16 + 3: two tile launches + one sub-pixel pass per group, or one resolve pass per pair
    assert ctx.launches - n0 == (2 + 2 if subpixel else 2 + n)
    for q, wnt in zip(pairs, want):
        np.testing.assert_array_equal(q.download(p), wnt)
    for q in pairs:
        q.close()
    # host form: grouped pipeline with pinned buffers
    bufs, Ls, Rs, outs = [], (capi.SrImage * n)(), (capi.SrImage * n)(), []
    for i, (L, R) in enumerate(data):
        Lp, a = capi.pinned_empty(L.shape, np.uint8); Rp, b = capi.pinned_empty(R.shape, np.uint8); o, c = capi.pinned_empty((h, w), np.float64)
        Lp[...] = L; Rp[...] = R; o[...] = -5.0
        bufs += [a, b, c]
        _, Ls[i] = capi.image(Lp); _, Rs[i] = capi.image(Rp)
        outs.append((Lp, Rp, o))
    try:
        optrs = (C.c_void_p * n)(*[o.ctypes.data for _, _, o in outs])
        ctx.check(ctx._l.sr_block_search_batch(ctx.h, n, Ls, Rs, C.byref(p), optrs, C.c_size_t(w * 8)))
        for (_, _, o), wnt in zip(outs, want):
            np.testing.assert_array_equal(np.asarray(o), wnt)
    finally:
        for x in bufs:
            capi.free_pinned(x)


@pytest.mark.parametrize("cost", [0, 1, 2])
@pytest.mark.parametrize("shape", [
    # (w1, h1, w2, h2, block, min_d, max_d, quant)
    (280, 50, 280, 50, 9, 0, 40, False),
    (280, 50, 280, 50, 8, 0, 40, True),      # even block, frequent ties
    (523, 61, 523, 61, 17, 2, 70, False),    # min_disp > 0, widest tile window
    (300, 48, 340, 48, 5, 0, 33, True),      # w1 < w2: still the mirrored tile search
    (340, 48, 300, 48, 7, 0, 33, False),     # w1 > w2: per-pixel fallback (partner windows beyond the right image)
])
def test_ext_left_referenced_map(sr, ctx, orc, cost, shape):
    """The left-referenced search behind the LR check (mirror-packed tile search) equals the oracle's
    left_ref search pixel for pixel -- not only where the consistency test looks."""
    w1, h1, w2, h2, block, dlo, dhi, quant = shape
    w, h = max(w1, w2), max(h1, h2)
    L, R = orc.synth_pair(29, 0, w, h, dlo, dhi, quant)
    L = np.ascontiguousarray(L[:h1, :w1]); R = np.ascontiguousarray(R[:h2, :w2])
    L[7:12, 100:130] = 0                        # black reference pixels on the left image are skipped
    R[20:25, 50:90] = 0
    capi = sr.capi
    pair = capi.Pair(ctx, w1, h1, w2, h2)
    pair.upload(L, R)
    p = capi.search_params(block=block, min_disp=dlo, max_disp=dhi, smooth=1.0, cost=cost, lr_check=True)
    pair.run(p)
    got = pair.download(p)
    got_l = pair.left_disparity()
    dr, _, _ = orc.search_ext(L, R, block, dlo, dhi, cost)
    dl, _, _ = orc.search_ext(L, R, block, dlo, dhi, cost, left_ref=True)
    np.testing.assert_array_equal(got_l, dl)
    np.testing.assert_array_equal(got, orc.lr_check(dr, dl).astype(np.float64))
    if w1 <= w2:
        assert "mirror" in ctx.last_kernel
    pair.close()


@pytest.mark.parametrize("shape", [
    # (w1, h1, w2, h2, block, min_d, max_d, quant)
    (420, 70, 420, 70, 9, 0, 100, False),     # two passes of 64 disparities
    (420, 70, 420, 70, 5, 3, 40, True),       # min_disp > 0, quantised images (flat windows: zero variance -> +inf costs)
    (333, 64, 333, 64, 15, 0, 70, False),     # halo from two neighbour lanes
    (333, 64, 333, 64, 17, 0, 30, False),
    (300, 48, 340, 48, 8, 0, 50, False),      # even block, w1 < w2
    (340, 48, 300, 48, 3, 0, 20, True),       # w1 > w2, smallest window
])
def test_ext_ncc_sliding_kernel(sr, ctx, orc, shape):
    """NCC through the sliding-sum kernel (interior) + per-pixel kernel (border frame) against the oracle:
    integer winners and sub-pixel values, compared bit-exact (same fp32 operation order)."""
    w1, h1, w2, h2, block, dlo, dhi, quant = shape
    w, h = max(w1, w2), max(h1, h2)
    L, R = orc.synth_pair(31, 0, w, h, dlo, dhi, quant)
    L = np.ascontiguousarray(L[:h1, :w1]); R = np.ascontiguousarray(R[:h2, :w2])
    R[20:25, 50:90] = 0
    L[30:40, 200:230] = 77                      # a flat patch: zero left variance
    d, _, sp = orc.search_ext(L, R, block, dlo, dhi, orc.COST_NCC, want_subpix=True)
    bs = sr.BlockSearch(L, R, block, dlo, dhi, cost=sr.capi.COST_NCC, ctx=ctx)
    got = bs.computeDisparityMapRight(1.0)
    assert "ncc-tile" in ctx.last_kernel
    np.testing.assert_array_equal(got, d.astype(np.float64))
    gots = bs.computeDisparityMapRight(1.0, subpixel=True)
    np.testing.assert_array_equal(gots, sp.astype(np.float64))
    assert np.abs(gots - sp).max() <= 1e-3       # north_star tolerance (met exactly)


@pytest.mark.parametrize("case", ["u8_integers", "subpixel_narrow", "wide_dynamic_range", "with_inf", "all_zero", "k_larger_than_image",
                                  "near_exact_limit", "even_width"])
def test_box_blur_order_free_and_sequential_paths(sr, ctx, orc, case):
    """cv::blur(32F, k x k) of removeDisparityOutliers: the prefix-sum path (taken when every partial sum is
    exactly representable in fp64) and the sequential OpenCV recurrences (any other input) both equal the
    oracle's literal RowSum/ColumnSum restatement bit for bit."""
    rng = np.random.default_rng(17)
    h, w, k = 77, 133, 21
    if case == "u8_integers":
        d = rng.integers(0, 256, (h, w)).astype(np.float32)
    elif case == "subpixel_narrow":
        d = (rng.integers(1 << 10, 1 << 19, (h, w)) / 1024.0).astype(np.float32)       # [1, 512), 10 fractional bits
    elif case == "wide_dynamic_range":
        d = (rng.standard_normal((h, w)) * 10.0 ** rng.integers(-12, 12, (h, w))).astype(np.float32)
    elif case == "with_inf":
        d = rng.integers(0, 256, (h, w)).astype(np.float32); d[5, 7] = np.inf; d[40, 100] = -np.inf
    elif case == "all_zero":
        d = np.zeros((h, w), np.float32)
    elif case == "near_exact_limit":
        # 46 bits between the largest value and the smallest set bit: window sums (<= 2 k^2 terms) still fit 53 bits,
        # a column prefix over (h + k - 1) * k terms does not -- the exactness proof has to count the prefix length
        k = 3
        d = (rng.integers(1, 1 << 24, (h, w)).astype(np.float64) * 2.0 ** (22 * rng.integers(0, 2, (h, w)))).astype(np.float32)
    elif case == "even_width":
        w = 134                                                                          # 16-byte loads of two columns
        d = rng.integers(0, 256, (h, w)).astype(np.float32)
    else:
        d = rng.integers(0, 64, (h, w)).astype(np.float32); k = 301                    # window wider and taller than the map
    with np.errstate(invalid="ignore", over="ignore"):
        want = orc.box_blur(d, k)
    got = sr.reconstruction.boxBlur(d, k, ctx=ctx)
    assert got.tobytes() == want.tobytes()



@pytest.mark.parametrize("thr", [0.37, 1.0, 2.0, 3.0000002, 1e-3, 7.5e5, 0.0, -1.0, float("inf"), float("nan"), 1e-30, 3e19])
def test_mesh_validity_threshold_is_exact_without_square_roots(sr, ctx, orc, thr):
    """CheckTriangularValidity (reconstruction.cpp:46-69) compares sqrtf(dx^2 + dy^2 + dz^2) > thr; the kernels compare the
    squared length with S(thr) = max{s : sqrtf(s) <= thr}.  Vertices whose edge lengths crowd around thr within a few
    ulps (both sides), plus exact hits, must give the oracle's face list for ordinary, tiny, huge, zero, negative,
    infinite and NaN thresholds."""
    R = sr.reconstruction
    rng = np.random.default_rng(99)
    w, h = 257, 129
    v = np.zeros(w * h, sr.capi.VERTEX_DTYPE)
    t = np.float32(thr) if np.isfinite(thr) and 0 < thr < 1e18 else np.float32(1.0)
    # a 0.8 thr x 0.6 thr grid: the diagonal of every cell is thr up to rounding (3-4-5), the sides are shorter -- so
    # validity hangs on sqrtf(s) > thr with s within a few ulps of thr^2, on both sides
    gx, gy = np.meshgrid(np.arange(w, dtype=np.float64), np.arange(h, dtype=np.float64))
    pos = np.stack([gx * 0.8 * float(t), gy * 0.6 * float(t), np.zeros_like(gx), np.ones_like(gx)], -1)
    p32 = pos.astype(np.float32)
    # nudge a third of the x coordinates by +-1..3 ulps
    nud = rng.integers(-3, 4, (h, w)).astype(np.int32) * (rng.random((h, w)) < 0.33)
    p32[..., 0] = (p32[..., 0].view(np.int32) + nud).view(np.float32)
    p32[rng.random((h, w)) < 0.02] = -np.inf             # a few invalid vertices
    v["pos"] = p32.reshape(-1, 4)
    got = R.meshFaces(v, w, h, thr, ctx=ctx)
    want = orc.mesh_faces(v, w, h, thr)
    np.testing.assert_array_equal(got, want)
    if np.isfinite(thr) and 1e-3 <= thr <= 1e6:
        assert 0.1 < len(want) / (2.0 * (w - 1) * (h - 1)) < 0.9      # the threshold really cuts through this set


def test_mesh_faces_large_map(sr, ctx, orc):
    """Face list of a map with more than 4096 scan blocks (several sweeps of the block-total scan, ragged tail)."""
    R = sr.reconstruction
    bgr, disp, K = _recon_inputs(orc, 1203, 997)
    h, w = disp.shape
    depth = R.convertDisparityToDepth(disp, K[0, 0], 1.0, ctx=ctx)
    v = R.backproject(bgr, depth, K, ctx=ctx)
    want_v = orc.backproject(bgr, depth, K[0, 0], K[1, 1], K[0, 2], K[1, 2])
    assert v.tobytes() == want_v.tobytes()
    for thr in (2.0, 1e9):
        np.testing.assert_array_equal(R.meshFaces(v, w, h, thr, ctx=ctx), orc.mesh_faces(want_v, w, h, thr))
    d = disp.copy(); R.removeDisparityOutliers(d, 75, 1.5, 0.8, ctx=ctx)
    np.testing.assert_array_equal(d, orc.remove_outliers(disp, 75, 1.5, 0.8))
    # the fused pipeline (vertices and validity flags from one pass) on the same map
    got = R.reconstructFused(bgr, disp, K, K[0, 0], 1.0, kernelSize=0, thrMesh=1e9, ctx=ctx)
    assert got["vertices"].tobytes() == want_v.tobytes()
    np.testing.assert_array_equal(got["depth"], depth)
    np.testing.assert_array_equal(got["faces"], orc.mesh_faces(want_v, w, h, 1e9))
