"""Parity at BASELINE.json's full sizes.  The oracle cannot sweep 4e9 evaluations in test time, so these
tests combine (1) the oracle on row samples spread over the image (top border, middle, bottom border),
computed on crops that carry the window halo, with (2) size-independent properties of the path: a row-band
split, the host row-chunk pipeline and the int16 wire format must all reproduce the one-launch result
exactly."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def sr():
    import stereo_reconstruction_b200 as s
    return s


@pytest.fixture(scope="module")
def ctx(sr):
    return sr.capi.default_context()


def _row_samples(h, n, pad):
    """(crop_row0, crop_row1, compare_row0, compare_row1): top border, middle, bottom border."""
    mid = h // 2 - n // 2
    return [(0, n + pad, 0, n), (mid - pad, mid + n + pad, mid, mid + n), (h - n - pad, h, h - n, h)]


def _check_right_rows(orc, L, R, block, D, cost, got, n=10):
    h = R.shape[0]
    for c0, c1, r0, r1 in _row_samples(h, n, block):
        want = orc.block_search_right_fast(L[c0:c1], R[c0:c1], block, 0, D, 1.0, cost=cost, rows=(r0 - c0, r1 - c0))
        np.testing.assert_array_equal(got[r0:r1], want[r0 - c0:r1 - c0], err_msg=f"rows {r0}..{r1}")


def test_config4_4k512_sad(sr, ctx, orc):
    """BASELINE.json configs[3]: 3840x2160, 512 disparities, 9x9 SAD (and the reference's SSD cost)."""
    capi = sr.capi
    w, h, D, block = 3840, 2160, 512, 9
    L, R = orc.synth_pair(0xB200, 0, w, h, 0, D)
    pair = capi.Pair(ctx, w, h, w, h)
    pair.upload(L, R)
    for cost in (capi.COST_SAD, capi.COST_SSD_L2):
        p = capi.search_params(block=block, min_disp=0, max_disp=D, smooth=1.0, cost=cost)
        pair.run(p)
        full = pair.download(p)
        assert ctx.last_kernel == "tile"
        # the WHOLE map against the oracle port (1-3 s on the box's host cores), not a row sample
        want = orc.block_search_right_fast(L, R, block, 0, D, 1.0, cost=cost)
        bad = int((full != want).sum())
        print(f"4K/512 cost {cost}: {bad} mismatching pixels of {full.size}")
        assert bad == 0
        assert full.min() >= 0 and full.max() < D
        if cost != capi.COST_SAD:
            continue
        # row bands (the multi-GPU sharding) reassemble to the same map
        banded = np.full((h, w), -1.0)
        for rank in range(3):
            r0, r1 = capi.band_rows(h, rank, 3)
            pair.run(p, r0, r1)
            pair.download(p, out=banded, row0=r0, row1=r1)
        np.testing.assert_array_equal(banded, full)
        # int16 wire format of the gather
        pair.run(p)
        pair.pack16(0, h)
        pair.run(capi.search_params(block=5, min_disp=0, max_disp=64, smooth=1.0, cost=cost), 0, 64)   # clobber rows
        pair.unpack16(0, h)
        np.testing.assert_array_equal(pair.download(p), full)
        # one-shot host call: row-chunk pipeline with pinned buffers
        Lp, a = capi.pinned_empty(L.shape, np.uint8); Rp, b = capi.pinned_empty(R.shape, np.uint8)
        outp, c = capi.pinned_empty((h, w), np.float64)
        try:
            Lp[...] = L; Rp[...] = R
            n0 = ctx.launches
            capi.block_search(ctx, Lp, Rp, p, out=outp)
            assert ctx.launches - n0 > 8            # several chunks
            np.testing.assert_array_equal(np.asarray(outp), full)
        finally:
            for x in (a, b, c):
                capi.free_pinned(x)
    pair.close()


def test_config3_1080p256_ncc_and_smoothing(sr, ctx, orc):
    """BASELINE.json configs[2] shape: NCC (sliding-sum kernel + border frame) against the oracle on row samples,
    and the reference's smoothFactor rule (raster dependency from row 0) on the top rows."""
    capi = sr.capi
    w, h, D, block = 1920, 1080, 256, 9
    L, R = orc.synth_pair(0xB200, 0, w, h, 0, D)
    got = sr.BlockSearch(L, R, block, 0, D, cost=capi.COST_NCC, ctx=ctx).computeDisparityMapRight(1.0)
    assert "ncc-tile" in ctx.last_kernel
    half = (block - 1) // 2
    n = 6
    for c0, c1, r0, r1 in _row_samples(h, n, block):
        d, _, _ = orc.search_ext(L[c0:c1], R[c0:c1], block, 0, D, orc.COST_NCC)
        # rows of the crop whose windows do not touch an artificial crop edge
        a = r0 if c0 == 0 else max(r0, c0 + half + 1)
        b = r1 if c1 == h else min(r1, c1 - half - 1)
        np.testing.assert_array_equal(got[a:b], d[a - c0:b - c0].astype(np.float64), err_msg=f"NCC rows {a}..{b}")
    # smoothing rule: the raster pass starts at row 0, so the first rows of the image are a complete sub-problem
    top = 24
    gs = sr.BlockSearch(L, R, 17, 0, D, ctx=ctx).computeDisparityMapRight(0.9)
    want = orc.block_search_right_fast(L[:top + 17], R[:top + 17], 17, 0, D, 0.9, rows=(0, top))
    np.testing.assert_array_equal(gs[:top], want[:top])


def test_config5_batch_720p_subpixel(sr, ctx, orc):
    """BASELINE.json configs[4] shape: 1280x720, 128 disparities, 7x7 SAD + sub-pixel through the batch API."""
    capi = sr.capi
    w, h, D, block, n = 1280, 720, 128, 7, 5
    pairs = [orc.synth_pair(0xB200, i, w, h, 0, D) for i in range(n)]
    outs = [np.zeros((h, w), np.float64) for _ in range(n)]
    p = capi.search_params(block=block, min_disp=0, max_disp=D, smooth=1.0, cost=capi.COST_SAD, subpixel=True)
    capi.block_search_batch(ctx, [a for a, _ in pairs], [b for _, b in pairs], p, outs)
    half = (block - 1) // 2
    for i in range(n):
        L, R = pairs[i]
        _, _, sp = orc.search_ext_fast(L, R, block, 0, D, orc.COST_SAD, want_subpix=True)   # whole maps
        assert np.abs(outs[i] - sp).max() <= 1e-3                                # north_star tolerance
        np.testing.assert_array_equal(outs[i], sp.astype(np.float64))            # and in fact bit-exact
    # the literal O(window) definition on row samples of one pair (ties the fast oracle to it at this size too)
    for i in (0,):
        L, R = pairs[i]
        for c0, c1, r0, r1 in _row_samples(h, 6, block):
            _, _, sp = orc.search_ext(L[c0:c1], R[c0:c1], block, 0, D, orc.COST_SAD, want_subpix=True)
            a = r0 if c0 == 0 else max(r0, c0 + half + 1)
            b = r1 if c1 == h else min(r1, c1 - half - 1)
            np.testing.assert_array_equal(outs[i][a:b], sp[a - c0:b - c0].astype(np.float64))
