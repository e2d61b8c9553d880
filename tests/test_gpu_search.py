"""Parity of the CUDA WindowSearch path (through the C-ABI) with the CPU oracle and the committed
cv2-executed golden vectors.  Bar: bit-exact (integer disparities stored as CV_64F)."""
import os

import numpy as np
import pytest

from conftest import golden_cases

pytestmark = pytest.mark.gpu
os.environ.setdefault("SR_PEER_TIMEOUT_MS", "400")   # band-gather waits give up quickly in the protocol tests (default 10 s)


@pytest.fixture(scope="module")
def sr():
    import stereo_reconstruction_b200 as s
    return s


@pytest.fixture(scope="module")
def ctx(sr):
    return sr.capi.default_context()


def synth(orc, w, h, min_d, max_d, seed=0xB200, pair=0, quant=False):
    return orc.synth_pair(seed, pair, w, h, min_d, max_d, quant)


@pytest.mark.parametrize("name,z", golden_cases("right"), ids=lambda v: v if isinstance(v, str) else "")
def test_right_golden(sr, ctx, name, z):
    bs = sr.BlockSearch(z["left"], z["right"], int(z["block"]), int(z["min_d"]), int(z["max_d"]), ctx=ctx)
    out = bs.computeDisparityMapRight(float(z["smooth"]), bool(z["var_block"]), float(z["thres"]))
    np.testing.assert_array_equal(out, z["out"])
    assert bs.maxBlockSize == int(z["max_block"])


@pytest.mark.parametrize("block", [3, 5, 7, 8, 9, 11, 13, 15, 17, 19, 21])
@pytest.mark.parametrize("smooth", [1.0, 0.9])
def test_right_vs_oracle_synth(sr, ctx, orc, block, smooth):
    w, h, D = 700, 150, 70
    L, R = synth(orc, w, h, 0, D)
    want = orc.block_search_right_fast(L, R, block, 0, D, smooth)
    got = sr.BlockSearch(L, R, block, 0, D, ctx=ctx).computeDisparityMapRight(smooth)
    np.testing.assert_array_equal(got, want)


@pytest.mark.parametrize("cost", [0, 1])
@pytest.mark.parametrize("quant", [False, True])
def test_right_ties_and_sad(sr, ctx, orc, cost, quant):
    """4-level quantised images make ties frequent: first minimum (smallest d) must win."""
    w, h, D = 530, 97, 40
    L, R = synth(orc, w, h, 0, D, seed=7, quant=quant)
    R[10:20, 30:50] = 0  # black pixels are skipped (BlockSearch.cpp:105)
    want = orc.block_search_right_fast(L, R, 9, 0, D, 1.0, cost=cost)
    got = sr.BlockSearch(L, R, 9, 0, D, cost=cost, ctx=ctx).computeDisparityMapRight(1.0)
    np.testing.assert_array_equal(got, want)


@pytest.mark.parametrize("w1,h1,w2,h2", [(300, 80, 280, 80), (280, 80, 300, 80), (300, 90, 300, 80), (64, 33, 64, 33), (17, 9, 17, 9)])
def test_right_unequal_sizes(sr, ctx, orc, w1, h1, w2, h2):
    rng = np.random.default_rng(3)
    L = rng.integers(0, 256, (h1, w1, 3), dtype=np.uint8)
    R = rng.integers(0, 256, (h2, w2, 3), dtype=np.uint8)
    for smooth in (1.0, 0.8):
        want = orc.block_search_right_fast(L, R, 7, 0, 50, smooth)
        got = sr.BlockSearch(L, R, 7, 0, 50, ctx=ctx).computeDisparityMapRight(smooth)
        np.testing.assert_array_equal(got, want)


def test_right_min_disp_and_wide_range(sr, ctx, orc):
    L, R = synth(orc, 400, 60, 5, 300)
    for mn, mx in ((5, 300), (0, 1), (3, 4), (390, 500), (0, 1000)):
        want = orc.block_search_right_fast(L, R, 9, mn, mx, 0.9)
        got = sr.BlockSearch(L, R, 9, mn, mx, ctx=ctx).computeDisparityMapRight(0.9)
        np.testing.assert_array_equal(got, want, err_msg=f"range {mn},{mx}")


def test_right_row_band(sr, ctx, orc):
    """Row bands (config 4 sharding): a band equals the same rows of the full map."""
    L, R = synth(orc, 600, 120, 0, 64)
    want = orc.block_search_right_fast(L, R, 9, 0, 64, 1.0)
    pair = sr.capi.Pair(ctx, 600, 120, 600, 120)
    pair.upload(L, R)
    p = sr.capi.search_params(block=9, min_disp=0, max_disp=64, smooth=1.0)
    for r0, r1 in ((0, 40), (40, 95), (95, 120), (7, 8)):
        pair.run(p, r0, r1)
        got = pair.download(p, row0=r0, row1=r1)
        np.testing.assert_array_equal(got[r0:r1], want[r0:r1])


def test_errors(sr, ctx):
    L = np.zeros((40, 60, 3), np.uint8)
    R = np.zeros((50, 60, 3), np.uint8)
    with pytest.raises(sr.capi.SrError) as e:  # h1 < h2: the reference throws cv::Exception
        sr.BlockSearch(L, R, 9, 0, 16, ctx=ctx).computeDisparityMapRight(1.0)
    assert e.value.code == sr.capi.E_SHAPE
    with pytest.raises(sr.capi.SrError):
        sr.BlockSearch(L, L, 0, 0, 16, ctx=ctx).computeDisparityMapRight(1.0)
    with pytest.raises(sr.capi.SrError):
        sr.BlockSearch(L[:, :, :2], L, 9, 0, 16, ctx=ctx).computeDisparityMapRight(1.0)


@pytest.mark.parametrize("cost,block,min_d", [(0, 9, 0), (1, 9, 0), (0, 17, 2), (1, 5, 0)])
@pytest.mark.parametrize("chunks", ["3", "7"])
def test_one_shot_row_chunk_pipeline(sr, ctx, orc, monkeypatch, cost, block, min_d, chunks):
    """sr_block_search with pinned host buffers runs the row-chunk pipeline (H2D / search / D2H
    overlapped): the result must equal the oracle and the plain (pageable, single-launch) call."""
    capi = sr.capi
    w, h, D = 640, 333, 48
    L, R = synth(orc, w, h, min_d, D, seed=11)
    R[100:120, 200:260] = 0
    # a pitch wider than the row, as cv::Mat ROI views have
    Lp, lh = capi.pinned_empty((h, w + 5, 3), np.uint8)
    Rp, rh = capi.pinned_empty((h, w + 5, 3), np.uint8)
    outp, oh_ = capi.pinned_empty((h, w + 3), np.float64)
    try:
        Lp[:, :w] = L
        Rp[:, :w] = R
        outp[...] = -7.0
        monkeypatch.setenv("SR_PIPE_CHUNKS", chunks)
        p = capi.search_params(block=block, min_disp=min_d, max_disp=D, smooth=1.0, cost=cost)
        n0 = ctx.launches
        got = capi.block_search(ctx, Lp[:, :w], Rp[:, :w], p, out=outp[:, :w])
        assert ctx.launches - n0 >= 3 * 2, "row-chunk pipeline did not run (expected several chunk launches)"
        want = orc.block_search_right_fast(L, R, block, min_d, D, 1.0, cost=cost)
        np.testing.assert_array_equal(np.asarray(got), want)
        assert (outp[:, w:] == -7.0).all(), "pipeline wrote outside the output rows' extent"
        plain = capi.block_search(ctx, L, R, p)
        np.testing.assert_array_equal(plain, want)
    finally:
        capi.free_pinned(lh); capi.free_pinned(rh); capi.free_pinned(oh_)


@pytest.mark.parametrize("pin", ["none", "left", "out"])
def test_row_chunk_pipeline_pageable_sides(sr, ctx, orc, monkeypatch, pin):
    """Pageable host buffers (cv::Mat memory) take the same pipeline through the staging threads
    (host_stage.cu): any mix of pinned and pageable sides, row pitches wider than the rows, chunks with
    fewer rows than there are staging threads."""
    capi = sr.capi
    w, h, D, block = 520, 260, 32, 7
    L, R = synth(orc, w, h, 0, D, seed=23)
    want = orc.block_search_right_fast(L, R, block, 0, D, 1.0, cost=1)
    monkeypatch.setenv("SR_PIPE_CHUNKS", "6")
    p = capi.search_params(block=block, min_disp=0, max_disp=D, smooth=1.0, cost=capi.COST_SAD)
    Lw = np.zeros((h, w + 7, 3), np.uint8); Rw = np.zeros((h, w + 2, 3), np.uint8)
    Lw[:, :w] = L; Rw[:, :w] = R
    outw = np.full((h, w + 3), -7.0)
    hs = []
    try:
        if pin == "left":
            Lw, hh = capi.pinned_empty((h, w + 7, 3), np.uint8); hs.append(hh)
            Lw[:, :w] = L
        if pin == "out":
            outw, hh = capi.pinned_empty((h, w + 3), np.float64); hs.append(hh)
            outw[...] = -7.0
        for _ in range(2):       # the second call reuses the staging buffers and the thread pool
            n0 = ctx.launches
            capi.block_search(ctx, Lw[:, :w], Rw[:, :w], p, out=outw[:, :w])
            assert ctx.launches - n0 >= 6 * 2
            np.testing.assert_array_equal(np.asarray(outw[:, :w]), want)
            assert (np.asarray(outw[:, w:]) == -7.0).all()
    finally:
        for hh in hs:
            capi.free_pinned(hh)


@pytest.mark.parametrize("pinned", [True, False])
def test_pair_search_band_from_host(sr, ctx, orc, monkeypatch, pinned):
    """sr_pair_search_band (one rank's share of a row-band sharded pair): partial upload with halo rows,
    band search, optional download -- the bands of a 3-way split reassemble to the oracle's full map."""
    capi = sr.capi
    w, h, D, block = 600, 301, 40, 9
    L, R = synth(orc, w, h, 0, D, seed=13)
    want = orc.block_search_right_fast(L, R, block, 0, D, 1.0, cost=1)
    p = capi.search_params(block=block, min_disp=0, max_disp=D, smooth=1.0, cost=capi.COST_SAD)
    monkeypatch.setenv("SR_PIPE_CHUNKS", "2")
    hs = []
    try:
        if pinned:
            Lh, a = capi.pinned_empty(L.shape, np.uint8); Rh, b = capi.pinned_empty(R.shape, np.uint8)
            out, c = capi.pinned_empty((h, w), np.float64); hs = [a, b, c]
            Lh[...] = L; Rh[...] = R
        else:
            Lh, Rh, out = L, R, np.zeros((h, w), np.float64)
        out[...] = -3.0
        for rank in range(3):
            r0, r1 = capi.band_rows(h, rank, 3)
            pair = capi.Pair(ctx, w, h, w, h)        # a fresh pair per "rank": nothing uploaded before
            if rank == 1:                            # device-resident band + separate download
                pair.search_band(Lh, Rh, p, r0, r1, out=None)
                pair.download(p, out=out, row0=r0, row1=r1)
            else:
                pair.search_band(Lh, Rh, p, r0, r1, out=out)
            pair.close()
        np.testing.assert_array_equal(np.asarray(out), want)
    finally:
        for x in hs:
            capi.free_pinned(x)


def test_int16_wire_format(sr, ctx, orc):
    """sr_search_pack16 / sr_search_unpack16: integer maps (incl. the reference's negative "-x" values) survive
    the int16 round trip bit-exactly; a sub-pixel map is refused."""
    capi = sr.capi
    w1, w2, h, D = 90, 120, 40, 16                   # w1 < w2: columns without a candidate store -x
    L, R = synth(orc, w2, h, 0, D, seed=5)
    L = np.ascontiguousarray(L[:, :w1])
    pair = capi.Pair(ctx, w1, h, w2, h)
    pair.upload(L, R)
    p = capi.search_params(block=9, min_disp=0, max_disp=D, smooth=1.0)
    pair.run(p)
    want = pair.download(p)
    assert (want < 0).any()
    pair.pack16(0, h)
    pair.run(capi.search_params(block=5, min_disp=0, max_disp=D, smooth=1.0))   # overwrite the CV_64F result
    pair.unpack16(3, h - 2)
    got = pair.download(p)
    np.testing.assert_array_equal(got[3:h - 2], want[3:h - 2])
    ps = capi.search_params(block=9, min_disp=0, max_disp=D, smooth=1.0, cost=capi.COST_SAD, subpixel=True)
    pair.run(ps)
    pair.pack16(0, h)
    with pytest.raises(capi.SrError):
        pair.download(ps)
    pair.close()


def test_host_register_enables_the_pipeline(sr, ctx, orc, monkeypatch):
    """sr_host_register pins caller-owned buffers in place (a cv::Mat's data): sr_block_search then takes the
    row-chunk pipeline exactly as with sr_host_alloc memory."""
    capi = sr.capi
    w, h, D = 640, 300, 40
    L, R = synth(orc, w, h, 0, D, seed=19)
    out = np.zeros((h, w), np.float64)
    monkeypatch.setenv("SR_PIPE_CHUNKS", "4")
    p = capi.search_params(block=9, min_disp=0, max_disp=D, smooth=1.0, cost=capi.COST_SAD)
    for a in (L, R, out):
        capi.host_register(a)
    try:
        n0 = ctx.launches
        capi.block_search(ctx, L, R, p, out=out)
        assert ctx.launches - n0 >= 4 * 2
        np.testing.assert_array_equal(out, orc.block_search_right_fast(L, R, 9, 0, D, 1.0, cost=1))
    finally:
        for a in (L, R, out):
            capi.host_unregister(a)


@pytest.mark.parametrize("w", [404, 403])       # even width: fused wire-format resolve; odd: search + push kernel
def test_peer_memory_band_gather_protocol(sr, orc, w):
    """Band gather over peer memory (capi_peer.cu) with the peers standing in as pair objects of one process:
    three "ranks" search their bands, ranks 1 and 2 store int16 rows into the root's window, the root collects.
    Several steps on changing images exercise the step flags and the acknowledgement that guards the window; then a
    failing step (sticky status on every rank, no acknowledgement) and a second generation on the same pairs."""
    capi = sr.capi
    h, D, block, world = 157, 24, 7, 3
    ctxs = [capi.Context(0) for _ in range(world)]
    pairs = [capi.Pair(c, w, h, w, h) for c in ctxs]
    p = capi.search_params(block=block, min_disp=0, max_disp=D, smooth=1.0, cost=capi.COST_SAD)

    def step(seq, fused):
        L, R = synth(orc, w, h, 0, D, seed=40 + seq)
        if seq == 2:
            R[50:60, 100:140] = 0                      # black pixels: 0.0 in the map
        want = orc.block_search_right_fast(L, R, block, 0, D, 1.0, cost=1)   # holds negative entries (-x) too
        for r in range(world):
            pairs[r].upload(L, R)
        # pushes first: on one GPU the collect kernel's waiting blocks would otherwise keep the SMs busy
        for r in (2, 1):
            r0, r1 = capi.band_rows(h, r, world)
            if fused:
                n0 = ctxs[r].launches
                pairs[r].run_push(p, r0, r1, r, seq)   # the resolve pass writes the wire format and publishes the step
                fused_launches = ctxs[r].launches - n0
            else:
                pairs[r].run(p, r0, r1)
                pairs[r].push_band(r0, r1, r, seq)
        pairs[0].run(p, *capi.band_rows(h, 0, world))
        pairs[0].collect_bands(world, seq)
        np.testing.assert_array_equal(pairs[0].download(p), want, err_msg=f"step {seq}")
        return fused_launches if fused else None

    try:
        handle = pairs[0].peer_export()
        assert len(handle) == 64
        for r in (1, 2):
            pairs[r].peer_attach_local(pairs[0])
        step(1, False)
        n_fused = step(2, True)
        step(3, True)
        # memset is not a kernel: search + (wait-ack + wire resolve) when fused, search + resolve + push otherwise
        assert n_fused == 3
        # step 4: rank 1 holds values outside the int16 wire format (a sub-pixel map) -> refused on rank 1, never
        # published; rank 2 stays silent -> the root's collect kernel gives up instead of hanging and does NOT
        # acknowledge the step; every later call on the failed ranks reports the (sticky) failure
        ps = capi.search_params(block=block, min_disp=0, max_disp=D, smooth=1.0, cost=capi.COST_SAD, subpixel=True)
        pairs[1].run(ps)
        pairs[1].push_band(0, h, 1, 4)
        with pytest.raises(capi.SrError) as e:
            pairs[1].peer_status()
        assert e.value.code == capi.E_BADARG
        with pytest.raises(capi.SrError):
            pairs[1].push_band(0, h, 1, 5)             # sticky
        with pytest.raises(capi.SrError):
            ctxs[1].sync()                             # and visible at the context's synchronisation
        pairs[0].run(p, *capi.band_rows(h, 0, world))
        pairs[0].collect_bands(world, 4)
        with pytest.raises(capi.SrError, match="did not arrive"):
            pairs[0].download(p)
        with pytest.raises(capi.SrError, match="did not arrive"):
            pairs[0].collect_bands(world, 5)           # sticky on the root as well
        # rank 2 was never told that step 4 failed; its next push waits for an acknowledgement that never comes
        r0, r1 = capi.band_rows(h, 2, world)
        pairs[2].run(p, r0, r1)
        pairs[2].push_band(r0, r1, 2, 5)
        with pytest.raises(capi.SrError, match="did not arrive"):
            pairs[2].peer_status()
        # a second generation on the same pairs: export again (header and status reset), attach again, steps restart at 1
        pairs[0].peer_export()
        for r in (1, 2):
            pairs[r].peer_detach()
            pairs[r].peer_attach_local(pairs[0])
        step(1, True)
        step(2, False)
    finally:
        for r in (1, 2):
            pairs[r].peer_detach()
        for q in pairs:
            q.close()
        for c in ctxs:
            c.close()


@pytest.mark.parametrize("chunks", ["2", "5"])
@pytest.mark.parametrize("smooth", [0.9, 1.3])
def test_one_shot_call_chunks_the_smoothing_rule(sr, ctx, orc, monkeypatch, chunks, smooth):
    """sr_block_search with smoothFactor != 1 runs as a row-chunk pipeline too: search, bit planes and copies of
    different chunks overlap, the raster passes resolve in row order (each starts from the last row of the chunk above).
    Same map as the oracle's, from pinned and from pageable buffers, for a bonus (adder-style chain) and a penalty
    (generic chain) factor, incl. a width of more than 32 words per row."""
    monkeypatch.setenv("SR_PIPE_CHUNKS", chunks)
    capi = sr.capi
    w, h, D = 1100, 230, 40
    L, R = orc.synth_pair(0xB200, 21, w, h, 0, D, True)
    R[100:108, 300:420] = 0
    want = orc.block_search_right_fast(L, R, 9, 0, D, smooth)
    p = capi.search_params(block=9, min_disp=0, max_disp=D, smooth=smooth)
    got = capi.block_search(ctx, L, R, p)                      # pageable buffers (staging threads)
    np.testing.assert_array_equal(got, want)
    Lp, a1 = capi.pinned_empty(L.shape, np.uint8); Lp[...] = L
    Rp, a2 = capi.pinned_empty(R.shape, np.uint8); Rp[...] = R
    o, a3 = capi.pinned_empty((h, w), np.float64); o[...] = -3.0
    try:
        capi.block_search(ctx, Lp, Rp, p, out=o)
        np.testing.assert_array_equal(np.asarray(o), want)
        capi.block_search(ctx, Lp, Rp, p, out=o)               # again: buffers of the first call are reused
        np.testing.assert_array_equal(np.asarray(o), want)
    finally:
        for a in (a1, a2, a3):
            capi.free_pinned(a)
