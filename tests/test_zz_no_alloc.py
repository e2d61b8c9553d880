"""sr_box_blur_static: the allocation-free variant of sr_box_blur (scratch in a file-scope __device__ array, see
include/sr_capi.h) against the existing entry point.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_PIXELS = 1 << 20      # SR_BLUR_STATIC_MAX_PIXELS


@pytest.fixture(scope="module")
def sr():
    import stereo_reconstruction_b200 as s
    return s


@pytest.fixture(scope="module")
def ctx(sr):
    return sr.capi.default_context()


def _blur_static(sr, ctx, d, k):
    """d (host, float32 h x w) -> device (torch owns the buffers; the call itself allocates nothing) -> blur -> host."""
    import torch
    h, w = d.shape
    src = torch.from_numpy(np.ascontiguousarray(d)).to("cuda:0")
    dst = torch.full((h, w), -7.0, dtype=torch.float32, device="cuda:0")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    rc = sr.capi.lib().sr_box_blur_static(ctx.h, C.c_void_p(src.data_ptr()), C.c_void_p(dst.data_ptr()),
                                          C.c_int(w), C.c_int(h), C.c_int(k), C.c_void_p(None))
    ctx.check(rc)
    ctx.sync()
    free1 = torch.cuda.mem_get_info()[0]
    return dst.cpu().numpy(), free0, free1


def test_static_blur_equals_box_blur_on_two_calls_in_a_row(sr, ctx):
    rng = np.random.default_rng(23)
    # first call: a large map (fills most of the static array); second call: a smaller map with other content and a
    # window larger than the image -- what the first call left in the array must not show; third: floats outside the
    # exact-sum proof, where sr_box_blur runs the sequential recurrences too
    a = rng.integers(0, 256, (900, 1100)).astype(np.float32)
    b = (rng.integers(1 << 10, 1 << 19, (77, 133)) / 1024.0).astype(np.float32)
    c = (rng.standard_normal((201, 64)) * 10.0 ** rng.integers(-12, 12, (201, 64))).astype(np.float32)
    for i, (d, k) in enumerate([(a, 21), (b, 301), (c, 5), (a, 4)]):
        want = sr.reconstruction.boxBlur(d, k, ctx=ctx)              # existing entry point (sr_box_blur)
        got, free0, free1 = _blur_static(sr, ctx, d, k)
        assert got.dtype == want.dtype and got.shape == want.shape
        assert got.tobytes() == want.tobytes(), f"call {i}: max diff {np.nanmax(np.abs(got - want))}"   # tolerance 0: bit-identical
        if i > 0:                                                        # (the first call may load the module)
            assert free1 >= free0, "the call changed the free device memory"


def test_static_blur_rejects_maps_beyond_the_static_scratch(sr, ctx):
    import torch
    w, h = 1025, 1024                                                    # one column more than the maximum
    assert w * h > MAX_PIXELS
    src = torch.zeros((h, w), dtype=torch.float32, device="cuda:0")
    dst = torch.zeros_like(src)
    l = sr.capi.lib()
    rc = l.sr_box_blur_static(ctx.h, C.c_void_p(src.data_ptr()), C.c_void_p(dst.data_ptr()), w, h, 5, C.c_void_p(None))
    assert rc == sr.capi.E_SHAPE
    assert b"exceed the static scratch" in l.sr_last_error(ctx.h)
    # exactly the maximum is accepted
    rc = l.sr_box_blur_static(ctx.h, C.c_void_p(src.data_ptr()), C.c_void_p(dst.data_ptr()), 1024, 1024, 5, C.c_void_p(None))
    ctx.check(rc)
    ctx.sync()
    assert float(dst.abs().max()) == 0.0
    # host pointers are refused instead of faulting
    hbuf = np.zeros((8, 8), np.float32)
    rc = l.sr_box_blur_static(ctx.h, C.c_void_p(hbuf.ctypes.data), C.c_void_p(dst.data_ptr()), 8, 8, 3, C.c_void_p(None))
    assert rc == sr.capi.E_BADARG
