"""ctypes binding of libsr_b200.so (the C-ABI declared in include/sr_capi.h).

There is no CPU fallback: loading fails loudly when the shared library is missing, and every
call fails with SR_E_CUDA when no B200 is present.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SR_B200_LIB") or os.path.join(_HERE, "libsr_b200.so")

OK, E_BADARG, E_SHAPE, E_CUDA, E_NCCL, E_IO = 0, -1, -2, -3, -4, -5
VARIANT_RIGHT, VARIANT_LEFT, VARIANT_LINEAR = 0, 1, 2
COST_SSD_L2, COST_SAD, COST_NCC = 0, 1, 2

VERTEX_DTYPE = np.dtype({"names": ["pos", "rgba"], "formats": [(np.float32, 4), (np.uint8, 4)],
                         "offsets": [0, 16], "itemsize": 32})


class SrImage(C.Structure):
    _fields_ = [("bgr", C.c_void_p), ("w", C.c_int), ("h", C.c_int), ("pitch", C.c_size_t)]


class SrSearchParams(C.Structure):
    _fields_ = [("variant", C.c_int), ("block", C.c_int), ("min_disp", C.c_int), ("max_disp", C.c_int),
                ("smooth", C.c_double), ("var_block", C.c_int), ("thres", C.c_double), ("cost", C.c_int),
                ("subpixel", C.c_int), ("lr_check", C.c_int)]


class SrEval(C.Structure):
    _fields_ = [("n", C.c_int), ("bad", C.c_int), ("invalid", C.c_int), ("serr", C.c_double),
                ("bad_pct", C.c_float), ("invalid_pct", C.c_float), ("total_bad_pct", C.c_float),
                ("avg_err", C.c_float)]


class SrError(RuntimeError):
    """Non-zero status from the C-ABI (the C++ shim throws std::runtime_error / cv::Exception)."""

    def __init__(self, code, msg):
        super().__init__(f"libsr_b200 error {code}: {msg}")
        self.code = code


# every symbol include/sr_capi.h declares (tests check that the library exports all of them)
SYMBOLS = [
    "sr_ctx_create", "sr_ctx_destroy", "sr_last_error", "sr_version", "sr_host_alloc", "sr_host_free",
    "sr_ctx_stream", "sr_ctx_sync", "sr_timer_start", "sr_timer_stop", "sr_kernel_launches", "sr_last_kernel",
    "sr_block_search", "sr_block_search_batch", "sr_pair_create", "sr_pair_destroy", "sr_pair_upload",
    "sr_search_run", "sr_search_download", "sr_search_result_dev", "sr_search_max_block", "sr_search_left_disparity", "sr_pair_search_band", "sr_host_register", "sr_host_unregister", "sr_search_pack16", "sr_search_unpack16", "sr_search_result16_dev",
    "sr_pair_peer_export", "sr_pair_peer_attach", "sr_pair_peer_attach_local", "sr_pair_peer_detach", "sr_search_push_band", "sr_search_collect_bands",
    "sr_search_effective_evals", "sr_remove_outliers", "sr_box_blur", "sr_scale", "sr_disparity_to_depth",
    "sr_backproject", "sr_mesh_faces", "sr_reconstruct", "sr_pair_reconstruct", "sr_write_mesh",
    "sr_warp_nearest_f64", "sr_evaldisp", "sr_write_pfm", "sr_write_ply", "sr_band_rows", "sr_l2_flush",
    "sr_search_kernel_ms", "sr_int_peak", "sr_pair_reconstruct_warp", "sr_average_disparity", "sr_pfm_info", "sr_read_pfm",
    "sr_read_calib", "sr_search_run_push", "sr_pair_peer_status", "sr_search_run_batch", "sr_box_blur_static", "sr_warp_linear_bgr", "sr_pair_upload_warped",
]

_lib = None


def lib():
    """Loads libsr_b200.so (built by __graft_entry__.build() / csrc/Makefile)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(there is no CPU fallback)")
        l = C.CDLL(LIB_PATH)
        l.sr_last_error.restype = C.c_char_p
        l.sr_last_error.argtypes = [C.c_void_p]
        l.sr_version.restype = C.c_char_p
        l.sr_last_kernel.restype = C.c_char_p
        l.sr_last_kernel.argtypes = [C.c_void_p]
        l.sr_kernel_launches.restype = C.c_uint64
        l.sr_kernel_launches.argtypes = [C.c_void_p]
        l.sr_ctx_stream.restype = C.c_void_p
        l.sr_ctx_stream.argtypes = [C.c_void_p]
        l.sr_search_result_dev.restype = C.c_void_p
        l.sr_search_result_dev.argtypes = [C.c_void_p, C.c_void_p]
        l.sr_search_result16_dev.restype = C.c_void_p
        l.sr_search_result16_dev.argtypes = [C.c_void_p, C.c_void_p]
        l.sr_ctx_destroy.restype = None
        l.sr_ctx_destroy.argtypes = [C.c_void_p]
        l.sr_pair_destroy.restype = None
        l.sr_pair_destroy.argtypes = [C.c_void_p, C.c_void_p]
        _lib = l
    return _lib


def image(a):
    """numpy HxWx3 uint8 (BGR, any row stride) -> (keep-alive array, SrImage)."""
    a = np.asarray(a)
    if a.dtype != np.uint8 or a.ndim != 3 or a.shape[2] != 3:
        raise SrError(E_BADARG, "image must be HxWx3 uint8 (CV_8UC3)")
    if a.strides[2] != 1 or a.strides[1] != 3:
        a = np.ascontiguousarray(a)
    return a, SrImage(a.ctypes.data, a.shape[1], a.shape[0], a.strides[0])


class Context:
    """Owns one sr_ctx (one CUDA stream + buffers on one GPU)."""

    def __init__(self, device=0):
        self._l = lib()
        h = C.c_void_p()
        rc = self._l.sr_ctx_create(int(device), C.byref(h))
        if rc != OK:
            raise SrError(rc, self._l.sr_last_error(None).decode())
        self.h = h
        self.device = device

    def close(self):
        if getattr(self, "h", None):
            self._l.sr_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def check(self, rc):
        if rc != OK:
            raise SrError(rc, self._l.sr_last_error(self.h).decode())

    @property
    def launches(self):
        return int(self._l.sr_kernel_launches(self.h))

    @property
    def last_kernel(self):
        return self._l.sr_last_kernel(self.h).decode()

    def sync(self):
        self.check(self._l.sr_ctx_sync(self.h))

    def timer_start(self):
        self.check(self._l.sr_timer_start(self.h))

    def timer_stop(self):
        ms = C.c_float()
        self.check(self._l.sr_timer_stop(self.h, C.byref(ms)))
        return ms.value

    def l2_flush(self):
        self.check(self._l.sr_l2_flush(self.h))

    def search_kernel_ms(self):
        ms = C.c_float()
        self.check(self._l.sr_search_kernel_ms(self.h, C.byref(ms)))
        return ms.value

    def int_peak(self):
        """(dual-pipe, single-pipe) measured integer issue peaks in lane-ops/s."""
        a, b = C.c_double(), C.c_double()
        self.check(self._l.sr_int_peak(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    @property
    def stream(self):
        return int(self._l.sr_ctx_stream(self.h))


_default_ctx = {}


def default_context(device=0):
    if device not in _default_ctx:
        _default_ctx[device] = Context(device)
    return _default_ctx[device]


def pinned_empty(shape, dtype):
    """numpy array backed by pinned host memory (sr_host_alloc); freed with free_pinned."""
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    p = C.c_void_p()
    rc = lib().sr_host_alloc(C.byref(p), C.c_size_t(max(n, 1)))
    if rc != OK:
        raise SrError(rc, "sr_host_alloc failed")
    buf = (C.c_uint8 * max(n, 1)).from_address(p.value)
    arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
    arr.flags.writeable = True
    return arr, p


def host_register(a):
    """Pins a numpy array's memory in place (sr_host_register); pair with host_unregister."""
    rc = lib().sr_host_register(C.c_void_p(a.ctypes.data), C.c_size_t(a.nbytes))
    if rc != OK:
        raise SrError(rc, "sr_host_register failed")


def host_unregister(a):
    lib().sr_host_unregister(C.c_void_p(a.ctypes.data))


def free_pinned(p):
    lib().sr_host_free(p)


class Pair:
    """Device-resident stereo pair (sr_pair)."""

    def __init__(self, ctx, w1, h1, w2, h2):
        self.ctx = ctx
        h = C.c_void_p()
        ctx.check(ctx._l.sr_pair_create(ctx.h, w1, h1, w2, h2, C.byref(h)))
        self.h = h
        self.w1, self.h1, self.w2, self.h2 = w1, h1, w2, h2

    def close(self):
        if getattr(self, "h", None) and self.ctx.h:
            self.ctx._l.sr_pair_destroy(self.ctx.h, self.h)
        self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload(self, left, right):
        la, li = image(left)
        ra, ri = image(right)
        self.ctx.check(self.ctx._l.sr_pair_upload(self.ctx.h, self.h, C.byref(li), C.byref(ri)))
        self.ctx.sync()  # la/ra may be temporaries

    def upload_warped(self, left, H, right, Hp):
        """Un-rectified images up, rectifying warp (rectification.cpp:486-493) on the device into this pair, whose
        sizes are the rectified sizes (sr_pair_upload_warped)."""
        la, li = image(left)
        ra, ri = image(right)
        h = np.ascontiguousarray(H, np.float64).reshape(9); hp = np.ascontiguousarray(Hp, np.float64).reshape(9)
        self.ctx.check(self.ctx._l.sr_pair_upload_warped(self.ctx.h, self.h, C.byref(li), C.c_void_p(h.ctypes.data),
                                                         C.byref(ri), C.c_void_p(hp.ctypes.data)))
        self.ctx.sync()

    def run(self, params, row0=0, row1=None):
        if row1 is None:
            row1 = self.h1 if params.variant == VARIANT_LEFT else self.h2
        self.ctx.check(self.ctx._l.sr_search_run(self.ctx.h, self.h, C.byref(params), int(row0), int(row1)))

    def result_ptr(self):
        """Device pointer of the CV_64F result (tight rows of the output width)."""
        return int(self.ctx._l.sr_search_result_dev(self.ctx.h, self.h))

    def download(self, params, out=None, row0=0, row1=None):
        oh, ow = (self.h1, self.w1) if params.variant == VARIANT_LEFT else (self.h2, self.w2)
        if row1 is None:
            row1 = oh
        if out is None:
            out = np.zeros((oh, ow), np.float64)
        self.ctx.check(self.ctx._l.sr_search_download(self.ctx.h, self.h, C.c_void_p(out.ctypes.data),
                                                      C.c_size_t(out.strides[0]), int(row0), int(row1)))
        return out

    def pack16(self, row0, row1):
        """Narrows rows of the CV_64F result to the int16 wire format (sr_search_pack16)."""
        self.ctx.check(self.ctx._l.sr_search_pack16(self.ctx.h, self.h, int(row0), int(row1)))

    def unpack16(self, row0, row1):
        self.ctx.check(self.ctx._l.sr_search_unpack16(self.ctx.h, self.h, int(row0), int(row1)))

    def result16_ptr(self):
        return int(self.ctx._l.sr_search_result16_dev(self.ctx.h, self.h))

    # ---- band gather over peer memory (include/sr_capi.h: sr_pair_peer_*) ----
    def peer_export(self):
        """Root: creates the gather window, returns its CUDA IPC handle (64 bytes)."""
        buf = C.create_string_buffer(64)
        self.ctx.check(self.ctx._l.sr_pair_peer_export(self.ctx.h, self.h, buf))
        return buf.raw

    def peer_attach(self, handle):
        """Other ranks (other processes): maps the root's window."""
        self.ctx.check(self.ctx._l.sr_pair_peer_attach(self.ctx.h, self.h, C.c_char_p(bytes(handle))))

    def peer_attach_local(self, root):
        """Same-process stand-in for peer_attach (single-GPU tests of the protocol)."""
        self.ctx.check(self.ctx._l.sr_pair_peer_attach_local(self.ctx.h, self.h, root.h))

    def peer_detach(self):
        self.ctx.check(self.ctx._l.sr_pair_peer_detach(self.ctx.h, self.h))

    def push_band(self, row0, row1, rank, seq):
        self.ctx.check(self.ctx._l.sr_search_push_band(self.ctx.h, self.h, int(row0), int(row1), int(rank), C.c_uint(int(seq))))

    def run_push(self, params, row0, row1, rank, seq):
        """sr_search_run_push: band search whose resolve pass stores the int16 wire format into the root's window."""
        self.ctx.check(self.ctx._l.sr_search_run_push(self.ctx.h, self.h, C.byref(params), int(row0), int(row1), int(rank), C.c_uint(int(seq))))

    def peer_status(self):
        """Synchronises and raises if a step of the gather protocol failed on this rank (sticky)."""
        self.ctx.check(self.ctx._l.sr_pair_peer_status(self.ctx.h, self.h))

    def collect_bands(self, nranks, seq):
        self.ctx.check(self.ctx._l.sr_search_collect_bands(self.ctx.h, self.h, int(nranks), C.c_uint(int(seq))))

    def search_band(self, left, right, params, row0, row1, out=None):
        """sr_pair_search_band: rows [row0,row1) from host images; out=None keeps the band in HBM."""
        la, li = image(left)
        ra, ri = image(right)
        optr, pitch = (C.c_void_p(out.ctypes.data), out.strides[0]) if out is not None else (C.c_void_p(None), 0)
        self.ctx.check(self.ctx._l.sr_pair_search_band(self.ctx.h, self.h, C.byref(li), C.byref(ri), C.byref(params),
                                                       int(row0), int(row1), optr, C.c_size_t(pitch)))
        return out

    def left_disparity(self):
        """int32 left-referenced map of the last lr_check run (sr_search_left_disparity)."""
        out = np.zeros((self.h1, self.w1), np.int32)
        self.ctx.check(self.ctx._l.sr_search_left_disparity(self.ctx.h, self.h, C.c_void_p(out.ctypes.data), C.c_size_t(out.strides[0])))
        return out

    def max_block(self):
        v = C.c_int()
        self.ctx.check(self.ctx._l.sr_search_max_block(self.ctx.h, self.h, C.byref(v)))
        return v.value

    def effective_evals(self):
        v = C.c_uint64()
        self.ctx.check(self.ctx._l.sr_search_effective_evals(self.ctx.h, self.h, C.byref(v)))
        return int(v.value)


def band_rows(h, rank, nranks):
    r0, r1 = C.c_int(), C.c_int()
    rc = lib().sr_band_rows(int(h), int(rank), int(nranks), C.byref(r0), C.byref(r1))
    if rc != OK:
        raise SrError(rc, "sr_band_rows: bad arguments")
    return r0.value, r1.value


def search_params(variant=VARIANT_RIGHT, block=17, min_disp=0, max_disp=200, smooth=1.0, var_block=False,
                  thres=19.0, cost=COST_SSD_L2, subpixel=False, lr_check=False):
    return SrSearchParams(int(variant), int(block), int(min_disp), int(max_disp), float(smooth), int(bool(var_block)),
                          float(thres), int(cost), int(bool(subpixel)), int(bool(lr_check)))


def search_run_batch(ctx, pairs, params):
    """sr_search_run_batch: kernels of n uploaded pairs, one tile-kernel launch per group of up to 16 pairs of one shape."""
    arr = (C.c_void_p * len(pairs))(*[q.h for q in pairs])
    ctx.check(ctx._l.sr_search_run_batch(ctx.h, arr, len(pairs), C.byref(params)))


def block_search(ctx, left, right, params, out=None, return_max_block=False):
    """One-shot host call (sr_block_search): H2D, kernels, D2H."""
    la, li = image(left)
    ra, ri = image(right)
    oh, ow = (li.h, li.w) if params.variant == VARIANT_LEFT else (ri.h, ri.w)
    if out is None:
        out = np.zeros((oh, ow), np.float64)
    mb = C.c_int(0)
    ctx.check(ctx._l.sr_block_search(ctx.h, C.byref(li), C.byref(ri), C.byref(params), C.c_void_p(out.ctypes.data),
                                     C.c_size_t(out.strides[0]), C.byref(mb)))
    return (out, mb.value) if return_max_block else out


def block_search_batch(ctx, lefts, rights, params, outs):
    """sr_block_search_batch: n pairs of one shape, pipelined H2D / kernels / D2H (config 5).
    lefts/rights: sequences of HxWx3 uint8 arrays; outs: sequence of HxW float64 arrays (written)."""
    n = len(lefts)
    if len(rights) != n or len(outs) != n:
        raise SrError(E_BADARG, "block_search_batch: lefts, rights and outs differ in length")
    if n == 0:
        return outs
    Ls = (SrImage * n)()
    Rs = (SrImage * n)()
    keep = []
    for i in range(n):
        la, li = image(lefts[i])
        ra, ri = image(rights[i])
        keep += [la, ra]
        Ls[i] = li
        Rs[i] = ri
    pitch = outs[0].strides[0]
    for o in outs:
        if o.dtype != np.float64 or o.strides[0] != pitch or o.strides[1] != 8:
            raise SrError(E_BADARG, "block_search_batch: outputs must be float64 with one common row pitch")
    optrs = (C.c_void_p * n)(*[o.ctypes.data for o in outs])
    ctx.check(ctx._l.sr_block_search_batch(ctx.h, n, Ls, Rs, C.byref(params), optrs, C.c_size_t(pitch)))
    return outs
