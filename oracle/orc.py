"""ctypes binding of the CPU ORACLE (oracle/libsr_oracle.so).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; the product package (stereo_reconstruction_b200) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

COST_SSD_L2, COST_SAD, COST_NCC = 0, 1, 2
OK, E_BADARG, E_SHAPE, E_IO = 0, -1, -2, -5

VERTEX_DTYPE = np.dtype({"names": ["pos", "rgba"], "formats": [(np.float32, 4), (np.uint8, 4)],
                         "offsets": [0, 16], "itemsize": 32})


class OracleError(RuntimeError):
    def __init__(self, code, what):
        super().__init__(f"{what}: oracle error {code}")
        self.code = code


def build(force=False):
    so = os.path.join(_HERE, "libsr_oracle.so")
    src = [os.path.join(_HERE, f) for f in ("sr_oracle.c", "sr_oracle.h")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "libsr_oracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        _LIB.orc_average_disparity.restype = C.c_float
    return _LIB


def _img(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    assert a.ndim == 3 and a.shape[2] == 3
    return a, a.ctypes.data_as(C.c_void_p), a.shape[1], a.shape[0], C.c_size_t(a.strides[0])


def _chk(rc, what):
    if rc != 0:
        raise OracleError(rc, what)


def block_search_right(left, right, block, min_disp, max_disp, smooth, var_block=False, thres=19.0, return_max_block=False):
    L, lp, w1, h1, p1 = _img(left)
    R, rp, w2, h2, p2 = _img(right)
    out = np.empty((h2, w2), np.float64)
    mb = C.c_int(0)
    _chk(lib().orc_block_search_right(lp, w1, h1, p1, rp, w2, h2, p2, int(block), int(min_disp), int(max_disp),
                                      C.c_double(smooth), int(bool(var_block)), C.c_double(thres),
                                      out.ctypes.data_as(C.c_void_p), C.byref(mb)), "block_search_right")
    return (out, mb.value) if return_max_block else out


def block_search_right_fast(left, right, block, min_disp, max_disp, smooth, cost=COST_SSD_L2, rows=None, nthreads=0):
    L, lp, w1, h1, p1 = _img(left)
    R, rp, w2, h2, p2 = _img(right)
    out = np.zeros((h2, w2), np.float64)
    r0, r1 = rows if rows is not None else (0, h2)
    _chk(lib().orc_block_search_right_fast(lp, w1, h1, p1, rp, w2, h2, p2, int(block), int(min_disp), int(max_disp),
                                           C.c_double(smooth), int(cost), int(r0), int(r1),
                                           out.ctypes.data_as(C.c_void_p), int(nthreads)), "block_search_right_fast")
    return out


def block_search_left(left, right, block, max_disp, smooth):
    L, lp, w1, h1, p1 = _img(left)
    R, rp, w2, h2, p2 = _img(right)
    out = np.empty((h1, w1), np.float64)
    _chk(lib().orc_block_search_left(lp, w1, h1, p1, rp, w2, h2, p2, int(block), int(max_disp), C.c_double(smooth),
                                     out.ctypes.data_as(C.c_void_p)), "block_search_left")
    return out


def linear_search(left, right, smooth):
    L, lp, w1, h1, p1 = _img(left)
    R, rp, w2, h2, p2 = _img(right)
    out = np.empty((h2, w2), np.float64)
    _chk(lib().orc_linear_search(lp, w1, h1, p1, rp, w2, h2, p2, C.c_double(smooth),
                                 out.ctypes.data_as(C.c_void_p)), "linear_search")
    return out


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def box_blur(src, ksize):
    s = _f32(src)
    d = np.empty_like(s)
    _chk(lib().orc_box_blur(s.ctypes.data_as(C.c_void_p), d.ctypes.data_as(C.c_void_p), s.shape[1], s.shape[0], int(ksize)), "box_blur")
    return d


def remove_outliers(disp, ksize=500, thr_front=1.5, thr_back=0.8):
    d = _f32(disp).copy()
    _chk(lib().orc_remove_outliers(d.ctypes.data_as(C.c_void_p), d.shape[1], d.shape[0], int(ksize),
                                   C.c_float(thr_front), C.c_float(thr_back)), "remove_outliers")
    return d


def scale(disp, f):
    d = _f32(disp).copy()
    _chk(lib().orc_scale(d.ctypes.data_as(C.c_void_p), d.shape[1], d.shape[0], C.c_float(f)), "scale")
    return d


def disparity_to_depth(disp, focal, baseline=1.0):
    d = _f32(disp)
    o = np.empty_like(d)
    _chk(lib().orc_disparity_to_depth(d.ctypes.data_as(C.c_void_p), o.ctypes.data_as(C.c_void_p), d.shape[1], d.shape[0],
                                      C.c_float(focal), C.c_float(baseline)), "disparity_to_depth")
    return o


def backproject(bgr, depth, fx, fy, cx, cy):
    I, ip, w, h, p = _img(bgr)
    d = _f32(depth)
    assert d.shape == (h, w)
    v = np.zeros(h * w, VERTEX_DTYPE)
    _chk(lib().orc_backproject(ip, p, d.ctypes.data_as(C.c_void_p), w, h, C.c_float(fx), C.c_float(fy), C.c_float(cx),
                               C.c_float(cy), v.ctypes.data_as(C.c_void_p)), "backproject")
    return v


def mesh_faces(vertices, w, h, thr):
    v = np.ascontiguousarray(vertices)
    tri = np.empty((max(2 * (w - 1) * (h - 1), 1), 3), np.uint32)
    n = C.c_uint32(0)
    _chk(lib().orc_mesh_faces(v.ctypes.data_as(C.c_void_p), w, h, C.c_float(thr), tri.ctypes.data_as(C.c_void_p), C.byref(n)), "mesh_faces")
    return tri[: n.value].copy()


def write_off(path, vertices, w, h, thr):
    v = np.ascontiguousarray(vertices)
    _chk(lib().orc_write_off(os.fsencode(path), v.ctypes.data_as(C.c_void_p), w, h, C.c_float(thr)), "write_off")


def write_pfm(path, data):
    d = _f32(data)
    _chk(lib().orc_write_pfm(os.fsencode(path), d.ctypes.data_as(C.c_void_p), d.shape[1], d.shape[0]), "write_pfm")


def write_ply(path, vertices, w, h):
    v = np.ascontiguousarray(vertices)
    _chk(lib().orc_write_ply(os.fsencode(path), v.ctypes.data_as(C.c_void_p), w, h), "write_ply")


class Eval(C.Structure):
    _fields_ = [("n", C.c_int), ("bad", C.c_int), ("invalid", C.c_int), ("serr", C.c_float), ("bad_pct", C.c_float),
                ("invalid_pct", C.c_float), ("total_bad_pct", C.c_float), ("avg_err", C.c_float)]


def evaldisp(disp, gt, mask, badthresh, maxdisp, rounddisp=0):
    d, g = _f32(disp), _f32(gt)
    m = np.ascontiguousarray(mask, np.uint8)
    e = Eval()
    _chk(lib().orc_evaldisp(d.ctypes.data_as(C.c_void_p), g.ctypes.data_as(C.c_void_p), m.ctypes.data_as(C.c_void_p),
                            d.shape[1], d.shape[0], C.c_float(badthresh), C.c_float(maxdisp), int(rounddisp), C.byref(e)), "evaldisp")
    return e


def average_disparity(disp):
    d = _f32(disp)
    return float(lib().orc_average_disparity(d.ctypes.data_as(C.c_void_p), d.shape[1], d.shape[0]))


def warp_nearest_f64(src, M, dsize):
    s = np.ascontiguousarray(src, np.float64)
    dw, dh = dsize
    dst = np.empty((dh, dw), np.float64)
    m = np.ascontiguousarray(M, np.float64).reshape(9)
    _chk(lib().orc_warp_nearest_f64(s.ctypes.data_as(C.c_void_p), s.shape[1], s.shape[0], dst.ctypes.data_as(C.c_void_p),
                                    dw, dh, m.ctypes.data_as(C.c_void_p)), "warp")
    return dst


def warp_linear_bgr(src, M, dsize):
    """cv::warpPerspective(image, dst, M, dsize) with the default flags on CV_8UC3 (rectification.cpp:486-493)."""
    s = np.ascontiguousarray(src, np.uint8)
    assert s.ndim == 3 and s.shape[2] == 3
    dw, dh = dsize
    dst = np.empty((dh, dw, 3), np.uint8)
    m = np.ascontiguousarray(M, np.float64).reshape(9)
    _chk(lib().orc_warp_linear_bgr(s.ctypes.data_as(C.c_void_p), s.shape[1], s.shape[0], C.c_size_t(s.strides[0]),
                                   dst.ctypes.data_as(C.c_void_p), dw, dh, C.c_size_t(dst.strides[0]),
                                   m.ctypes.data_as(C.c_void_p)), "warp_linear")
    return dst


def quantize_u8(disp):
    d = np.ascontiguousarray(disp, np.float64)
    o = np.empty(d.shape, np.float32)
    _chk(lib().orc_quantize_u8(d.ctypes.data_as(C.c_void_p), d.size, o.ctypes.data_as(C.c_void_p)), "quantize")
    return o


def search_ext(left, right, block, min_disp, max_disp, cost, left_ref=False, want_cost=False, want_subpix=False):
    L, lp, w1, h1, p1 = _img(left)
    R, rp, w2, h2, p2 = _img(right)
    shape = (h1, w1) if left_ref else (h2, w2)
    disp = np.zeros(shape, np.int32)
    ci = np.zeros(shape, np.int64) if want_cost else None
    sp = np.zeros(shape, np.float32) if want_subpix else None
    _chk(lib().orc_search_ext(lp, w1, h1, p1, rp, w2, h2, p2, int(block), int(min_disp), int(max_disp), int(cost),
                              int(bool(left_ref)), disp.ctypes.data_as(C.c_void_p),
                              ci.ctypes.data_as(C.c_void_p) if want_cost else None,
                              sp.ctypes.data_as(C.c_void_p) if want_subpix else None), "search_ext")
    return disp, ci, sp


def search_ext_fast(left, right, block, min_disp, max_disp, cost, left_ref=False, want_cost=False, want_subpix=False, nthreads=0):
    """search_ext for SSD / SAD through sliding box sums + OpenMP: same results, whole pairs in seconds."""
    L, lp, w1, h1, p1 = _img(left)
    R, rp, w2, h2, p2 = _img(right)
    shape = (h1, w1) if left_ref else (h2, w2)
    disp = np.zeros(shape, np.int32)
    ci = np.zeros(shape, np.int64) if want_cost else None
    sp = np.zeros(shape, np.float32) if want_subpix else None
    _chk(lib().orc_search_ext_fast(lp, w1, h1, p1, rp, w2, h2, p2, int(block), int(min_disp), int(max_disp), int(cost),
                                   int(bool(left_ref)), disp.ctypes.data_as(C.c_void_p),
                                   ci.ctypes.data_as(C.c_void_p) if want_cost else None,
                                   sp.ctypes.data_as(C.c_void_p) if want_subpix else None, int(nthreads)), "search_ext_fast")
    return disp, ci, sp


def lr_check(disp_r, disp_l):
    dr = np.ascontiguousarray(disp_r, np.int32)
    dl = np.ascontiguousarray(disp_l, np.int32)
    out = np.empty_like(dr)
    _chk(lib().orc_lr_check(dr.ctypes.data_as(C.c_void_p), dr.shape[1], dr.shape[0], dl.ctypes.data_as(C.c_void_p),
                            dl.shape[1], dl.shape[0], out.ctypes.data_as(C.c_void_p)), "lr_check")
    return out


def synth_pair(seed, pair, w, h, min_disp, max_disp, quant=False):
    left = np.empty((h, w, 3), np.uint8)
    right = np.empty((h, w, 3), np.uint8)
    lib().orc_synth_pair(C.c_uint64(seed), C.c_uint64(pair), w, h, int(min_disp), int(max_disp), int(bool(quant)),
                         left.ctypes.data_as(C.c_void_p), right.ctypes.data_as(C.c_void_p))
    return left, right
