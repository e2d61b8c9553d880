"""ctypes binding of oracle/_ref/libsr_ref.so: the reference's OWN hot-path sources (BlockSearch.cpp, LinearSearch.cpp,
reconstruction.cpp), compiled unchanged against the header stand-ins of oracle/refstub.  TEST INFRASTRUCTURE ONLY --
only tests/, __graft_entry__.smoke()/build() and bench.py's cpu_baseline / --impl reference legs may import this.

The library is built where /root/reference exists (the build container; `make -C oracle ref`) and travels to the GPU
box as a git-ignored .so; nothing here reads /root/reference at run time."""
import ctypes as C
import os
import subprocess

import numpy as np

from .orc import VERTEX_DTYPE, _img, _f32

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_ref", "libsr_ref.so")
_LIB = None
OK, E_BADARG, E_SHAPE, E_IO = 0, -1, -2, -5


class RefError(RuntimeError):
    def __init__(self, code, what, msg=""):
        super().__init__(f"{what}: reference error {code} {msg}")
        self.code = code


def available():
    return os.path.exists(_SO) or os.path.isdir("/root/reference/src")


def build():
    """(Re)build where the reference tree exists; otherwise the prebuilt .so is used as it came."""
    if os.path.isdir("/root/reference/src"):
        subprocess.check_call(["make", "-C", _HERE, "ref"], stdout=subprocess.DEVNULL)
    return _SO


def lib():
    global _LIB
    if _LIB is None:
        build()
        _LIB = C.CDLL(_SO)
        _LIB.ref_last_error.restype = C.c_char_p
    return _LIB


def _chk(rc, what):
    if rc < 0:
        raise RefError(rc, what, lib().ref_last_error().decode(errors="replace"))
    return rc


def max_threads():
    return int(lib().ref_max_threads())


def block_search_right(left, right, block, min_disp, max_disp, smooth, var_block=False, thres=19.0, return_max_block=False):
    """BlockSearch(left, right, block, min_disp, max_disp).computeDisparityMapRight(smooth, var_block, thres)"""
    L, lp, w1, h1, p1 = _img(left)
    R, rp, w2, h2, p2 = _img(right)
    out = np.empty((h2, w2), np.float64)
    mb = C.c_int(0)
    _chk(lib().ref_block_search_right(lp, w1, h1, p1, rp, w2, h2, p2, int(block), int(min_disp), int(max_disp), C.c_double(smooth),
                                      int(bool(var_block)), C.c_double(thres), out.ctypes.data_as(C.c_void_p), C.byref(mb)), "block_search_right")
    return (out, mb.value) if return_max_block else out


def block_search_right_rows(left, right, block, min_disp, max_disp, rows, nthreads=0):
    """smoothFactor 1 on output rows [rows[0], rows[1]) only, row bands on `nthreads` threads; -> (map, threads used)"""
    L, lp, w1, h1, p1 = _img(left)
    R, rp, w2, h2, p2 = _img(right)
    out = np.zeros((h2, w2), np.float64)
    used = _chk(lib().ref_block_search_right_rows(lp, w1, h1, p1, rp, w2, h2, p2, int(block), int(min_disp), int(max_disp),
                                                  int(rows[0]), int(rows[1]), int(nthreads), out.ctypes.data_as(C.c_void_p)), "block_search_right_rows")
    return out, used


def block_search_left(left, right, block, max_disp, smooth):
    L, lp, w1, h1, p1 = _img(left)
    R, rp, w2, h2, p2 = _img(right)
    out = np.empty((h1, w1), np.float64)
    _chk(lib().ref_block_search_left(lp, w1, h1, p1, rp, w2, h2, p2, int(block), int(max_disp), C.c_double(smooth),
                                     out.ctypes.data_as(C.c_void_p)), "block_search_left")
    return out


def linear_search(left, right, smooth, return_printed=False):
    L, lp, w1, h1, p1 = _img(left)
    R, rp, w2, h2, p2 = _img(right)
    out = np.empty((h2, w2), np.float64)
    buf = C.create_string_buffer(4096)
    _chk(lib().ref_linear_search(lp, w1, h1, p1, rp, w2, h2, p2, C.c_double(smooth), out.ctypes.data_as(C.c_void_p), buf, 4096), "linear_search")
    return (out, buf.value.decode()) if return_printed else out


def remove_outliers(disp, ksize=500, thr_front=1.5, thr_back=0.8):
    d = _f32(disp).copy()
    _chk(lib().ref_remove_outliers(d.ctypes.data_as(C.c_void_p), d.shape[1], d.shape[0], int(ksize), C.c_float(thr_front), C.c_float(thr_back)), "remove_outliers")
    return d


def scale(disp, f):
    d = _f32(disp).copy()
    _chk(lib().ref_scale(d.ctypes.data_as(C.c_void_p), d.shape[1], d.shape[0], C.c_float(f)), "scale")
    return d


def disparity_to_depth(disp, focal, baseline=1.0):
    d = _f32(disp)
    o = np.empty_like(d)
    _chk(lib().ref_disparity_to_depth(d.ctypes.data_as(C.c_void_p), o.ctypes.data_as(C.c_void_p), d.shape[1], d.shape[0],
                                      C.c_float(focal), C.c_float(baseline)), "disparity_to_depth")
    return o


def triangle_valid(vertices, a, b, c, thr):
    v = np.ascontiguousarray(vertices)
    return bool(lib().ref_triangle_valid(v.ctypes.data_as(C.c_void_p), int(a), int(b), int(c), C.c_float(thr)))


def write_mesh(path, vertices, w, h, thr):
    v = np.ascontiguousarray(vertices)
    assert v.dtype == VERTEX_DTYPE and v.ctypes.data % 16 == 0
    _chk(lib().ref_write_mesh(v.ctypes.data_as(C.c_void_p), int(w), int(h), os.fsencode(path), C.c_float(thr)), "write_mesh")


def reconstruction(bgr, depth, fx, fy, cx, cy, thr, root):
    """reconstruction(bgr, depth, K, thr) -> path of the mesh file it wrote (<root>/results/reconstruction_mesh.off)"""
    I, ip, w, h, p = _img(bgr)
    d = _f32(depth)
    os.makedirs(os.path.join(root, "a", "b"), exist_ok=True)
    os.makedirs(os.path.join(root, "results"), exist_ok=True)
    _chk(lib().ref_reconstruction(ip, p, d.ctypes.data_as(C.c_void_p), w, h, C.c_float(fx), C.c_float(fy), C.c_float(cx), C.c_float(cy),
                                  C.c_float(thr), os.fsencode(os.path.join(root, "a", "b"))), "reconstruction")
    return os.path.join(root, "results", "reconstruction_mesh.off")


# ---- the stand-in's OpenCV functions (held against cv2 in tests/test_ref_build.py) ----
def cv_absdiff_norm(a, b):
    A, ap, w, h, _ = _img(a)
    B, bp, _, _, _ = _img(b)
    diff = np.empty_like(A)
    n = C.c_double(0)
    lib().ref_cv_absdiff_norm(ap, bp, w, h, diff.ctypes.data_as(C.c_void_p), C.byref(n))
    return diff, n.value


def cv_mean_subtract(a):
    A, ap, w, h, _ = _img(a)
    sub = np.empty_like(A)
    mean = (C.c_double * 3)()
    n = C.c_double(0)
    lib().ref_cv_mean_subtract(ap, w, h, mean, sub.ctypes.data_as(C.c_void_p), C.byref(n))
    return np.array(list(mean)), sub, n.value


def cv_blur(src, k):
    s = _f32(src)
    d = np.empty_like(s)
    lib().ref_cv_blur(s.ctypes.data_as(C.c_void_p), d.ctypes.data_as(C.c_void_p), s.shape[1], s.shape[0], int(k))
    return d


def roi_throws(w, h, x, y, rw, rh):
    return bool(lib().ref_roi_throws(w, h, x, y, rw, rh))
