"""Host-side mirror of the reference's Reconstruction interface (reconstruction.h:13-34): same
function names, argument order, defaults and in-place behaviour, forwarding to the C-ABI.

    removeDisparityOutliers(disp, 500, 1.5, 0.8)          # in place, CV_32FC1
    scaleDisparityMap(disp, f)                            # in place
    depth = convertDisparityToDepth(disp, focal, baseline=1.0)
    reconstruction(bgr, depth, K, thrMesh=1.0)            # writes ../../results/reconstruction_mesh.off
    WriteMesh(vertices, w, h, filename, edgeThreshold) ; CheckTriangularValidity(vertices, a, b, c, thr)

numpy float32 maps stand for cv::Mat CV_32FC1, a 3x3 array for Eigen Matrix3f, a structured array of
capi.VERTEX_DTYPE for Vertex[].
"""
import ctypes as C

import numpy as np

from . import capi

MINF = -np.inf   # Eigen.h:20-22


def _ctx(ctx):
    return ctx or capi.default_context()


def _f32_inplace(a, name):
    if not (isinstance(a, np.ndarray) and a.dtype == np.float32 and a.ndim == 2 and a.strides[1] == 4):
        raise capi.SrError(capi.E_BADARG, f"{name}: needs a 2-D float32 map (CV_32FC1)")
    return a


def removeDisparityOutliers(disparityMap, kernelSize=500, thrFront=1.5, thrBack=0.8, *, ctx=None):
    c = _ctx(ctx)
    d = _f32_inplace(disparityMap, "removeDisparityOutliers")
    c.check(c._l.sr_remove_outliers(c.h, C.c_void_p(d.ctypes.data), d.shape[1], d.shape[0], C.c_size_t(d.strides[0]),
                                    int(kernelSize), C.c_float(thrFront), C.c_float(thrBack)))


def boxBlur(src, kernelSize, *, ctx=None):
    """cv::blur(src, Size(k,k)) on a CV_32FC1 map (the filter inside removeDisparityOutliers)."""
    c = _ctx(ctx)
    s = np.ascontiguousarray(src, dtype=np.float32)
    out = np.empty_like(s)
    c.check(c._l.sr_box_blur(c.h, C.c_void_p(s.ctypes.data), C.c_void_p(out.ctypes.data), s.shape[1], s.shape[0], int(kernelSize)))
    return out


def scaleDisparityMap(disparityMap, scalingFactor, *, ctx=None):
    c = _ctx(ctx)
    d = _f32_inplace(disparityMap, "scaleDisparityMap")
    c.check(c._l.sr_scale(c.h, C.c_void_p(d.ctypes.data), d.shape[1], d.shape[0], C.c_size_t(d.strides[0]), C.c_float(scalingFactor)))


def convertDisparityToDepth(dispImage, focalLength, baseline=1.0, *, ctx=None):
    c = _ctx(ctx)
    d = np.ascontiguousarray(dispImage, dtype=np.float32)
    out = np.empty_like(d)
    c.check(c._l.sr_disparity_to_depth(c.h, C.c_void_p(d.ctypes.data), C.c_void_p(out.ctypes.data), d.shape[1], d.shape[0],
                                       C.c_float(focalLength), C.c_float(baseline)))
    return out


def _intrinsics(K):
    K = np.asarray(K, dtype=np.float32)
    return float(K[0, 0]), float(K[1, 1]), float(K[0, 2]), float(K[1, 2])   # reconstruction.cpp:154-157


def backproject(bgrImage, depthValues, intrinsics, *, ctx=None):
    """The vertex loop of reconstruction() (reconstruction.cpp:169-196) -> Vertex[]."""
    c = _ctx(ctx)
    img, si = capi.image(bgrImage)
    d = np.ascontiguousarray(depthValues, dtype=np.float32)
    if d.shape != img.shape[:2]:
        raise capi.SrError(capi.E_BADARG, "reconstruction: bgrImage and depthValues sizes differ")
    fx, fy, cx, cy = _intrinsics(intrinsics)
    out = np.zeros(d.shape[0] * d.shape[1], dtype=capi.VERTEX_DTYPE)
    c.check(c._l.sr_backproject(c.h, C.c_void_p(img.ctypes.data), C.c_size_t(img.strides[0]), C.c_void_p(d.ctypes.data),
                                d.shape[1], d.shape[0], C.c_float(fx), C.c_float(fy), C.c_float(cx), C.c_float(cy),
                                C.c_void_p(out.ctypes.data)))
    return out


def meshFaces(vertices, width, height, edgeThreshold, *, ctx=None):
    """Face list of WriteMesh in the reference's order (reconstruction.cpp:92-113), n x 3 uint32."""
    c = _ctx(ctx)
    v = np.ascontiguousarray(vertices, dtype=capi.VERTEX_DTYPE)
    cells = max(width - 1, 0) * max(height - 1, 0)
    tri = np.zeros((max(cells, 1) * 2, 3), np.uint32)
    nf = C.c_uint32(0)
    c.check(c._l.sr_mesh_faces(c.h, C.c_void_p(v.ctypes.data), int(width), int(height), C.c_float(edgeThreshold),
                               C.c_void_p(tri.ctypes.data), C.byref(nf)))
    return tri[:nf.value].copy()


def CheckTriangularValidity(vertices, one, two, three, threshold, *, ctx=None):
    """reconstruction.cpp:46-69 for one triangle (evaluated on the GPU through a 2x2 grid)."""
    v = np.zeros(4, dtype=capi.VERTEX_DTYPE)
    v[0], v[2], v[1] = vertices[one], vertices[two], vertices[three]   # cell (a,b,c) = (0, 2, 1)
    v[3]["pos"] = MINF
    return len(meshFaces(v, 2, 2, threshold, ctx=ctx)) == 1


def WriteMesh(vertices, width, height, filename, edgeThreshold, *, ctx=None):
    c = _ctx(ctx)
    v = np.ascontiguousarray(vertices, dtype=capi.VERTEX_DTYPE)
    rc = c._l.sr_write_mesh(c.h, C.c_void_p(v.ctypes.data), int(width), int(height), str(filename).encode(), C.c_float(edgeThreshold))
    if rc == capi.E_IO:
        return False      # reconstruction.cpp:117
    c.check(rc)
    return True


def reconstruction(bgrImage, depthValues, intrinsics, thrMesh=1.0, *, filename="../../results/reconstruction_mesh.off", ctx=None):
    """reconstruction.cpp:152-208: back-project, then WriteMesh to the reference's fixed path."""
    v = backproject(bgrImage, depthValues, intrinsics, ctx=ctx)
    h, w = np.asarray(depthValues).shape
    if not WriteMesh(v, w, h, filename, thrMesh, ctx=ctx):
        print("Failed to write mesh!\nCheck file path!")    # :202-203


def reconstructFused(bgrImage, disparity, intrinsics, focalLength, baseline=1.0, kernelSize=500, thrFront=1.5, thrBack=0.8,
                     thrMesh=1.0, *, ctx=None):
    """main.cpp:53-65 in one device pipeline (sr_reconstruct): returns dict(disp, depth, vertices, faces)."""
    c = _ctx(ctx)
    img, si = capi.image(bgrImage)
    d = np.ascontiguousarray(disparity, dtype=np.float32)
    h, w = d.shape
    fx, fy, cx, cy = _intrinsics(intrinsics)
    disp_f = np.empty_like(d); depth = np.empty_like(d)
    vert = np.zeros(w * h, dtype=capi.VERTEX_DTYPE)
    tri = np.zeros((max((w - 1) * (h - 1), 1) * 2, 3), np.uint32)
    nf = C.c_uint32(0)
    c.check(c._l.sr_reconstruct(c.h, C.c_void_p(img.ctypes.data), C.c_size_t(img.strides[0]), C.c_void_p(d.ctypes.data), w, h,
                                int(kernelSize), C.c_float(thrFront), C.c_float(thrBack), C.c_float(focalLength), C.c_float(baseline),
                                C.c_float(fx), C.c_float(fy), C.c_float(cx), C.c_float(cy), C.c_float(thrMesh),
                                C.c_void_p(disp_f.ctypes.data), C.c_void_p(depth.ctypes.data), C.c_void_p(vert.ctypes.data),
                                C.c_void_p(tri.ctypes.data), C.byref(nf)))
    return {"disp": disp_f, "depth": depth, "vertices": vert, "faces": tri[:nf.value].copy()}


def pairReconstruct(pair, intrinsics, focalLength, baseline=1.0, quantize_u8=True, kernelSize=500, thrFront=1.5, thrBack=0.8,
                    thrMesh=1.0):
    """Search result already in HBM -> (8-bit hand-off) -> outliers -> depth -> vertices -> faces (sr_pair_reconstruct)."""
    c = pair.ctx
    h, w = pair.h2, pair.w2
    fx, fy, cx, cy = _intrinsics(intrinsics)
    disp_f = np.empty((h, w), np.float32); depth = np.empty((h, w), np.float32)
    vert = np.zeros(w * h, dtype=capi.VERTEX_DTYPE)
    tri = np.zeros((max((w - 1) * (h - 1), 1) * 2, 3), np.uint32)
    nf = C.c_uint32(0)
    c.check(c._l.sr_pair_reconstruct(c.h, pair.h, int(bool(quantize_u8)), int(kernelSize), C.c_float(thrFront), C.c_float(thrBack),
                                     C.c_float(focalLength), C.c_float(baseline), C.c_float(fx), C.c_float(fy), C.c_float(cx),
                                     C.c_float(cy), C.c_float(thrMesh), C.c_void_p(disp_f.ctypes.data), C.c_void_p(depth.ctypes.data),
                                     C.c_void_p(vert.ctypes.data), C.c_void_p(tri.ctypes.data), C.byref(nf)))
    return {"disp": disp_f, "depth": depth, "vertices": vert, "faces": tri[:nf.value].copy()}


def pairReconstructWarp(pair, M, dsize, colorImage, intrinsics, focalLength, baseline=1.0, quantize_u8=True, kernelSize=500,
                        thrFront=1.5, thrBack=0.8, thrMesh=1.0):
    """rectification.cpp:82-87 -> main.cpp:41-65 with the disparity map never leaving HBM (sr_pair_reconstruct_warp):
    search result in the rectified frame -> cv::warpPerspective(disp, M = H_.inv(), dsize, INTER_NEAREST) -> (8-bit
    hand-off) -> outliers -> depth -> vertices (coloured from colorImage, the un-rectified right image) -> faces.
    Returns dict(warped, disp, depth, vertices, faces)."""
    c = pair.ctx
    dw, dh = dsize
    fx, fy, cx, cy = _intrinsics(intrinsics)
    m = np.ascontiguousarray(M, dtype=np.float64).reshape(9)
    warped = np.empty((dh, dw), np.float64)
    disp_f = np.empty((dh, dw), np.float32); depth = np.empty((dh, dw), np.float32)
    vert = np.zeros(dw * dh, dtype=capi.VERTEX_DTYPE)
    tri = np.zeros((max((dw - 1) * (dh - 1), 1) * 2, 3), np.uint32)
    nf = C.c_uint32(0)
    if colorImage is not None:
        img, si = capi.image(colorImage)
        col = C.byref(si)
    else:
        col = None
    c.check(c._l.sr_pair_reconstruct_warp(c.h, pair.h, C.c_void_p(m.ctypes.data), int(dw), int(dh), col, int(bool(quantize_u8)),
                                          int(kernelSize), C.c_float(thrFront), C.c_float(thrBack), C.c_float(focalLength),
                                          C.c_float(baseline), C.c_float(fx), C.c_float(fy), C.c_float(cx), C.c_float(cy),
                                          C.c_float(thrMesh), C.c_void_p(warped.ctypes.data), C.c_void_p(disp_f.ctypes.data),
                                          C.c_void_p(depth.ctypes.data), C.c_void_p(vert.ctypes.data), C.c_void_p(tri.ctypes.data),
                                          C.byref(nf)))
    return {"warped": warped, "disp": disp_f, "depth": depth, "vertices": vert, "faces": tri[:nf.value].copy()}


def computeAverageDisparity(disparityMap, *, ctx=None, return_count=False):
    """utils.cpp:8-20: the float sum of the non-infinite pixels in raster order / their count (bit-identical)."""
    c = _ctx(ctx)
    d = np.ascontiguousarray(disparityMap, dtype=np.float32)
    avg = C.c_float(0); cnt = C.c_int(0)
    c.check(c._l.sr_average_disparity(c.h, C.c_void_p(d.ctypes.data), d.shape[1], d.shape[0], C.c_size_t(d.strides[0]), C.byref(avg), C.byref(cnt)))
    return (np.float32(avg.value), cnt.value) if return_count else np.float32(avg.value)


def warpPerspectiveNearest(src, M, dsize, *, ctx=None):
    """cv::warpPerspective(src, dst, M, dsize, INTER_NEAREST) on CV_64F (rectification.cpp:82-87)."""
    c = _ctx(ctx)
    s = np.ascontiguousarray(src, dtype=np.float64)
    m = np.ascontiguousarray(M, dtype=np.float64).reshape(9)
    dw, dh = dsize
    out = np.empty((dh, dw), np.float64)
    c.check(c._l.sr_warp_nearest_f64(c.h, C.c_void_p(s.ctypes.data), s.shape[1], s.shape[0], C.c_void_p(out.ctypes.data), int(dw), int(dh),
                                     C.c_void_p(m.ctypes.data)))
    return out


def warpPerspectiveLinear(image, M, dsize, *, ctx=None):
    """cv::warpPerspective(image, dst, M, dsize) with the default INTER_LINEAR / BORDER_CONSTANT(0) on CV_8UC3: the
    rectifying warp of the input images (rectification.cpp:486-493); bit-identical to cv2.warpPerspective."""
    c = _ctx(ctx)
    s = np.asarray(image)
    if s.dtype != np.uint8 or s.ndim != 3 or s.shape[2] != 3:
        raise capi.SrError(capi.E_BADARG, "image must be HxWx3 uint8 (CV_8UC3)")
    if s.strides[2] != 1 or s.strides[1] != 3:
        s = np.ascontiguousarray(s)
    m = np.ascontiguousarray(M, dtype=np.float64).reshape(9)
    dw, dh = dsize
    out = np.empty((dh, dw, 3), np.uint8)
    c.check(c._l.sr_warp_linear_bgr(c.h, C.c_void_p(s.ctypes.data), s.shape[1], s.shape[0], C.c_size_t(s.strides[0]),
                                    C.c_void_p(out.ctypes.data), int(dw), int(dh), C.c_size_t(out.strides[0]), C.c_void_p(m.ctypes.data)))
    return out


def rectifyImages(leftImage, rightImage, H, Hp, leftSize, rightSize, *, ctx=None):
    """The two warps at the end of ImageRectifier::computeRectification (rectification.cpp:483-493): left image by H_,
    right image by Hp_, into the rectified sizes (cols, rows) that function derives from the transformed corners."""
    return warpPerspectiveLinear(leftImage, H, leftSize, ctx=ctx), warpPerspectiveLinear(rightImage, Hp, rightSize, ctx=ctx)


def evaldisp(disp, gtdisp, mask, badthresh, maxdisp, rounddisp=0, *, ctx=None):
    """utils.cpp:123-168 -> dict(n, bad, invalid, bad_pct, invalid_pct, total_bad_pct, avg_err)."""
    c = _ctx(ctx)
    d = np.ascontiguousarray(disp, dtype=np.float32); g = np.ascontiguousarray(gtdisp, dtype=np.float32)
    m = np.ascontiguousarray(mask, dtype=np.uint8)
    e = capi.SrEval()
    c.check(c._l.sr_evaldisp(c.h, C.c_void_p(d.ctypes.data), C.c_void_p(g.ctypes.data), C.c_void_p(m.ctypes.data), d.shape[1], d.shape[0],
                             C.c_float(badthresh), C.c_float(maxdisp), int(rounddisp), C.byref(e)))
    return {k: getattr(e, k) for k, _ in capi.SrEval._fields_}


def writePFM(path, data):
    d = np.ascontiguousarray(data, dtype=np.float32)
    rc = capi.lib().sr_write_pfm(str(path).encode(), C.c_void_p(d.ctypes.data), d.shape[1], d.shape[0])
    if rc != capi.OK:
        raise capi.SrError(rc, f"cannot write {path}")


def writePLY(path, vertices, width, height):
    v = np.ascontiguousarray(vertices, dtype=capi.VERTEX_DTYPE)
    rc = capi.lib().sr_write_ply(str(path).encode(), C.c_void_p(v.ctypes.data), int(width), int(height))
    if rc != capi.OK:
        raise capi.SrError(rc, f"cannot write {path}")
