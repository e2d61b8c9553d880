#!/usr/bin/env python
"""Generates tests/golden/*.npz by EXECUTING the reference's arithmetic through cv2.

The reference's hot path cannot be compiled in this image (no OpenCV C++/Eigen), but Python cv2
exposes the very functions its arithmetic lives in.  Each function below is a line-by-line
transliteration of the cited reference lines that calls cv2.absdiff / cv2.norm / cv2.mean /
cv2.subtract / cv2.blur / cv2.warpPerspective / cv2.imwrite on ROI views, exactly like the C++.
It is slow (~1e5 candidates/s) so the vectors are small crops.

Run in the build container (needs /root/reference for the real-image crops):
    python tests/golden/make_golden.py
The outputs are committed; the GPU box only reads the .npz files.
"""
import os
import sys
import tempfile

import cv2
import numpy as np

REF = "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))
DBL_MAX = np.finfo(np.float64).max


def ref_right(left, right, block, min_d, max_d, smooth, var_block=False, thres=19.0):
    """BlockSearch::computeDisparityMapRight, src/WindowSearch/BlockSearch.cpp:88-179."""
    h1, w1 = left.shape[:2]
    h2, w2 = right.shape[:2]
    height = min(h1, h2)
    D = np.zeros((h2, w2), np.float64)
    max_block = block
    for y in range(height):
        for x in range(w2):
            if not right[y, x].any():
                continue
            min_x, best = 0, DBL_MAX
            bs = block
            l, r = min(x, (bs - 1) // 2), min(w2 - x - 1, (bs - 1) // 2)
            u, dn = min(y, (bs - 1) // 2), min(h2 - y - 1, (bs - 1) // 2)
            rw = right[y - u:y + dn, x - l:x + r]
            if rw.size:
                mean = cv2.mean(rw)
                s = cv2.subtract(rw, mean)
            else:
                s = rw
            guard = 0
            while var_block and (cv2.norm(s, cv2.NORM_L2) if s.size else 0.0) < thres:
                prev = (l, r, u, dn)
                bs += 4
                l, r = min(x, (bs - 1) // 2), min(w2 - x - 1, (bs - 1) // 2)
                u, dn = min(y, (bs - 1) // 2), min(h2 - y - 1, (bs - 1) // 2)
                rw = right[y - u:y + dn, x - l:x + r]
                mean = cv2.mean(rw)
                s = cv2.subtract(rw, mean)
                if prev == (l, r, u, dn):   # the reference loops forever here; documented cap
                    break
            max_block = max(max_block, bs)
            for cx in range(x + min_d, x + max_d):
                if cx + r >= w1:
                    break
                lw = left[y - u:y + dn, cx - l:cx + r]
                if lw.shape != rw.shape:
                    raise RuntimeError("cv::Exception: ROI outside the left image")
                if lw.size:
                    dist = cv2.norm(cv2.absdiff(lw, rw), cv2.NORM_L2)
                else:
                    dist = 0.0
                with np.errstate(divide="ignore", invalid="ignore"):
                    dist = np.float64(dist) / np.float64((l + r) * (u + dn))
                if y >= 1 and D[y - 1, x] == float(x - cx):
                    dist *= smooth
                if x >= 1 and D[y, x - 1] == float(x - cx):
                    dist *= smooth
                if dist < best:
                    min_x, best = cx, dist
            D[y, x] = float(min_x - x)
    return D, max_block


def ref_left(left, right, block, max_d, smooth):
    """BlockSearch::computeDisparityMapLeft, src/WindowSearch/BlockSearch.cpp:24-86."""
    h1, w1 = left.shape[:2]
    h2, w2 = right.shape[:2]
    height = min(h1, h2)
    hb = (block - 1) // 2
    D = np.zeros((h1, w1), np.float64)
    for y in range(hb, height - hb):
        for x in range(hb, w1 - hb):
            if not left[y, x].any():
                continue
            lw = left[y - hb:y - hb + block, x - hb:x - hb + block]
            min_x, best = 0, DBL_MAX
            for cx in range(x - max_d, x):
                if cx < hb or cx >= w2 - hb:
                    continue
                rw = right[y - hb:y - hb + block, cx - hb:cx - hb + block]
                dist = cv2.norm(cv2.absdiff(lw, rw), cv2.NORM_L2)
                if y >= 1 and D[y - 1, x] == float(x - cx):
                    dist *= smooth
                if x >= 1 and D[y, x - 1] == float(x - cx):
                    dist *= smooth
                if dist < best:
                    min_x, best = cx, dist
            D[y, x] = float(x - min_x)
    return D


def ref_linear(left, right, smooth):
    """LinearSearch::computeDisparityMap, src/WindowSearch/LinearSearch.cpp:10-59 (k clamped to < w1)."""
    h1, w1 = left.shape[:2]
    h2, w2 = right.shape[:2]
    D = np.zeros((h2, w2), np.float64)
    for i in range(h2):
        for j in range(w2):
            if not left[i, j].any():
                continue
            col, best = 0, DBL_MAX
            pr = right[i, j].astype(np.int64)
            for k in range(j, min(j + 200, w1)):
                pl = left[i, k].astype(np.int64)
                dist = float(np.sqrt(float((pr[0] - pl[0]) ** 2) + float((pr[1] - pl[1]) ** 2) + float((pr[2] - pl[2]) ** 2)))
                if i >= 1 and D[i - 1, j] == float(j - k):
                    dist *= smooth
                if j >= 1 and D[i, j - 1] == float(j - k):
                    dist *= smooth
                if dist < best:
                    col, best = k, dist
            D[i, j] = float(col - j)
    return D


def ref_remove_outliers(disp, k, front, back):
    """removeDisparityOutliers, src/Reconstruction/reconstruction.cpp:5-18."""
    d = disp.astype(np.float32).copy()
    b = cv2.blur(d, (k, k))
    front, back = np.float32(front), np.float32(back)
    m = (d > front * b) | (d < back * b)
    d[m] = b[m]
    return d, b


def crop(img, y0, x0, h, w):
    return np.ascontiguousarray(img[y0:y0 + h, x0:x0 + w])


def main():
    rng = np.random.default_rng(0xB200)
    cases = {}

    teddyL = cv2.imread(f"{REF}/results/Rectified/trainingH/Teddy/rectifiedLeft.png")
    teddyR = cv2.imread(f"{REF}/results/Rectified/trainingH/Teddy/rectifiedRight.png")
    motoL = cv2.imread(f"{REF}/results/Rectified/trainingH/MotorcycleE/rectifiedLeft.png")
    motoR = cv2.imread(f"{REF}/results/Rectified/trainingH/MotorcycleE/rectifiedRight.png")
    adirL = cv2.imread(f"{REF}/data/MiddEval3/trainingH/Adirondack/im0.png")
    adirR = cv2.imread(f"{REF}/data/MiddEval3/trainingH/Adirondack/im1.png")

    def add_right(name, L, R, block, mn, mx, s, vb=False, th=19.0):
        D, mb = ref_right(L, R, block, mn, mx, s, vb, th)
        cases[name] = dict(kind="right", left=L, right=R, block=block, min_d=mn, max_d=mx, smooth=s, var_block=int(vb),
                           thres=th, out=D, max_block=mb)
        print(name, D.shape, "nonzero", int((D != 0).sum()), "maxblock", mb, flush=True)

    # real rectified crops (equal sizes), interior + image-border windows
    L = crop(teddyL, 300, 380, 40, 96); R = crop(teddyR, 300, 380, 40, 72)
    add_right("right_teddy_b9_s1", L, R, 9, 0, 24, 1.0)
    add_right("right_teddy_b9_s09", L, R, 9, 0, 24, 0.9)
    add_right("right_teddy_b17_s09", L, R, 17, 0, 20, 0.9)
    add_right("right_teddy_b5_min3", L, R, 5, 3, 19, 0.9)
    add_right("right_teddy_b8_even", L, R, 8, 0, 16, 0.9)
    # dataset pair (im0/im1 are already rectified), flat region -> ties and zero-disparity runs
    L = crop(adirL, 100, 200, 36, 80); R = crop(adirR, 100, 200, 36, 64)
    add_right("right_adir_b9_s09", L, R, 9, 0, 16, 0.9)
    add_right("right_adir_b7_s05", L, R, 7, 0, 16, 0.5)
    add_right("right_adir_b7_s12", L, R, 7, 0, 16, 1.2)
    # black borders, unequal sizes (w1 < w2 triggers the "no candidate" D = -x case; h1 > h2)
    L = crop(motoL, 10, 700, 40, 60); R = crop(motoR, 10, 700, 36, 70)
    add_right("right_moto_black_uneq", L, R, 9, 0, 12, 0.9)
    L = crop(motoL, 500, 1400, 40, 81); R = crop(motoR, 500, 1405, 36, 90)
    add_right("right_moto_edge_uneq", L, R, 9, 0, 20, 0.9)
    # quantised image -> many exact ties
    q = (rng.integers(0, 3, (28, 60, 3)) * 100 + 20).astype(np.uint8)
    Lq = np.roll(q, 5, axis=1).copy(); Rq = q[:, :52].copy()
    add_right("right_ties_b5", Lq, Rq, 5, 0, 12, 0.9)
    add_right("right_ties_b3", Lq, Rq, 3, 0, 12, 1.0)
    add_right("right_ties_b1_empty_window", Lq[:6, :20].copy(), Rq[:6, :16].copy(), 1, 0, 4, 0.9)
    # varBlock
    L = crop(adirL, 400, 600, 30, 72); R = crop(adirR, 400, 600, 30, 60)
    add_right("right_adir_varblock", L, R, 5, 0, 12, 0.9, True, 19.0)
    add_right("right_adir_varblock_t60", L, R, 5, 0, 12, 1.0, True, 60.0)

    def add_left(name, L, R, block, mx, s):
        D = ref_left(L, R, block, mx, s)
        cases[name] = dict(kind="left", left=L, right=R, block=block, max_d=mx, smooth=s, out=D)
        print(name, D.shape, "nonzero", int((D != 0).sum()), flush=True)

    L = crop(teddyL, 300, 380, 36, 80); R = crop(teddyR, 300, 380, 36, 80)
    add_left("left_teddy_b5_s09", L, R, 5, 20, 0.9)
    add_left("left_teddy_b9_s1", L, R, 9, 20, 1.0)
    add_left("left_teddy_b9_s09", L, R, 9, 20, 0.9)
    add_left("left_ties_b3", Lq, np.ascontiguousarray(q[:, :56]), 3, 10, 0.7)
    L = crop(motoL, 10, 690, 34, 64); R = crop(motoR, 10, 690, 38, 58)
    add_left("left_moto_black_uneq", L, R, 5, 14, 0.9)

    L = crop(teddyL, 300, 380, 12, 260); R = crop(teddyR, 300, 380, 12, 230)
    D = ref_linear(L, R, 0.9)
    cases["linear_teddy"] = dict(kind="linear", left=L, right=R, smooth=0.9, out=D)
    print("linear_teddy", D.shape, flush=True)
    L = crop(motoL, 16, 660, 10, 120); R = crop(motoR, 16, 660, 10, 100)
    cases["linear_moto_black"] = dict(kind="linear", left=L, right=R, smooth=1.0, out=ref_linear(L, R, 1.0))

    # Reconstruction: cv::blur-based outlier removal on integer-valued (u8-derived) and fractional maps
    tr = cv2.imread(f"{REF}/results/PerceptualWindowSearch/test_result.png", cv2.IMREAD_GRAYSCALE)
    dsmall = tr[200:320, 300:460].astype(np.float32)
    for k in (15, 50, 500):
        d, b = ref_remove_outliers(dsmall, k, 1.5, 0.8)
        cases[f"outliers_int_k{k}"] = dict(kind="outliers", disp=dsmall, k=k, front=1.5, back=0.8, out=d, blur=b)
    dfrac = (dsmall * np.float32(0.37)).astype(np.float32)
    d, b = ref_remove_outliers(dfrac, 21, 1.5, 0.8)
    cases["outliers_frac_k21"] = dict(kind="outliers", disp=dfrac, k=21, front=1.5, back=0.8, out=d, blur=b)
    # full-size blur statistic (main.cpp:53 on the reference's own 900x750 test image): keep a checksum + a sparse sample
    full = tr.astype(np.float32)
    d, b = ref_remove_outliers(full, 500, 1.5, 0.8)
    cases["outliers_full_teddy_k500"] = dict(kind="outliers_full", disp_u8=tr, k=500, front=1.5, back=0.8,
                                             out_sample=d[::25, ::25].copy(), blur_sample=b[::25, ::25].copy(),
                                             out_sum=np.float64(d.astype(np.float64).sum()), changed=int((d != full).sum()))

    # nearest-neighbour homography warp of a CV_64F map (rectification.cpp:82-87)
    src = rng.integers(0, 200, (60, 90)).astype(np.float64)
    M = np.array([[1.02, 0.03, -4.0], [-0.02, 0.98, 3.0], [1e-4, -2e-4, 1.0]])
    dst = cv2.warpPerspective(src, M, (100, 70), flags=cv2.INTER_NEAREST)
    cases["warp_nearest"] = dict(kind="warp", src=src, M=M, dw=100, dh=70, out=dst)

    # imwrite(CV_64F) -> u8 -> float32 hand-off (main.cpp:41-50)
    vals = np.array([[0, 1.4, 1.5, 2.5, 3.5, 199, 254.5, 255, 255.5, 300, -3, -0.5, 0.5, 127.5, 128.5, 1e6]], np.float64)
    with tempfile.TemporaryDirectory() as td:
        p = os.path.join(td, "d.png")
        cv2.imwrite(p, vals)
        back = cv2.imread(p, cv2.IMREAD_GRAYSCALE).astype(np.float32)
    cases["quantize_u8"] = dict(kind="quantize", vals=vals, out=back)

    # cv2.subtract(u8, scalar mean) semantics used by varBlock
    w = rng.integers(0, 256, (9, 7, 3)).astype(np.uint8)
    mean = cv2.mean(w)
    cases["subtract_mean"] = dict(kind="subtract", win=w, out=cv2.subtract(w, mean), norm=np.float64(cv2.norm(cv2.subtract(w, mean), cv2.NORM_L2)))

    for name, c in cases.items():
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **c)
    print("wrote", len(cases), "fixtures to", OUT)


if __name__ == "__main__":
    sys.exit(main())
