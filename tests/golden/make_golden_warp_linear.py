"""Golden vectors for the bilinear rectifying warp (rectification.cpp:486-493: cv::warpPerspective(image, dst, H, size)
with the default INTER_LINEAR / BORDER_CONSTANT on CV_8UC3), produced by the real cv2.warpPerspective in the build
container (OpenCV's fixed-point remap path: 1/32-pixel positions, 15-bit weights): every one of the 1024 weight-table
entries, borders on all four sides, strong perspective with a vanishing denominator, singular and near-singular
matrices, saturated short coordinates, images narrower / lower than OpenCV's evaluation block, a crop of a shipped
MiddEval3 image under a rectification-like homography.
Run: python tests/golden/make_golden_warp_linear.py   (writes tests/golden/warplin_*.npz)"""
import os

import cv2
import numpy as np

OUT = os.path.dirname(os.path.abspath(__file__))
rng = np.random.default_rng(20261)
I = np.eye(3)
cv2.setNumThreads(1)


def case(name, M, src, dw, dh):
    M = np.asarray(M, np.float64)
    dst = cv2.warpPerspective(src, M, (dw, dh))
    np.savez_compressed(os.path.join(OUT, f"warplin_{name}.npz"), kind="warplin", src=src, M=M, dw=dw, dh=dh, out=dst)
    print(name, dst.shape, int((dst != 0).any(-1).sum()), "non-border pixels")


def noise(h, w):
    return rng.integers(0, 256, (h, w, 3)).astype(np.uint8)


# source point (x + y/32, y + x/32): the fractions are (y mod 32, x mod 32) -- all 32 x 32 table entries
case("all_fractions", np.linalg.inv(np.array([[1, 1 / 32, 0], [1 / 32, 1, 0], [0, 0, 1.0]])), noise(80, 80), 70, 70)
case("identity", I, noise(33, 47), 47, 33)
case("half_pixel_shift", [[1, 0, 0.5], [0, 1, -0.5], [0, 0, 1]], noise(40, 50), 60, 50)
case("borders_all_sides", [[0.8, 0.1, 10.3], [-0.05, 0.9, 7.7], [0, 0, 1]], noise(50, 60), 90, 80)
case("perspective", [[1.2, 0.1, -3.3], [0.05, 0.9, 4.1], [1e-4, -2e-4, 1.0]], noise(150, 200), 220, 160)
case("strong_perspective", [[1.0, 0.2, 5.0], [0.1, 1.1, -4.0], [4e-3, 3e-3, 1.0]], noise(120, 160), 200, 150)
case("vanishing_denominator", [[1.0, 0.0, 0.0], [0.0, 1.0, 0.0], [-1.0 / 64, 0.0, 1.0]], noise(60, 90), 130, 40)
case("singular", [[1, 2, 3], [2, 4, 6], [0, 0, 1]], noise(30, 30), 40, 30)
case("near_singular", [[1, 1, 0], [1, 1 + 1e-9, 0], [0, 0, 1]], noise(30, 30), 40, 30)
case("huge_scale", [[1e-6, 0, 0], [0, 1e-6, 0], [0, 0, 1]], noise(20, 20), 30, 20)          # saturated short coordinates
case("rotation", [[np.cos(0.3), -np.sin(0.3), 20], [np.sin(0.3), np.cos(0.3), -10], [0, 0, 1]], noise(90, 120), 140, 110)
case("wide_low", [[1.01, 0.002, -2.5], [0.001, 0.99, 1.5], [1e-5, 0, 1]], noise(9, 1500), 1400, 7)     # one block row, many block columns
case("narrow_tall", [[0.97, 0.01, 1.5], [0.02, 1.03, -2.5], [0, 2e-5, 1]], noise(300, 21), 19, 280)
# a crop of a shipped image under a homography of the size rectification produces
try:
    import sys
    sys.path.insert(0, os.path.dirname(OUT))
    from refdata import load_pair
    L, _ = load_pair("middeval", "Teddy")
    crop = np.ascontiguousarray(L[40:200, 100:340])
    case("teddy_crop_rectify", [[0.98, -0.04, 12.0], [0.03, 1.01, -6.0], [-6e-5, 2e-5, 1.0]], crop, 260, 170)
except Exception as e:   # fixtures missing: the synthetic cases stand
    print("no shipped image:", e)
