"""More golden vectors for the nearest-neighbour inverse warp of the disparity map (rectification.cpp:82-87:
cv::warpPerspective(disp_rect, out, H_.inv(), size, INTER_NEAREST) on CV_64F), produced by the real
cv2.warpPerspective in the build container: half-pixel landings, strong perspective (one landing point is exactly
k + 0.5 and OpenCV's block-wise evaluation order decides it), near-singular and singular matrices, rotation, a
rectification-sized map, maps narrower than OpenCV's 64-column block and lower than its 16-row block.
Run: python tests/golden/make_golden_warp.py   (writes tests/golden/warp_*.npz; sources are small integer maps)"""
import os

import cv2
import numpy as np

OUT = os.path.dirname(os.path.abspath(__file__))
rng = np.random.default_rng(20260)
I = np.eye(3)


def case(name, M, sw, sh, dw, dh):
    src = rng.integers(1, 250, (sh, sw)).astype(np.uint8)
    M = np.asarray(M, np.float64)
    dst = cv2.warpPerspective(src.astype(np.float64), M, (dw, dh), flags=cv2.INTER_NEAREST)
    assert np.array_equal(dst, dst.astype(np.uint8))
    np.savez_compressed(os.path.join(OUT, f"warp_{name}.npz"), kind="warp", src=src, M=M, dw=dw, dh=dh, out=dst.astype(np.uint8))
    print(name, dst.shape, int((dst != 0).sum()), "non-border pixels")


T = I.copy(); T[0, 2] = 0.5; T[1, 2] = -0.5
case("half_pixel_shift", T, 70, 50, 80, 60)
T = I.copy(); T[0, 2] = 2.5; T[1, 2] = 1.5; T[0, 0] = 0.5; T[1, 1] = 0.5
case("scale_half", T, 90, 70, 80, 60)
case("mild_perspective", [[1.2, 0.1, -3.3], [0.05, 0.9, 4.1], [1e-4, -2e-4, 1.0]], 200, 150, 220, 160)
case("strong_perspective", [[1.0, 0.2, 5], [0.1, 1.1, -7], [2e-3, 1.5e-3, 1.0]], 300, 200, 320, 240)
case("near_singular", [[1.0, 2.0, 3], [2.0, 4.0000001, 6.1], [0.001, 0.002, 1.0]], 100, 100, 120, 90)
case("singular", [[1.0, 2.0, 3], [2.0, 4.0, 6], [0.0, 0.0, 1.0]], 40, 30, 50, 40)
case("rotation", [[0.8660254, -0.5, 40], [0.5, 0.8660254, -20], [0, 0, 1.0]], 320, 240, 320, 240)
case("rectification_teddy", [[1.0021, 0.0153, -12.7], [-0.0112, 0.9987, 8.3], [1.3e-5, -2.1e-6, 1.0]], 899, 750, 900, 750)
case("zoom2", [[2.0, 0, 0.0], [0, 2.0, 0.0], [0, 0, 1.0]], 64, 64, 128, 128)
case("zoom3_shift", [[3.0, 0, 1.5], [0, 3.0, 1.5], [0, 0, 1.0]], 64, 64, 190, 190)
case("narrow_block", [[1.01, 0.02, -1.5], [0.03, 0.97, 0.5], [3e-4, 1e-4, 1.0]], 50, 40, 45, 38)
case("low_block", [[0.99, 0.01, 2.5], [0.0, 1.0, 0.5], [1e-4, 2e-4, 1.0]], 300, 12, 290, 10)
