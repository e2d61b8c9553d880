"""Full-size golden maps from the reference ITSELF: oracle/_ref/libsr_ref.so (the reference's BlockSearch.cpp compiled
unchanged, see oracle/Makefile) run on the shipped pairs with main.cpp:40's parameters.  Run once in the build
container (~30 min on 8 cores); the outputs are committed as tests/golden/ref_full/*.npz (int16, compressed) and are
what the GPU parity tests of BASELINE.json configs 1-2 compare against, next to the oracle.

  MiddEval3 trainingH, 15 pairs: BlockSearch(im0, im1, 17, 0, ndisp).computeDisparityMapRight(0.9) and (1.0)
  results/Rectified/trainingH, 11 pairs (the reference's own rectified inputs; 2 of unequal size):
                                 BlockSearch(left, right, 17, 0, 200).computeDisparityMapRight(0.9)   [main.cpp:40]
"""
import multiprocessing as mp
import os
import sys
import time

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
DATA = os.path.join(ROOT, "tests", "data")
OUT = os.path.join(ROOT, "tests", "golden", "ref_full")


def ndisp_of(calib):
    for line in open(calib):
        if line.startswith("ndisp="):
            return int(line.split("=")[1])
    raise ValueError(calib)


def tasks():
    t = []
    mid = os.path.join(DATA, "MiddEval3", "trainingH")
    for scene in sorted(os.listdir(mid)):
        nd = ndisp_of(os.path.join(mid, scene, "calib.txt"))
        for s in (0.9, 1.0):
            t.append(("middeval", scene, os.path.join(mid, scene, "im0.png"), os.path.join(mid, scene, "im1.png"), nd, s))
    rec = os.path.join(DATA, "Rectified", "trainingH")
    for scene in sorted(os.listdir(rec)):
        t.append(("rectified", scene, os.path.join(rec, scene, "rectifiedLeft.png"), os.path.join(rec, scene, "rectifiedRight.png"), 200, 0.9))
    return t


def run(task):
    from oracle import ref
    kind, scene, lp, rp, max_d, smooth = task
    name = f"{kind}_{scene}_b17_d{max_d}_s{str(smooth).replace('.', '')}"
    dst = os.path.join(OUT, name + ".npz")
    if os.path.exists(dst):
        return name, 0.0
    L, R = cv2.imread(lp, cv2.IMREAD_COLOR), cv2.imread(rp, cv2.IMREAD_COLOR)
    t0 = time.time()
    try:
        out, mb = ref.block_search_right(L, R, 17, 0, max_d, smooth, return_max_block=True)
    except ref.RefError as e:   # Rectified/MotorcycleE: left image lower than the right one -> cv::Exception from the left ROI
        return name + f" -> the reference throws ({e.code})", time.time() - t0
    assert mb == 17 and np.array_equal(out, np.rint(out)) and np.abs(out).max() < 32768
    np.savez_compressed(dst, kind=f"ref_full_{kind}", scene=scene, block=17, min_d=0, max_d=max_d, smooth=smooth,
                        out=out.astype(np.int16), left_shape=L.shape[:2], right_shape=R.shape[:2])
    return name, time.time() - t0


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    ts = tasks()
    ts.sort(key=lambda t: -cv2.imread(t[3]).size * t[4])
    with mp.Pool(int(sys.argv[1]) if len(sys.argv) > 1 else os.cpu_count()) as pool:
        for name, dt in pool.imap_unordered(run, ts):
            print(f"{name}: {dt:.0f} s", flush=True)
