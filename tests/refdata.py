"""Shared access to the committed data fixtures (tests/data, copied from the reference's shipped data by
tests/data/fetch_fixtures.py) and to the full-size maps computed by the compiled reference (tests/golden/ref_full)."""
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
DATA = os.path.join(HERE, "data")
MID = os.path.join(DATA, "MiddEval3", "trainingH")
RECT = os.path.join(DATA, "Rectified", "trainingH")
REF_FULL = os.path.join(HERE, "golden", "ref_full")

MIDDEVAL_SCENES = sorted(os.listdir(MID)) if os.path.isdir(MID) else []
RECTIFIED_SCENES = sorted(os.listdir(RECT)) if os.path.isdir(RECT) else []


def ref_full_cases():
    return sorted(f[:-4] for f in os.listdir(REF_FULL) if f.endswith(".npz")) if os.path.isdir(REF_FULL) else []


def load_pair(kind, scene):
    """BGR u8 images exactly as cv::imread(IMREAD_COLOR) gives them to the reference."""
    import cv2
    if kind.endswith("middeval"):
        return (cv2.imread(os.path.join(MID, scene, "im0.png"), cv2.IMREAD_COLOR), cv2.imread(os.path.join(MID, scene, "im1.png"), cv2.IMREAD_COLOR))
    return (cv2.imread(os.path.join(RECT, scene, "rectifiedLeft.png"), cv2.IMREAD_COLOR),
            cv2.imread(os.path.join(RECT, scene, "rectifiedRight.png"), cv2.IMREAD_COLOR))


def ndisp(scene):
    for line in open(os.path.join(MID, scene, "calib.txt")):
        if line.startswith("ndisp="):
            return int(line.split("=")[1])
    raise ValueError(scene)


def ref_full(kind, scene, max_d, smooth):
    p = os.path.join(REF_FULL, f"{kind}_{scene}_b17_d{max_d}_s{str(smooth).replace('.', '')}.npz")
    return np.load(p)["out"].astype(np.float64) if os.path.exists(p) else None
