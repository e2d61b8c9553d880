"""Minimal host-side mirror of the reference's DataLoader (src/DataLoader/data_loader.cpp) -- just what feeds the hot
path in BASELINE.json configs 1-2: scene enumeration (alphabetical = the reference's scenarioIdx, data_loader.cpp:5-30),
im0/im1 (cv::imread IMREAD_COLOR -> BGR u8, :64-90), the calib.txt parser (:141-164) and the ground-truth PFMs with the
Middlebury SDK's layout (:110-125).  PNG decoding is OpenCV's on both sides (cv2 here, cv::imread there); the PFM
reader and the calib parser are the C-ABI's own (sr_read_pfm, sr_read_calib)."""
import ctypes as C
import os

import numpy as np

from . import capi


class SrCalib(C.Structure):
    _fields_ = [("doffs", C.c_double), ("baseline", C.c_double), ("dyavg", C.c_double), ("dymax", C.c_double),
                ("width", C.c_int), ("height", C.c_int), ("ndisp", C.c_int), ("isint", C.c_int), ("vmin", C.c_int), ("vmax", C.c_int)]


def readPFM(path):
    """Middlebury PFM -> float32 (h, w), rows top to bottom, +inf = unknown, scale magnitude not applied."""
    l = capi.lib()
    w, h = C.c_int(0), C.c_int(0)
    rc = l.sr_pfm_info(os.fsencode(path), C.byref(w), C.byref(h))
    if rc != capi.OK:
        raise capi.SrError(rc, f"cannot read PFM header of {path}")
    out = np.empty((h.value, w.value), np.float32)
    rc = l.sr_read_pfm(os.fsencode(path), C.c_void_p(out.ctypes.data), w.value, h.value)
    if rc != capi.OK:
        raise capi.SrError(rc, f"cannot read {path}")
    return out


def readCameraMatrices(scenarioPath):
    """DataLoader::readCameraMatrices: (cameraLeft, cameraRight) as 3x3 float32 + the calib.txt scalars as a dict."""
    cam0 = (C.c_float * 9)(); cam1 = (C.c_float * 9)()
    cal = SrCalib()
    rc = capi.lib().sr_read_calib(os.fsencode(os.path.join(scenarioPath, "calib.txt")), cam0, cam1, C.byref(cal))
    if rc != capi.OK:
        raise capi.SrError(rc, f"cannot parse {scenarioPath}/calib.txt")
    return (np.array(list(cam0), np.float32).reshape(3, 3), np.array(list(cam1), np.float32).reshape(3, 3),
            {k: getattr(cal, k) for k, _ in SrCalib._fields_})


class Data:
    """data.h:10-60: the images, the intrinsics and (where shipped) ground truth + masks of one scenario."""

    def __init__(self, name, imageLeft, imageRight, cameraLeft, cameraRight, calib, path):
        self.name, self.path = name, path
        self.imageLeft, self.imageRight = imageLeft, imageRight
        self.cameraLeft, self.cameraRight = cameraLeft, cameraRight
        self.calib = calib
        self.ndisp = calib["ndisp"]

    def has_ground_truth(self):
        return os.path.exists(os.path.join(self.path, "disp1GT.pfm"))

    def loadDisparityMatrices(self):
        """DataLoader::loadDisparityMatrices (:110-125) -> (dispLeft, dispRight)"""
        return readPFM(os.path.join(self.path, "disp0GT.pfm")), readPFM(os.path.join(self.path, "disp1GT.pfm"))

    def loadMasks(self):
        import cv2
        return (cv2.imread(os.path.join(self.path, "mask0nocc.png"), cv2.IMREAD_GRAYSCALE),
                cv2.imread(os.path.join(self.path, "mask1nocc.png"), cv2.IMREAD_GRAYSCALE))


class DataLoader:
    """DataLoader(root): root holds MiddEval3/trainingH/<scene>/ (the reference scans ../../data relative to its CWD)."""

    def __init__(self, root):
        self.root = os.path.join(root, "MiddEval3", "trainingH")
        self.trainingScenarioPaths = sorted(os.path.join(self.root, d) for d in os.listdir(self.root)
                                            if os.path.isdir(os.path.join(self.root, d)))

    def getTrainingScenarioNames(self):
        return [os.path.basename(p) for p in self.trainingScenarioPaths]

    def loadTrainingScenario(self, scenarioIdx):
        if scenarioIdx < 0 or scenarioIdx >= len(self.trainingScenarioPaths):
            raise RuntimeError("Select scenario index in range of [0; 14]")   # data_loader.cpp:66
        import cv2
        p = self.trainingScenarioPaths[scenarioIdx]
        L = cv2.imread(os.path.join(p, "im0.png"), cv2.IMREAD_COLOR)
        R = cv2.imread(os.path.join(p, "im1.png"), cv2.IMREAD_COLOR)
        k0, k1, cal = readCameraMatrices(p)
        return Data(os.path.basename(p), L, R, k0, k1, cal, p)


def readGrayscaleImageAsDisparityMap(path):
    """data_loader.cpp:191-203: 8-bit PNG -> CV_32FC1."""
    import cv2
    return cv2.imread(path, cv2.IMREAD_GRAYSCALE).astype(np.float32)
