"""Host-side neighbours of the path: the Middlebury PFM reader and the calib.txt parser of the
C-ABI (no GPU needed), held against independent numpy / text parsing of the shipped fixtures."""
import os
import re

import numpy as np
import pytest

DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")
MID = os.path.join(DATA, "MiddEval3", "trainingH")

# scene -> (W, H, ndisp); index = the reference's scenarioIdx (alphabetical, data_loader.cpp:10)
SCENES = [("Adirondack", 1436, 992, 145), ("ArtL", 694, 554, 128), ("Jadeplant", 1318, 994, 320), ("Motorcycle", 1482, 994, 140),
          ("MotorcycleE", 1482, 994, 140), ("Piano", 1414, 962, 130), ("PianoL", 1414, 962, 130), ("Pipes", 1470, 970, 150),
          ("Playroom", 1398, 952, 165), ("Playtable", 1360, 926, 145), ("PlaytableP", 1362, 924, 145), ("Recycle", 1440, 972, 130),
          ("Shelves", 1476, 994, 120), ("Teddy", 900, 750, 128), ("Vintage", 1444, 960, 380)]


@pytest.fixture(scope="module")
def dl():
    from stereo_reconstruction_b200 import data_loader
    return data_loader


def _numpy_pfm(path):
    with open(path, "rb") as f:
        assert f.readline().strip() == b"Pf"
        w, h = map(int, f.readline().split())
        scale = float(f.readline())
        data = np.frombuffer(f.read(), "<f4" if scale < 0 else ">f4").reshape(h, w)
    return np.ascontiguousarray(data[::-1]).astype(np.float32)


@pytest.mark.parametrize("scene,name", [("ArtL", "disp0GT"), ("ArtL", "disp1GT"), ("Teddy", "disp0GT"), ("Teddy", "disp1GT")])
def test_read_pfm_shipped_ground_truth(dl, scene, name):
    p = os.path.join(MID, scene, name + ".pfm")
    got = dl.readPFM(p)
    want = _numpy_pfm(p)
    assert got.shape == want.shape
    np.testing.assert_array_equal(got.view(np.uint32), want.view(np.uint32))   # bitwise, incl. +inf
    assert np.isinf(got).any() and got[np.isfinite(got)].max() < 130          # scale NOT applied: values are pixels of disparity


def test_pfm_round_trip_and_big_endian(dl, tmp_path):
    from stereo_reconstruction_b200 import reconstruction
    rng = np.random.default_rng(0)
    a = rng.standard_normal((7, 11)).astype(np.float32)
    a[2, 3] = np.inf
    p = str(tmp_path / "a.pfm")
    reconstruction.writePFM(p, a)
    np.testing.assert_array_equal(dl.readPFM(p), a)
    q = str(tmp_path / "b.pfm")
    with open(q, "wb") as f:
        f.write(b"Pf\n11 7\n1.0\n")
        f.write(a[::-1].astype(">f4").tobytes())
    np.testing.assert_array_equal(dl.readPFM(q), a)
    with pytest.raises(RuntimeError):
        dl.readPFM(str(tmp_path / "missing.pfm"))


@pytest.mark.parametrize("idx", range(15))
def test_calib_and_scenario_table(dl, idx):
    scene, w, h, ndisp = SCENES[idx]
    loader = dl.DataLoader(DATA)
    assert loader.getTrainingScenarioNames()[idx] == scene
    d = loader.loadTrainingScenario(idx)
    assert d.imageLeft.shape == (h, w, 3) and d.imageRight.shape == (h, w, 3) and d.imageLeft.dtype == np.uint8
    assert d.ndisp == ndisp and d.calib["width"] == w and d.calib["height"] == h
    txt = open(os.path.join(MID, scene, "calib.txt")).read()
    for key, cam in (("cam0", d.cameraLeft), ("cam1", d.cameraRight)):
        nums = re.search(key + r"=\[(.*?)\]", txt).group(1).replace(";", " ").split()
        np.testing.assert_array_equal(cam, np.array([np.float32(v) for v in nums]).reshape(3, 3))
    assert d.has_ground_truth() == (scene in ("ArtL", "Teddy"))


def test_scenario_index_out_of_range(dl):
    with pytest.raises(RuntimeError):
        dl.DataLoader(DATA).loadTrainingScenario(15)
