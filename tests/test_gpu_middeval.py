"""BASELINE.json configs 1-2 on the REAL data at full size (north_star's first target: bit-exact SSD/SAD maps on all
MiddEval3 training pairs).  Every comparison is over the WHOLE map:
  * the CUDA path (through the C-ABI) against the CPU oracle on all 15 shipped pairs with main.cpp:40's parameters
    (block 17, minD 0, maxD = calib ndisp, smoothFactor 0.9 and 1.0), the reference's SSD-L2 cost and the SAD extension,
  * against the maps the compiled reference itself produced (tests/golden/ref_full, oracle/_ref),
  * LR check + convertDisparityToDepth(fx of cam1, 1) + vertices + WriteMesh face list per pair,
  * the reference's own rectified inputs (results/Rectified/trainingH) at maxD 200 incl. the unequal-size pairs,
  * Teddy against the shipped block-search result (soft golden, >= 90 % exact on the interior),
  * bad-2.0 / avgErr / invalid % on ArtL and Teddy (the only scenes with shipped ground truth) through sr_evaldisp.
SAD has no reference counterpart (parity unpinned by construction): its evidence is the oracle + its numpy witness."""
import json
import os

import numpy as np
import pytest

from refdata import DATA, MIDDEVAL_SCENES, RECT, RECTIFIED_SCENES, load_pair, ndisp, ref_full

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def sr():
    import stereo_reconstruction_b200 as s
    return s


@pytest.fixture(scope="module")
def ctx(sr):
    return sr.capi.default_context()


def _same(got, want, what):
    bad = int((got != want).sum())
    print(f"{what}: {bad} mismatching pixels of {got.size}")
    assert bad == 0, f"{what}: {bad} pixels differ"


@pytest.mark.parametrize("scene", MIDDEVAL_SCENES)
def test_middeval_right_ssd_whole_map(sr, ctx, orc, scene):
    """computeDisparityMapRight(17, 0, ndisp, s) as main.cpp:40 calls it, s = 0.9 (default) and 1.0."""
    L, R = load_pair("middeval", scene)
    D = ndisp(scene)
    for smooth in (0.9, 1.0):
        got = sr.BlockSearch(L, R, 17, 0, D, ctx=ctx).computeDisparityMapRight(smooth)
        _same(got, orc.block_search_right_fast(L, R, 17, 0, D, smooth), f"{scene} SSD s={smooth} vs oracle")
        want = ref_full("middeval", scene, D, smooth)
        assert want is not None, "tests/golden/ref_full is incomplete"
        _same(got, want, f"{scene} SSD s={smooth} vs the compiled reference")


@pytest.mark.parametrize("scene", MIDDEVAL_SCENES)
def test_middeval_right_sad_whole_map(sr, ctx, orc, scene):
    L, R = load_pair("middeval", scene)
    D = ndisp(scene)
    for block in (17, 9):
        got = sr.BlockSearch(L, R, block, 0, D, cost=sr.capi.COST_SAD, ctx=ctx).computeDisparityMapRight(1.0)
        _same(got, orc.block_search_right_fast(L, R, block, 0, D, 1.0, cost=orc.COST_SAD), f"{scene} SAD block {block} vs oracle")


@pytest.mark.parametrize("scene", MIDDEVAL_SCENES)
def test_middeval_lr_check_depth_vertices_faces(sr, ctx, orc, scene):
    """config 2: default window, left-right check, disparity -> depth -> point cloud -> face list, per pair."""
    capi = sr.capi
    d = sr.data_loader.DataLoader(DATA).loadTrainingScenario(MIDDEVAL_SCENES.index(scene))
    L, R, D = d.imageLeft, d.imageRight, d.ndisp
    h, w = R.shape[:2]
    pair = capi.Pair(ctx, w, h, w, h)
    pair.upload(L, R)
    p = capi.search_params(block=17, min_disp=0, max_disp=D, smooth=1.0, lr_check=True)
    pair.run(p)
    got = pair.download(p)
    dr, _, _ = orc.search_ext_fast(L, R, 17, 0, D, orc.COST_SSD_L2)
    dl, _, _ = orc.search_ext_fast(L, R, 17, 0, D, orc.COST_SSD_L2, left_ref=True)
    _same(pair.left_disparity(), dl, f"{scene} left-referenced map")
    want = orc.lr_check(dr, dl)
    _same(got, want.astype(np.float64), f"{scene} LR-checked map")
    K = d.cameraRight
    focal = float(K[0, 0])
    rec = sr.reconstruction.pairReconstruct(pair, K, focal, 1.0, quantize_u8=False, kernelSize=0, thrMesh=1.0)
    disp32 = want.astype(np.float32)
    depth = orc.disparity_to_depth(disp32, focal, 1.0)                       # main.cpp:56-58
    np.testing.assert_array_equal(rec["depth"], depth)
    verts = orc.backproject(R, depth, float(K[0, 0]), float(K[1, 1]), float(K[0, 2]), float(K[1, 2]))
    assert rec["vertices"].tobytes() == verts.tobytes()                       # 32-byte struct Vertex, byte for byte
    faces = orc.mesh_faces(verts, w, h, 1.0)
    np.testing.assert_array_equal(rec["faces"], faces)
    print(f"{scene}: {int((want != 0).sum())} consistent pixels, {len(faces)} faces")
    pair.close()


@pytest.mark.parametrize("scene", RECTIFIED_SCENES)
def test_rectified_pairs_main_cpp_call(sr, ctx, orc, scene):
    """rectifier.computeDisparityMapRight(17, 0, 200, 0.9) (main.cpp:40) on the reference's own rectified images,
    i.e. the exact inputs BlockSearch sees on main.cpp's path; MotorcycleE (left 1038 rows < right 1052) makes the
    reference throw cv::Exception from the left ROI -> SR_E_SHAPE here."""
    L, R = load_pair("rectified", scene)
    if L.shape[0] < R.shape[0]:
        with pytest.raises(sr.capi.SrError) as e:
            sr.BlockSearch(L, R, 17, 0, 200, ctx=ctx).computeDisparityMapRight(0.9)
        assert e.value.code == sr.capi.E_SHAPE
        with pytest.raises(orc.OracleError):
            orc.block_search_right_fast(L, R, 17, 0, 200, 0.9)
        return
    got = sr.BlockSearch(L, R, 17, 0, 200, ctx=ctx).computeDisparityMapRight(0.9)
    _same(got, orc.block_search_right_fast(L, R, 17, 0, 200, 0.9), f"rectified {scene} vs oracle")
    want = ref_full("rectified", scene, 200, 0.9)
    assert want is not None
    _same(got, want, f"rectified {scene} vs the compiled reference")


def test_teddy_soft_golden(sr, ctx):
    """results/PerceptualWindowSearch/test_result.png is a block-search result for Teddy after the inverse warp and the
    8-bit hand-off: >= 90 % of the interior pixels agree exactly with the search on the rectified pair."""
    import cv2
    L, R = load_pair("rectified", "Teddy")
    got = sr.BlockSearch(L, R, 17, 0, 200, ctx=ctx).computeDisparityMapRight(1.0)
    soft = cv2.imread(os.path.join(DATA, "PerceptualWindowSearch", "test_result.png"), cv2.IMREAD_GRAYSCALE)
    h, w = min(got.shape[0], soft.shape[0]), min(got.shape[1], soft.shape[1])
    a, b = got[20:h - 20, 20:w - 220], soft[20:h - 20, 20:w - 220].astype(np.float64)
    exact, near = float((a == b).mean()), float((np.abs(a - b) <= 1).mean())
    print(f"Teddy vs test_result.png: {100 * exact:.1f} % exact, {100 * near:.1f} % within 1")
    assert exact >= 0.90


@pytest.mark.parametrize("scene", ["ArtL", "Teddy"])
def test_bad2_on_ground_truth_scenes(sr, ctx, orc, scene):
    """evaldisp (utils.cpp:123-168) of the right-referenced map against disp1GT.pfm / mask1nocc.png: counts identical
    to the oracle's, percentages and avgErr equal; the numbers go to BASELINE.md (gpurun_out/bad2.jsonl)."""
    d = sr.data_loader.DataLoader(DATA).loadTrainingScenario(15 - 2 if scene == "Teddy" else 1)
    assert d.name == scene
    _, gt = d.loadDisparityMatrices()
    _, mask = d.loadMasks()
    rows = []
    for smooth, cost, tag in ((0.9, sr.capi.COST_SSD_L2, "ssd_s0.9"), (1.0, sr.capi.COST_SSD_L2, "ssd_s1.0"), (1.0, sr.capi.COST_SAD, "sad_s1.0")):
        disp = sr.BlockSearch(d.imageLeft, d.imageRight, 17, 0, d.ndisp, cost=cost, ctx=ctx).computeDisparityMapRight(smooth).astype(np.float32)
        for thr in (2.0, 1.0):
            got = sr.reconstruction.evaldisp(disp, gt, mask, thr, float(d.ndisp), 0, ctx=ctx)
            want = orc.evaldisp(disp, gt, mask, thr, float(d.ndisp), 0)
            assert (got["n"], got["bad"], got["invalid"]) == (want.n, want.bad, want.invalid)
            assert got["bad_pct"] == pytest.approx(want.bad_pct, rel=1e-6) and got["avg_err"] == pytest.approx(want.avg_err, rel=1e-4)
            rows.append({"scene": scene, "search": tag, "badthresh": thr, "evaluated": got["n"], "bad_pct": round(got["bad_pct"], 3),
                         "invalid_pct": round(got["invalid_pct"], 3), "total_bad_pct": round(got["total_bad_pct"], 3), "avg_err": round(got["avg_err"], 4)})
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    os.makedirs(out, exist_ok=True)
    with open(os.path.join(out, f"bad2_{scene}.jsonl"), "w") as f:
        for r in rows:
            f.write(json.dumps(r) + "\n")
            print(r)
