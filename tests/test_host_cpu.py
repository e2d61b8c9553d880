"""CPU-only checks of the host logic: synthetic generator vs the oracle's, evaluation counting,
the product package never importing the oracle."""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_synth_matches_oracle_generator(orc):
    from stereo_reconstruction_b200 import synth
    for (w, h, a, b, q, p) in [(200, 70, 0, 40, False, 0), (333, 45, 3, 100, True, 7), (640, 33, 0, 512, False, 1023)]:
        L, R = synth.synth_pair(0xB200, p, w, h, a, b, q)
        L2, R2 = orc.synth_pair(0xB200, p, w, h, a, b, q)
        assert np.array_equal(L, L2) and np.array_equal(R, R2)
        assert R.min() >= 16 and (L.reshape(-1, 3).max(axis=1) > 0).all()      # never black


def test_effective_evals_formula(orc):
    from stereo_reconstruction_b200 import synth
    w, h, D, B = 97, 20, 30, 9
    L, R = orc.synth_pair(1, 0, w, h, 0, D)
    # count candidates the literal oracle loop visits: d < min(D, w1 - x - r)
    half = (B - 1) // 2
    n = sum(max(0, min(D, w - x - min(w - x - 1, half))) for x in range(w)) * h
    assert synth.effective_evals(w, w, h, B, 0, D) == n


def test_product_never_imports_oracle():
    code = ("import sys; sys.path.insert(0, %r); import stereo_reconstruction_b200 as s; "
            "from stereo_reconstruction_b200 import reconstruction, sharding, synth, window_search; "
            "assert not any(m == 'oracle' or m.startswith('oracle.') for m in sys.modules), 'oracle imported'") % ROOT
    subprocess.check_call([sys.executable, "-c", code])
    for dirpath, _, files in os.walk(os.path.join(ROOT, "stereo_reconstruction_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "sr_oracle" not in txt and "import orc" not in txt and "from oracle" not in txt, f


def test_recon_roofline_summary_of_committed_launch_list():
    """tools/recon_roofline.py --summarise on the committed ncu launch list: every HBM-bound kernel of the 4K pipeline is
    there with its algorithmic bytes and a fraction of the HBM peak (no GPU needed: parsing only)."""
    import json
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "tools", "recon_roofline.py"), "--summarise",
                          os.path.join(root, "profiles", "r01_launches_recon_4k.csv"), "--peak", "6553"],
                         check=True, capture_output=True, text=True, env=dict(os.environ, SR_FACES="16026827", SR_RECON_U8="0")).stdout   # round-1 list: the general (float map) pipeline
    rows = {j["kernel"]: j for j in map(json.loads, out.splitlines())}
    n = 3840 * 2160
    assert rows["quantize_kernel"]["algorithmic_bytes"] == n * 12
    assert rows["depth_backproject_faces_kernel"]["algorithmic_bytes"] == n * 43 + 3839 * 2159
    for k in ("quantize_kernel", "blur_rows_fast_kernel", "blur_cols_seg_kernel", "blur_cols_prefix_kernel", "blur_cols_out_kernel",
              "outlier_replace_kernel", "depth_backproject_faces_kernel", "resolve_plain2_kernel", "face_scatter_kernel"):
        r = rows[k]
        assert r["us"] > 0 and r["dram_bytes"] > 0 and 0.1 < r["frac_of_hbm_peak"] < 1.5, (k, r)
        assert abs(r["algorithmic_GBps"] / (r["algorithmic_bytes"] / r["us"] / 1e3) - 1.0) < 1e-3     # us is printed rounded
