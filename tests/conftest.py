import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def golden_cases(kind):
    import numpy as np
    out = []
    for f in sorted(os.listdir(GOLDEN)):
        if f.endswith(".npz"):
            z = np.load(os.path.join(GOLDEN, f), allow_pickle=False)
            if str(z["kind"]) == kind:
                out.append((f[:-4], z))
    return out


@pytest.fixture(scope="session")
def orc():
    from oracle import orc as o
    o.lib()
    return o
