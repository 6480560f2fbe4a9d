import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "penr-oz-neural-network-v3-torch-ddp_b200")
for p in (PKG, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (sm_100a) device; run with -m gpu on the GPU box")


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    import json
    g = np.load(os.path.join(ROOT, "tests", "golden", "golden.npz"))
    meta = json.load(open(os.path.join(ROOT, "tests", "golden", "golden_meta.json")))
    return g, meta
