import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as o
    o.lib()
    return o


@pytest.fixture(scope="session")
def default_volume(oracle):
    """The reference's start-up bake (TestMain.cpp:43-92) by the oracle's restatement, 128^3 RGBA8.
    Cached on disk: it is ~10 s of scalar CPU work."""
    path = "/tmp/volrt_default_volume_v1.npy"
    if os.path.exists(path):
        v = np.load(path)
        if v.shape == (128, 128, 128, 4):
            return v
    v = oracle.bake_default(128)
    try:
        np.save(path, v)
    except OSError:
        pass
    return v
