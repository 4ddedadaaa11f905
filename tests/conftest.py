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
    """The reference's start-up bake (TestMain.cpp:43-92), 128^3 RGBA8.  Where oracle/_ref travelled along (the build
    container and the GPU box) these are the bytes of the REFERENCE'S OWN FastNoise2 at the CPU's best SIMD level — what
    the app itself would upload; elsewhere the oracle's restatement (~10 s of scalar work, cached on disk under a key
    that changes whenever the oracle library does)."""
    if oracle.have_ref():
        return oracle.ref_bake_default(128)[0]
    import hashlib
    so = os.path.join(ROOT, "oracle", "libvolrt_oracle.so")
    key = hashlib.sha256(open(so, "rb").read()).hexdigest()[:16]
    path = "/tmp/volrt_default_volume_%s.npy" % key
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
