import os
import sys

import numpy as np
import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "slow: long-running CPU test")


def golden(name):
    return dict(np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False))


@pytest.fixture(scope="session")
def sr():
    """The product package (import fails loudly if the CUDA library cannot be loaded)."""
    import sr459
    sr459._lib.load()
    return sr459


# tolerances of BASELINE.json's north_star, on the [0,255] scale the reference works in
TOL_ABS_255 = 1e-4 * 255.0        # max-abs <= 1e-4 on [0,1] images
TOL_PSNR_DB = 0.01
TOL_MSE_REL = 1e-4
