import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Make sure liblensrt.so and the oracle exist (nvcc cross-compiles without a GPU)."""
    import __graft_entry__ as g
    if not os.path.exists(g.LIB):
        g.build()
    yield


@pytest.fixture(scope="session")
def mi():
    import lensrt
    return lensrt


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    d = os.path.join(ROOT, "tests", "golden")
    return {k: np.load(os.path.join(d, k + ".npz")) for k in ("rng", "maths", "lens")}
