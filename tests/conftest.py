import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "layers_tiny.npz"))


@pytest.fixture(scope="session")
def criteo_cfg1():
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "criteo_cfg1.npz"))


@pytest.fixture(scope="session")
def lib():
    from recommender_system_b200 import build
    build.build_lib(verbose=False)
    from recommender_system_b200 import _lib
    return _lib.load()
