import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


PARITY_LOG = []       # filled by tests/test_gpu_parity.py::close_rel: norm-wise (asserted) and element-wise relative errors


def pytest_sessionfinish(session, exitstatus):
    if PARITY_LOG and os.path.isdir(os.path.join(ROOT, "gpurun_out")):
        import json
        with open(os.path.join(ROOT, "gpurun_out", "parity_relative_errors.json"), "w") as f:
            json.dump(PARITY_LOG, f, indent=1)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


class _RefGolden:
    """Inputs and weights of every tiny case from layers_tiny.npz, ``<case>/expected`` from ref_layers.npz -- the output of
    the REFERENCE'S OWN call() bodies (tests/golden/make_ref_golden.py), so the GPU parity tests that load a case compare
    with the reference itself, not with the restatement."""

    def __init__(self, arrays, ref):
        self._a, self._r = arrays, ref
        self.files = list(arrays.files)

    def __getitem__(self, key):
        if key.endswith("/expected"):
            return self._r[key]
        return self._a[key]


@pytest.fixture(scope="session")
def golden_oracle():
    """layers_tiny.npz as frozen: inputs, weights and the float64 ORACLE's outputs."""
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "layers_tiny.npz"))


@pytest.fixture(scope="session")
def ref_golden():
    """ref_layers.npz: outputs of the reference's unmodified layer / model code on the same cases."""
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "ref_layers.npz"))


@pytest.fixture(scope="session")
def golden(golden_oracle, ref_golden):
    return _RefGolden(golden_oracle, ref_golden)


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
