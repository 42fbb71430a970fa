"""pytest configuration: markers, repo root on sys.path, golden-vector loader."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    # the oracle calls scipy.spatial.distance_matrix because the reference does (ovehicle.py:94)
    config.addinivalue_line("filterwarnings", "ignore:.*is deprecated in favor of.*:DeprecationWarning")


@pytest.fixture(scope="session")
def golden():
    def load(name):
        return np.load(os.path.join(GOLDEN, name + ".npz"))
    return load


def relerr(a, b):
    """max |a-b| / max(|b|, tiny) over the array (matrix-level relative error)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b)) / max(float(np.max(np.abs(b))), 1e-300))
