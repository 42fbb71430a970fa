"""Helpers shared by the test modules (tests/ is on sys.path under pytest's default import mode)."""
import numpy as np


def relerr(a, b):
    """max |a-b| / max(|b|, tiny) over the array (matrix-level relative error)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    if a.size == 0 and b.size == 0:
        return 0.0
    return float(np.max(np.abs(a - b)) / max(float(np.max(np.abs(b))), 1e-300))
