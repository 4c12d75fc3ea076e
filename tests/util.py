import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pfb():
    return importlib.import_module("dune-parafields_b200")


def rel(a, b):
    """max |a-b| / max |b|; NaN / inf entries must coincide (the reference
    itself produces them, e.g. 1/lambda on clamped eigenvalues or a CG that
    iterates past exact convergence) and are left out of the norm."""
    import numpy as np

    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    fa, fb = np.isfinite(a), np.isfinite(b)
    if not np.array_equal(fa, fb):
        return float("inf")
    if not fb.any():
        return 0.0
    return float(np.max(np.abs(a[fa] - b[fb])) / max(np.max(np.abs(b[fb])), 1e-300))
