import numpy as np


def rel_err(a, ref):
    a = np.asarray(a, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    return np.abs(a - ref) / np.maximum(np.abs(ref), 1e-6)


def norm_err(A, ref):
    """max-abs error relative to the largest entry (covariances: elementwise relative error is ill-posed)."""
    ref = np.asarray(ref, dtype=np.float64)
    return float(np.max(np.abs(np.asarray(A) - ref)) / max(np.max(np.abs(ref)), 1e-300))


def projector_err(V, Vref):
    return float(np.linalg.norm(V @ V.T - Vref @ Vref.T))
