"""Import the UNMODIFIED reference components from /root/reference.

Only usable in the build container (the reference tree does not travel to the
GPU box).  Used by ``oracle/make_golden.py`` to generate ``tests/golden/*.npz``
and by the ``not gpu`` tests that are skipped when the tree is absent.

The one adaptation: ``CoreMahalanobis.py:3`` imports ``scipy.linalg.eigh`` and
calls it with the ``eigvals=(lo, hi)`` keyword (``CoreMahalanobis.py:103``) that
SciPy >= 1.9 removed.  We rebind the *module global* ``eigh`` to a forwarder that
maps it to ``subset_by_index``; the reference source is not touched.
"""
import importlib.util
import os
import sys

REFERENCE_ROOT = os.environ.get("SB200_REFERENCE_ROOT", "/root/reference")

_PATHS = {
    "CoreMahalanobis": "components/outlier-detection/mahalanobis/CoreMahalanobis.py",
    "EpsilonGreedy": "components/routers/epsilon-greedy/EpsilonGreedy.py",
    "ThompsonSampling": "components/routers/thompson-sampling/ThompsonSampling.py",
    "MnistCombiner": "examples/combiners/mnist_combiner/MnistCombiner.py",
}


def available() -> bool:
    return all(os.path.isfile(os.path.join(REFERENCE_ROOT, p)) for p in _PATHS.values())


def _load(name):
    path = os.path.join(REFERENCE_ROOT, _PATHS[name])
    spec = importlib.util.spec_from_file_location("sb200_ref_" + name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[spec.name] = mod
    spec.loader.exec_module(mod)
    return mod


def core_mahalanobis_class():
    import scipy.linalg

    mod = _load("CoreMahalanobis")

    def eigh_compat(a, eigvals=None, **kw):
        if eigvals is not None:
            kw["subset_by_index"] = [int(eigvals[0]), int(eigvals[1])]
        return scipy.linalg.eigh(a, **kw)

    mod.eigh = eigh_compat
    return mod.CoreMahalanobis


def epsilon_greedy_class():
    return _load("EpsilonGreedy").EpsilonGreedy


def thompson_sampling_class():
    return _load("ThompsonSampling").ThompsonSampling


def mnist_combiner_class():
    return _load("MnistCombiner").MnistCombiner
