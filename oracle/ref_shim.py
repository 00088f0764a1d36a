"""Import the UNMODIFIED reference components: from /root/reference in the build
container, else from the verbatim copy ``oracle/build.py`` left in the git-ignored
``oracle/_ref/`` (which travels to the GPU box with the snapshot).

Used by ``oracle/make_golden.py`` to generate ``tests/golden/*.npz``, by the
``--impl reference`` arm of ``bench.py`` (the CPU baseline) and by the tests that
drive the drop-in classes through the reference's own ``user_model.client_*``.

The one adaptation: ``CoreMahalanobis.py:3`` imports ``scipy.linalg.eigh`` and
calls it with the ``eigvals=(lo, hi)`` keyword (``CoreMahalanobis.py:103``) that
SciPy >= 1.9 removed.  We rebind the *module global* ``eigh`` to a forwarder that
maps it to ``subset_by_index``; the reference source is not touched.
"""
import importlib.util
import os
import sys

_COPY = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")
REFERENCE_ROOT = os.environ.get("SB200_REFERENCE_ROOT") or ("/root/reference" if os.path.isdir("/root/reference")
                                                            else _COPY)

_PATHS = {
    "CoreMahalanobis": "components/outlier-detection/mahalanobis/CoreMahalanobis.py",
    "EpsilonGreedy": "components/routers/epsilon-greedy/EpsilonGreedy.py",
    "ThompsonSampling": "components/routers/thompson-sampling/ThompsonSampling.py",
    "MnistCombiner": "examples/combiners/mnist_combiner/MnistCombiner.py",
    "user_model": "python/seldon_core/user_model.py",
    "metrics": "python/seldon_core/metrics.py",
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


def user_model_module():
    """The reference's ``seldon_core.user_model`` (the ``client_*`` dispatch the wrapper calls, user_model.py:81-361)
    with ``seldon_core.metrics`` (validate_metrics, metrics.py:65-89), both unmodified.  The two imports of user_model
    that cannot be satisfied here are stood in for: ``seldon_core.flask_utils`` (imports flask) by a module that only
    defines ``SeldonMicroserviceException`` with the reference's fields (flask_utils.py:67-81), and
    ``seldon_core.proto.prediction_pb2`` (imports tensorflow; used by user_model in type annotations only) by an empty
    module whose attributes are placeholder classes."""
    import types

    if "seldon_core.user_model" in sys.modules and getattr(sys.modules["seldon_core.user_model"], "_sb200_ref", False):
        return sys.modules["seldon_core.user_model"]
    pkg = types.ModuleType("seldon_core")
    pkg.__path__ = []
    fu = types.ModuleType("seldon_core.flask_utils")

    class SeldonMicroserviceException(Exception):
        status_code = 400

        def __init__(self, message, status_code=None, payload=None, reason="MICROSERVICE_BAD_DATA"):
            Exception.__init__(self)
            self.message = message
            if status_code is not None:
                self.status_code = status_code
            self.payload = payload
            self.reason = reason

        def to_dict(self):
            return {"status": {"status": 1, "info": self.message, "code": -1, "reason": self.reason}}

    fu.SeldonMicroserviceException = SeldonMicroserviceException
    proto = types.ModuleType("seldon_core.proto")
    proto.__path__ = []

    class _Pb2(types.ModuleType):
        def __getattr__(self, name):
            if name.startswith("__"):
                raise AttributeError(name)
            return type(name, (), {})

    pb2 = _Pb2("seldon_core.proto.prediction_pb2")
    proto.prediction_pb2 = pb2
    saved = {k: sys.modules.get(k) for k in ("seldon_core", "seldon_core.flask_utils", "seldon_core.proto",
                                             "seldon_core.proto.prediction_pb2", "seldon_core.metrics",
                                             "seldon_core.user_model")}
    sys.modules.update({"seldon_core": pkg, "seldon_core.flask_utils": fu, "seldon_core.proto": proto,
                        "seldon_core.proto.prediction_pb2": pb2})
    mods = {}
    for name in ("metrics", "user_model"):
        spec = importlib.util.spec_from_file_location("seldon_core." + name, os.path.join(REFERENCE_ROOT, _PATHS[name]))
        mod = importlib.util.module_from_spec(spec)
        sys.modules[spec.name] = mod
        spec.loader.exec_module(mod)
        setattr(pkg, name, mod)
        mods[name] = mod
    mods["user_model"]._sb200_ref = True
    mods["user_model"].metrics_module = mods["metrics"]
    return mods["user_model"]
