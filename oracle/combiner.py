"""CPU oracle for the combiner's aggregate (elementwise mean over child outputs) and the standardise transformer.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).

* ``mean_combiner`` restates ``examples/combiners/mnist_combiner/MnistCombiner.py:8-11``
  (two children, ``(Xs[0]+Xs[1])/2.0``) generalised to N children the way the engine's
  built-in unit does it: ``engine/src/main/java/io/seldon/engine/predictors/
  AverageCombinerUnit.java:35-79`` (shape checks, sum in child order, divide by N) with
  ``PredictorUtils.add`` L67-95.  The Java unit accumulates in ojalgo 47.0.0
  ``PrimitiveMatrix`` (fp64; ``engine/pom.xml:175-177``, not vendored, no JVM here),
  so the accumulation is restated in fp64 numpy and pinned by the known answers of
  ``engine/src/test/java/io/seldon/engine/predictors/AverageCombinerTest.java:41-225``
  (``tests/test_oracle_combiner.py``) and by running ``MnistCombiner`` itself.
* ``standardise`` restates ``components/outlier-detection/vae/CoreVAE.py:128``:
  ``(X - mu) / (sigma + 1e-10)``.
"""
import numpy as np


class CombinerError(ValueError):
    """Stands for the engine's APIException(ENGINE_INVALID_COMBINER_RESPONSE)."""


def mean_combiner(Xs):
    if len(Xs) == 0:
        raise CombinerError("Combiner received no inputs")  # AverageCombinerUnit.java:37-39
    first = np.asarray(Xs[0])
    if first.ndim != 2:
        raise CombinerError("Combiner received data that is not 2 dimensional")  # L47-49
    acc = np.zeros(first.shape, dtype=np.float64)
    for X in Xs:
        X = np.asarray(X)
        if X.ndim != 2:
            raise CombinerError("Combiner received data that is not 2 dimensional")
        if X.shape[0] != first.shape[0]:
            raise CombinerError("Expected batch length %d but found %d" % (first.shape[0], X.shape[0]))  # L63-65
        if X.shape[1] != first.shape[1]:
            raise CombinerError("Expected batch length %d but found %d" % (first.shape[1], X.shape[1]))  # L66-68
        acc += X  # PredictorUtils.add, child order
    acc /= len(Xs)  # L71
    return acc


def standardise(X, mu, sigma):
    return (np.asarray(X, dtype=np.float64) - mu) / (np.asarray(sigma, dtype=np.float64) + 1e-10)
