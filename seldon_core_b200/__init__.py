"""seldon_core_b200: the numeric data plane of the seldon_core Python wrapper, computed on NVIDIA B200.

Drop-in component classes (same constructor parameters and methods as the reference's):
  OutlierMahalanobisB200  - components/outlier-detection/mahalanobis/CoreMahalanobis.py
  EpsilonGreedyB200       - components/routers/epsilon-greedy/EpsilonGreedy.py
  ThompsonSamplingB200    - components/routers/thompson-sampling/ThompsonSampling.py
  MeanCombinerB200        - examples/combiners/mnist_combiner/MnistCombiner.py (+ engine AverageCombinerUnit)
  StandardiseB200         - components/outlier-detection/vae/CoreVAE.py:128
  RandomABTestB200        - engine/.../predictors/RandomABTestUnit.java (built-in RANDOM_ABTEST)
  GraphB200 / Unit        - engine/.../predictors/PredictiveUnitBean.java:113-197 (one device-resident graph pass)
plus `codec` (SeldonMessage tensor / ndarray payloads), `wire` (serialised SeldonMessage in -> out, the `*_raw` hooks)
and `dist` (row-sharded batches).
"""
from .errors import SeldonB200Error  # noqa: F401
from .outlier import OutlierMahalanobisB200  # noqa: F401
from .feedback import OutlierMahalanobisFeedbackB200  # noqa: F401
from .routers import EpsilonGreedyB200, ThompsonSamplingB200  # noqa: F401
from .combiner import MeanCombinerB200  # noqa: F401
from .transformer import StandardiseB200  # noqa: F401
from .graph import GraphB200, RandomABTestB200, Unit  # noqa: F401
from . import codec, dist, wire  # noqa: F401

__version__ = "0.1.0"
