"""The engine's built-in units as device kernels and a device-resident pass over a small inference graph
(SURVEY 8(f) row 3).

The reference walks a graph in the Java engine, one REST/gRPC hop per node with the payload serialised at every hop
(engine/src/main/java/io/seldon/engine/predictors/PredictiveUnitBean.java:113-197).  `GraphB200` keeps the payload on
the GPU between the nodes and follows the same order of operations:

    transformInput -> (no children: done) -> route -> selected children (all when the unit does not route)
                   -> aggregate -> transformOutput,

with the meta handling of `mergeMeta` (L354-372): the tags of the earlier message(s) first, then the latest message's
on top, puid carried along, metrics collected per unit (`addMetrics`, L96-110) and cleared from the message, the chosen
branch recorded in `routing` (L159) and the unit in `requestPath` (L123).

Built-ins (PredictiveUnitImplementation): `AVERAGE_COMBINER` = MeanCombinerB200 (AverageCombinerUnit.java:35-79),
`RANDOM_ABTEST` = RandomABTestB200 below (RandomABTestUnit.java:29-59).
"""
import ctypes
import threading

import numpy as np

from . import _cabi
from .errors import SeldonB200Error


class RandomABTestB200(object):
    """`RANDOM_ABTEST`: branch 0 with probability ratioA, else branch 1, from `java.util.Random(1337).nextFloat()`
    - the decisions are the engine's, call for call."""

    def __init__(self, ratioA=None, seed=1337, device=None):
        if ratioA is None:
            raise SeldonB200Error("Parameter 'ratioA' is missing.")   # RandomABTestUnit.java:39-41
        self.ratioA = float(np.float32(ratioA))
        self.seed = int(seed)
        self._device = device
        self._state = None
        self._restore = None
        self._lock = threading.RLock()

    def _ensure(self):
        import torch

        if self._state is None:
            if not torch.cuda.is_available():
                raise SeldonB200Error("no CUDA device: seldon_core_b200 has no CPU fallback", status_code=500)
            if self._device is None:
                self._device = torch.device("cuda", torch.cuda.current_device())
            s0 = self._restore if self._restore is not None else int(_cabi.lib().sb_abtest_initial_state(self.seed))
            self._state = torch.tensor([s0, 0], dtype=torch.int64, device=self._device)
            self._restore = None
        return self._state

    def run_batched(self, T):
        """T consecutive route() calls in one launch; returns the decisions as a CUDA int32 tensor."""
        import torch

        with self._lock:
            st = self._ensure()
            dec = torch.empty(int(T), dtype=torch.int32, device=st.device)
            with torch.cuda.device(st.device):
                _cabi.check(_cabi.lib().sb_abtest_route(st.data_ptr(), ctypes.c_float(self.ratioA), int(T), dec.data_ptr(),
                                                        torch.cuda.current_stream().cuda_stream))
            return dec

    def route(self, features=None, feature_names=None):
        return int(self.run_batched(1).item())

    def __getstate__(self):
        with self._lock:
            s = self._restore if self._state is None else int(self._state[0].item())
            return {"ratioA": self.ratioA, "seed": self.seed, "_restore": s}

    def __setstate__(self, state):
        self.ratioA, self.seed, self._restore = state["ratioA"], state["seed"], state["_restore"]
        self._device = self._state = None
        self._lock = threading.RLock()


class Unit(object):
    """One node of the graph: `impl` is a component instance (any object with some of transform_input / route /
    aggregate / transform_output / predict taking and returning CUDA tensors, plus optional tags() / metrics()),
    `children` the nodes below it.  `kind` follows PredictiveUnitType: MODEL, ROUTER, COMBINER, TRANSFORMER,
    OUTPUT_TRANSFORMER (it selects which methods the engine calls, PredictiveUnitBean.java:219-292)."""

    def __init__(self, name, kind, impl, children=(), image=""):
        self.name, self.kind, self.impl, self.children, self.image = name, kind, impl, list(children), image


class GraphB200(object):
    def __init__(self, root):
        self.root = root

    def predict(self, X, puid=""):
        """One request batch (a CUDA float32 matrix) through the graph.  Returns (output tensor, meta) with
        meta = {puid, tags, routing, requestPath, metrics}."""
        routing, path, metrics = {}, {}, {}
        out, tags = self._walk(self.root, X, {}, routing, path, metrics)
        return out, {"puid": puid, "tags": tags, "routing": routing, "requestPath": path, "metrics": metrics}

    # ------------------------------------------------------------------ PredictiveUnitBean.getOutputAsync
    @staticmethod
    def _call_meta(impl, unit, metrics):
        """tags and metrics the unit's microservice would attach to its response (utils.py:431-436)."""
        tags = {}
        if hasattr(impl, "tags"):
            tags = impl.tags() or {}
        if hasattr(impl, "metrics"):
            m = impl.metrics()
            if m:
                metrics.setdefault(unit.name, []).extend(m)                      # addMetrics, L96-110
        return tags

    def _walk(self, unit, X, tags_in, routing, path, metrics):
        path[unit.name] = unit.image                                              # L123
        impl = unit.impl
        cur, tags = X, dict(tags_in)
        if unit.kind in ("MODEL", "TRANSFORMER"):                                 # transformInput, L219-243
            fn = getattr(impl, "predict" if unit.kind == "MODEL" else "transform_input")
            cur = fn(X, None)
            tags.update(self._call_meta(impl, unit, metrics))                     # mergeMeta(latest, previous), L365-372
        if not unit.children:
            return cur, tags                                                      # L136-139
        branch = -1
        if unit.kind == "ROUTER":                                                 # L146-155
            branch = impl.route(cur, None)
            if not isinstance(branch, int):
                raise SeldonB200Error("Routing response must be int but got " + str(branch))
            if branch < -1 or branch >= len(unit.children):                       # sanityCheckRouting, L345-351
                raise SeldonB200Error("Invalid branch index. Router that caused the exception: id=%s name=%s"
                                      % (unit.name, unit.name))
        routing[unit.name] = branch                                               # L159
        selected = unit.children if branch == -1 else [unit.children[branch]]     # L162-170
        outs, child_tags = [], []
        for ch in selected:
            o, t = self._walk(ch, cur, tags, routing, path, metrics)
            outs.append(o)
            child_tags.append(t)
        if unit.kind == "COMBINER":                                               # aggregate, L186
            agg = impl.aggregate(outs, None)
            latest = self._call_meta(impl, unit, metrics)
        else:
            if len(outs) != 1:
                raise SeldonB200Error("unit %s has %d child outputs and no combiner" % (unit.name, len(outs)))
            agg, latest = outs[0], {}                                             # default aggregate: the single child
        merged = {}
        for t in child_tags:                                                      # mergeMeta(latest, List previous), L354-363
            merged.update(t)
        merged.update(latest)
        if unit.kind == "OUTPUT_TRANSFORMER":                                     # transformOutput, L190
            agg = impl.transform_output(agg, None)
            merged.update(self._call_meta(impl, unit, metrics))
        return agg, merged
