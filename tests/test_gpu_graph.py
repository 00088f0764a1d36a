"""SURVEY 8(f) row 3 on the device: RANDOM_ABTEST decisions bit-exact against the oracle (pinned by the reference's own
known answers), and a whole graph pass - transformer -> router -> combiner -> models - on device-resident tensors
against the numpy walk of oracle/graph.py (PredictiveUnitBean.java:113-197 semantics)."""
import pickle

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle import graph as ograph
from oracle.abtest import RandomABTestOracle
from oracle.combiner import mean_combiner
from oracle.mahalanobis import MahalanobisLean
from seldon_core_b200 import (GraphB200, MeanCombinerB200, OutlierMahalanobisB200, RandomABTestB200, Unit, synth)
from seldon_core_b200.errors import SeldonB200Error


def test_random_abtest_known_answers_and_batches():
    r = RandomABTestB200(ratioA=0.5)
    assert [r.route(None, None) for _ in range(3)] == [1, 0, 1]      # RandomABTestUnitInternalTest.java:59-74
    for ratio, T in ((0.5, 1), (0.3, 17), (0.7, 100003), (0.05, 3_000_000)):
        dev, ref = RandomABTestB200(ratioA=ratio), RandomABTestOracle(ratio)
        got = np.concatenate([dev.run_batched(T).cpu().numpy(), dev.run_batched(33).cpu().numpy()])
        want = ref.run(min(T, 200000))
        assert np.array_equal(got[:len(want)], want)
        if T > 200000:                                                # the tail: jump the oracle ahead like the kernel does
            from oracle.abtest import ADD, MASK, MUL

            s = ref.rand.s
            for _ in range(T - 200000):
                s = (s * MUL + ADD) & MASK
            ref.rand.s = s
        assert np.array_equal(got[T:], ref.run(33))                   # the state carried over to the next launch
    a = RandomABTestB200(ratioA=0.4)
    a.run_batched(1000)
    b = pickle.loads(pickle.dumps(a))
    assert np.array_equal(a.run_batched(500).cpu().numpy(), b.run_batched(500).cpu().numpy())
    with pytest.raises(SeldonB200Error, match="ratioA"):
        RandomABTestB200()


class DeviceModel(object):
    """A stand-in MODEL leaf (the models themselves are outside the data plane): rows -> softmax(x W)."""

    def __init__(self, W, tag):
        import torch

        self.W, self.tag = torch.tensor(W, dtype=torch.float32, device="cuda"), tag

    def predict(self, X, names):
        import torch

        return torch.softmax(X @ self.W, dim=1)

    def tags(self):
        return {"model": self.tag, self.tag: 1}


class HostModel(object):
    def __init__(self, W, tag):
        self.W, self.tag = W.astype(np.float32).astype(np.float64), tag

    def predict(self, X, names):
        z = X @ self.W
        e = np.exp(z - z.max(axis=1, keepdims=True))
        return e / e.sum(axis=1, keepdims=True)

    def tags(self):
        return {"model": self.tag, self.tag: 1}


class HostCombiner(object):
    def aggregate(self, Xs, names):
        return mean_combiner(Xs)


class DetAsTransformer(object):
    """The detector in transformer mode on device tensors (transform_input of the class API takes numpy)."""

    def __init__(self, det):
        self.det = det

    def transform_input(self, X, names):
        _, flag = self.det.score_device(X)
        self.det.prediction_meta = flag.cpu().numpy().astype(int)
        return X

    def tags(self):
        return self.det.tags()

    def metrics(self):
        return self.det.metrics()


def test_graph_pass_matches_the_numpy_walk():
    import torch

    rs = np.random.RandomState(3)
    Ws = [rs.randn(64, 8) * 0.1 for _ in range(4)]
    prime, _ = synth.maha_rows_numpy(5000, 64, 16, seed=91)
    det, ref = OutlierMahalanobisB200(n_components=16), MahalanobisLean(n_components=16)
    det.predict(prime, None), ref.predict(prime.astype(np.float64), None)
    ab_dev, ab_ref = RandomABTestB200(ratioA=0.5), RandomABTestOracle(0.5)
    # transformer(outlier) -> router(AB test) -> [ combiner(2 models), combiner(2 models) ]
    g_dev = GraphB200(Unit("outlier", "TRANSFORMER", DetAsTransformer(det), [
        Unit("ab", "ROUTER", ab_dev, [
            Unit("combA", "COMBINER", MeanCombinerB200(), [Unit("m0", "MODEL", DeviceModel(Ws[0], "m0")),
                                                           Unit("m1", "MODEL", DeviceModel(Ws[1], "m1"))]),
            Unit("combB", "COMBINER", MeanCombinerB200(), [Unit("m2", "MODEL", DeviceModel(Ws[2], "m2")),
                                                           Unit("m3", "MODEL", DeviceModel(Ws[3], "m3"))]),
        ])], image="outlier:0.1"))
    leaf = lambda n, i: {"name": n, "kind": "MODEL", "impl": HostModel(Ws[i], n)}
    g_ref = {"name": "outlier", "kind": "TRANSFORMER", "impl": ref, "image": "outlier:0.1", "children": [
        {"name": "ab", "kind": "ROUTER", "impl": ab_ref, "children": [
            {"name": "combA", "kind": "COMBINER", "impl": HostCombiner(), "children": [leaf("m0", 0), leaf("m1", 1)]},
            {"name": "combB", "kind": "COMBINER", "impl": HostCombiner(), "children": [leaf("m2", 2), leaf("m3", 3)]}]}]}
    seen = set()
    for step in range(6):
        X, _ = synth.maha_rows_numpy(2000 + step, 64, 16, seed=100 + step)
        out, meta = g_dev.predict(torch.from_numpy(X).cuda(), puid="p%d" % step)
        want, wmeta = ograph.predict(g_ref, X.astype(np.float64), puid="p%d" % step)
        assert np.allclose(out.cpu().numpy(), want, rtol=2e-5, atol=1e-7)
        assert meta["routing"] == wmeta["routing"] and meta["requestPath"] == wmeta["requestPath"]
        assert meta["puid"] == wmeta["puid"] and meta["tags"] == wmeta["tags"]     # outlier flags + the chosen models' tags
        assert sorted(meta["metrics"]) == sorted(wmeta["metrics"]) == ["outlier"]
        seen.add(meta["routing"]["ab"])
        assert meta["routing"]["combA" if meta["routing"]["ab"] == 0 else "combB"] == -1
    assert seen == {0, 1}


def test_graph_rejects_bad_routing():
    import torch

    class BadRouter(object):
        def route(self, X, names):
            return 5

    g = GraphB200(Unit("r", "ROUTER", BadRouter(), [Unit("m", "MODEL", DeviceModel(np.eye(4), "m"))]))
    with pytest.raises(SeldonB200Error, match="Invalid branch index"):
        g.predict(torch.zeros((2, 4), device="cuda"))


def test_random_abtest_against_the_committed_stream():
    """Decisions of the kernel against tests/golden/java_random_1337.npz (the first 1 000 nextFloat() values of
    java.util.Random(1337)) for several ratios: branch 0 iff nextFloat() <= ratioA (RandomABTestUnit.java:52)."""
    import os

    from conftest import GOLDEN

    f = np.load(os.path.join(GOLDEN, "java_random_1337.npz"))["next_float"]
    for ratio in (0.0, 0.1, 0.5, 0.9, 1.0, float(f[10])):
        dec = RandomABTestB200(ratioA=ratio).run_batched(1000).cpu().numpy()
        assert np.array_equal(dec, (f > np.float32(ratio)).astype(np.int32)), ratio
