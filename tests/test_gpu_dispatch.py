"""the drop-in classes driven through the reference's OWN dispatch code.

`user_model.client_*` (python/seldon_core/user_model.py:81-361) and `metrics.validate_metrics`
(python/seldon_core/metrics.py:65-89) are imported UNMODIFIED - from /root/reference in the build container, from the
verbatim copy oracle/build.py leaves in oracle/_ref/ on the GPU box - with only the two imports that need flask /
tensorflow stood in for (oracle/ref_shim.py).  What `seldon_methods` does around those calls (the `int` check on a
route, seldon_methods.py:231; the `[[r]]` wrapping, L234; the routing lookup of send_feedback, L96-97) is restated
inline because seldon_methods itself imports tensorflow."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle import ref_shim
from oracle.combiner import mean_combiner
from oracle.mahalanobis import MahalanobisLean
from seldon_core_b200 import (EpsilonGreedyB200, MeanCombinerB200, OutlierMahalanobisB200,
                              OutlierMahalanobisFeedbackB200, StandardiseB200, ThompsonSamplingB200, synth)

if not ref_shim.available():
    pytest.skip("reference files not present (run oracle/build.py where /root/reference exists)", allow_module_level=True)

um = ref_shim.user_model_module()


def test_detector_through_client_predict_tags_metrics():
    prime, _ = synth.maha_rows_numpy(4000, 64, 16, seed=71)
    X, _ = synth.maha_rows_numpy(9000, 64, 16, seed=72)
    det, ref = OutlierMahalanobisB200(n_components=16), MahalanobisLean(n_components=16)
    names = ["f%d" % i for i in range(64)]
    # client_predict first tries predict(features, names, meta=...) and falls back on TypeError (user_model.py:153-157)
    um.client_predict(det, prime.astype(np.float64), names, meta={"puid": "abc"})
    ref.predict(prime.astype(np.float64), names)
    out = um.client_predict(det, X.astype(np.float64), names, meta={})
    want = ref.predict(X.astype(np.float64), names)
    assert isinstance(out, np.ndarray) and out.shape == (9000,) and np.array_equal(out, want)
    assert um.client_class_names(det, out) == []                       # 1-D result: no names (user_model.py:130-131)
    assert um.client_custom_tags(det) is None                          # as the reference's tags() before transform_input (L168-177)
    met = um.client_custom_metrics(det)                                # -> validate_metrics
    assert [m["key"] for m in met] == ["is_outlier", "outlier_score", "nb_outliers", "fraction_outliers", "observation",
                                       "threshold"]
    for a, b in zip(met, ref.metrics()):
        assert abs(a["value"] - b["value"]) <= 1e-4 * max(abs(b["value"]), 1e-9)
    # transformer mode: X comes back unchanged, the flags travel as tags
    X2, _ = synth.maha_rows_numpy(300, 64, 16, seed=73)
    back = um.client_transform_input(det, X2.astype(np.float64), names, meta={})
    want2 = ref.predict(X2.astype(np.float64), names)
    assert np.array_equal(back, X2.astype(np.float64))
    assert um.client_custom_tags(det) == {"outlier-predictions": want2.tolist()}
    assert um.client_send_feedback(det, X2, names, 1.0, None, routing=None) == []
    assert um.client_feature_names(det, names) == names                # no feature_names(): the original ones


def test_feedback_detector_metrics_validate():
    det = OutlierMahalanobisFeedbackB200(n_components=3)
    X, _ = synth.maha_rows_numpy(200, 8, 3, seed=74)
    for i in range(60):
        um.client_predict(det, X[i:i + 1].astype(np.float64), [], meta={})
        um.client_send_feedback(det, X[i:i + 1], [], 0.0, np.array([i % 7 == 0]), routing=None)
        met = um.client_custom_metrics(det)
        assert len(met) == 23 and met[0]["key"] == "is_outlier" and met[-1]["key"] == "nb_labels_tot"


@pytest.mark.parametrize("cls", [EpsilonGreedyB200, ThompsonSamplingB200])
def test_router_through_client_route_and_feedback(cls):
    # the unmodified reference class beside it (it seeds the process-global generators: nothing else may draw from
    # `random` / `np.random` until the loop ends)
    ref_cls = ref_shim.epsilon_greedy_class() if cls is EpsilonGreedyB200 else ref_shim.thompson_sampling_class()
    r = cls(n_branches=5, seed=4)
    ref = ref_cls(n_branches=5, seed=4)
    feats = np.zeros((3, 4))
    for t in range(200):
        routing = um.client_route(r, feats, [])
        assert isinstance(routing, int)                               # seldon_methods.py:231-233
        assert np.array([[routing]]).shape == (1, 1)                  # L234: wrapped as [[r]]
        want = int(ref.route(feats, []))
        assert routing == want
        reward = float((t * 7) % 3 == 0)
        # seldon_methods.py:96-97: routing is looked up in the response's meta and passed as a keyword
        um.client_send_feedback(r, feats, [], reward, None, routing=routing)
        ref.send_feedback(feats, [], want, reward, None)                # component order (EpsilonGreedy.py:108)
    assert um.client_custom_metrics(r) == [] and um.client_custom_tags(r) == {}


def test_combiner_and_transformer_through_the_clients():
    kids = [synth.softmax_rows_numpy(129, 10, s).astype(np.float64) for s in range(4)]
    out = um.client_aggregate(MeanCombinerB200(), kids, [[], [], [], []])
    assert np.allclose(out, mean_combiner(kids), rtol=1e-6, atol=0)
    assert um.client_class_names(MeanCombinerB200(), out) == ["t:%d" % i for i in range(10)]
    rs = np.random.RandomState(0)
    mu, sigma = rs.randn(6), rs.rand(6) + 0.5
    X = rs.randn(50, 6)
    got = um.client_transform_input(StandardiseB200(mu, sigma), X, [], meta={})
    assert np.allclose(got, (X - mu) / (sigma + 1e-10), rtol=2e-6, atol=1e-6)
    assert np.array_equal(um.client_transform_output(StandardiseB200(mu, sigma), X, []), X)   # not implemented: passes through
