"""Feedback metrics on the device (csrc/feedback.cu through the C ABI) against the oracle and the reference fixture."""
import os
import pickle

import numpy as np
import pytest

from oracle.feedback import FeedbackOracle

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _update(state, metrics, pred, truth):
    import torch

    from seldon_core_b200 import _cabi

    p = torch.as_tensor(pred, dtype=torch.int32, device="cuda")
    t = torch.as_tensor(truth, dtype=torch.int32, device="cuda")
    _cabi.check(_cabi.lib().sb_feedback_update(p.data_ptr(), t.data_ptr(), p.numel(), state.data_ptr(),
                                               metrics.data_ptr(), torch.cuda.current_stream().cuda_stream))
    return metrics.cpu().numpy()


def _new_state(window):
    import torch

    from seldon_core_b200 import _cabi

    L = _cabi.lib()
    state = torch.empty(int(L.sb_feedback_state_bytes(window)) // 8, dtype=torch.int64, device="cuda")
    _cabi.check(L.sb_feedback_reset(state.data_ptr(), window, torch.cuda.current_stream().cuda_stream))
    return state, torch.empty(18, dtype=torch.float64, device="cuda")


def test_reference_stream_fixture():
    z = np.load(os.path.join(GOLD, "feedback_stream.npz"))
    state, metrics = _new_state(100)
    off = 0
    for i, nb in enumerate(z["sizes"]):
        m = _update(state, metrics, z["pred"][off:off + nb], z["truth"][off:off + nb])
        off += nb
        assert np.allclose(m, z["metrics"][i], rtol=1e-13, atol=0, equal_nan=True), (i, m, z["metrics"][i])


@pytest.mark.parametrize("window,sizes", [(1, [1, 1, 5]), (7, [3, 3, 3, 20, 1]), (100, [0, 250, 1, 99, 100, 101]),
                                          (4096, [100000, 5000])])
def test_against_oracle_random(window, sizes):
    rs = np.random.RandomState(window)
    state, metrics = _new_state(window)
    o = FeedbackOracle(window)
    for nb in sizes:
        y = (rs.rand(nb) < 0.3).astype(np.int64)
        p = np.where(rs.rand(nb) < 0.7, y, 1 - y)
        m = _update(state, metrics, p, y)
        if nb == 0 and not o._labels:
            continue  # the reference cannot be called before any feedback exists
        want = o.send_feedback(p, y) if nb else o.metric
        assert np.allclose(m, want, rtol=1e-13, atol=0, equal_nan=True), (nb, m, want)


def test_class_api_gauges_and_pickle():
    from seldon_core_b200 import OutlierMahalanobisFeedbackB200

    rs = np.random.RandomState(0)
    det = OutlierMahalanobisFeedbackB200(n_components=3)
    det.predict(rs.randn(500, 6), None)  # warm the state with a batch, no feedback for it
    o = FeedbackOracle(100)
    for i in range(130):
        x = rs.randn(1, 6) * (6.0 if i % 9 == 0 else 1.0)
        pred = det.predict(x, None)
        truth = np.array([1 if i % 9 == 0 else 0])
        assert det.send_feedback(x, None, 0.0, truth) == []
        want = o.send_feedback(pred, truth)
        assert np.allclose(det.metric, want, rtol=1e-13, equal_nan=True)
        if i == 60:
            det = pickle.loads(pickle.dumps(det))
    g = {d["key"]: d["value"] for d in det.metrics()}
    assert len(g) == 23 and all(d["type"] == "GAUGE" for d in det.metrics())
    assert g["true_negative"] + g["false_positive"] + g["false_negative"] + g["true_positive"] == 130
    assert g["nb_labels_tot"] == 15 and g["observation"] == det.n - 1 and g["label"] == 0
    assert g["is_outlier"] == int(pred[0]) and g["accuracy_roll"] == want[9]
    det.predict(rs.randn(4, 6), None)
    with pytest.raises(ValueError):
        det.metrics()  # single observations only, as in the reference
