"""The CPU restatement of the feedback metrics against the reference's own performance / outlier_stats
(tests/golden/feedback_stream.npz, written by oracle/make_golden.py from /root/reference)."""
import os

import numpy as np

from oracle.feedback import FeedbackOracle, outlier_stats, performance

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_oracle_matches_reference_stream():
    z = np.load(os.path.join(GOLD, "feedback_stream.npz"))
    o, off = FeedbackOracle(100), 0
    for i, nb in enumerate(z["sizes"]):
        m = o.send_feedback(z["pred"][off:off + nb], z["truth"][off:off + nb])
        off += nb
        assert np.allclose(m, z["metrics"][i], rtol=1e-13, atol=0, equal_nan=True), i
    assert off > 100  # the rolling window (100) has wrapped


def test_known_answers():
    y = np.array([0, 1, 1, 0, 1, 0, 0, 0, 1, 1])
    p = np.array([0, 1, 0, 0, 1, 1, 0, 0, 1, 0])
    # values printed by the reference's utils.performance / outlier_stats for this input (roll_window=4)
    assert performance(y, p, 4) == [4, 1, 2, 3, 0.7, 0.75, 0.6, 0.6666666666666666, 0.625, 0.75, 1.0, 0.5,
                                    0.6666666666666666, 0.5555555555555556]
    assert outlier_stats(y, p, 4) == (1, 2, 4, 5)
    # no positives at all: precision, recall and F are 0 (scikit-learn's zero_division default), accuracy 1
    z = np.zeros(5, dtype=int)
    assert performance(z, z, 100)[4:9] == [1.0, 0.0, 0.0, 0.0, 0.0]
