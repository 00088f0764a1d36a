"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): CPU restatement of the feedback metrics of the reference's
Mahalanobis outlier detector.

Follows components/outlier-detection/mahalanobis/utils.py:106-133 (`performance`: confusion matrix with labels [0, 1],
accuracy / precision / recall / F1 / F2 on the whole history and on the last `roll_window` entries) and :135-142
(`outlier_stats`), and OutlierMahalanobis.py:34-66 (`send_feedback`: append the last call's predictions and the truth
labels, recompute the 18 numbers) and :69-118 (`metrics`: the 23 gauges).  The scikit-learn functions the reference
calls are restated from their definitions for binary labels with pos_label = 1 and zero_division -> 0:
precision = tp/(tp+fp), recall = tp/(tp+fn), F_beta = (1+b^2) tp / ((1+b^2) tp + b^2 fn + fp), accuracy = mean(y == p).

Pinned by tests/golden/feedback_*.npz, which oracle/make_golden.py writes by calling the reference's own
`performance` and `outlier_stats` (imported unmodified from /root/reference, scikit-learn 1.9 in this container).
"""
import numpy as np


def _div(a, b):
    return 0.0 if b == 0 else a / b


def _scores(y_true, y_pred):
    y_true = np.asarray(y_true)
    y_pred = np.asarray(y_pred)
    ok = np.isin(y_true, (0, 1)) & np.isin(y_pred, (0, 1))
    tn = int(np.sum(ok & (y_true == 0) & (y_pred == 0)))
    fp = int(np.sum(ok & (y_true == 0) & (y_pred == 1)))
    fn = int(np.sum(ok & (y_true == 1) & (y_pred == 0)))
    tp = int(np.sum(ok & (y_true == 1) & (y_pred == 1)))
    acc = float(np.mean(y_true == y_pred)) if len(y_true) else float("nan")
    prec, rec = _div(tp, tp + fp), _div(tp, tp + fn)
    f1 = _div(2.0 * tp, 2.0 * tp + fn + fp)
    f2 = _div(5.0 * tp, 5.0 * tp + 4.0 * fn + fp)
    return (tn, fp, fn, tp), [acc, prec, rec, f1, f2]


def performance(y_true, y_pred, roll_window=100):
    """utils.py:106-133."""
    cm, tot = _scores(y_true, y_pred)
    _, roll = _scores(y_true[-roll_window:], y_pred[-roll_window:])
    return list(cm) + tot + roll


def outlier_stats(y_true, y_pred, roll_window=100):
    """utils.py:135-142."""
    return (int(np.sum(y_pred[-roll_window:])), int(np.sum(y_true[-roll_window:])), int(np.sum(y_pred)),
            int(np.sum(y_true)))


class FeedbackOracle:
    """The list-keeping part of OutlierMahalanobis.send_feedback (OutlierMahalanobis.py:46-64)."""

    def __init__(self, roll_window=100):
        self.roll_window = roll_window
        self._predictions, self._labels = [], []
        self.metric = [float("nan")] * 18

    def send_feedback(self, prediction, truth):
        self._predictions += [int(v) for v in np.ravel(prediction)]
        self._labels += [int(v) for v in np.ravel(truth)]
        y, p = np.asarray(self._labels), np.asarray(self._predictions)
        self.metric = [float(v) for v in performance(y, p, self.roll_window)] + \
                      [float(v) for v in outlier_stats(y, p, self.roll_window)]
        return self.metric
