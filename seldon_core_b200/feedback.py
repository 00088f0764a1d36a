"""Mahalanobis outlier detector with feedback metrics, mirroring the reference's `OutlierMahalanobis`
(components/outlier-detection/mahalanobis/OutlierMahalanobis.py:6-120): `send_feedback` records the truth labels of
the last scored batch against its predictions, `metrics` returns the 23 gauges.

The reference keeps every prediction, score and label in Python lists and re-runs scikit-learn over the whole
history on each call; here the history is four counters, two sums and a ring of the last `roll_window` pairs in
device memory (`sb_feedback_update`, csrc/feedback.cu), so a feedback call is one pass over the new flags.
"""
import ctypes

import numpy as np

from . import _cabi
from .errors import SeldonB200Error
from .outlier import OutlierMahalanobisB200


class OutlierMahalanobisFeedbackB200(OutlierMahalanobisB200):
    def __init__(self, threshold=25, n_components=3, n_stdev=3, start_clip=50, max_n=-1, device=None):
        super().__init__(threshold=threshold, n_components=n_components, n_stdev=n_stdev, start_clip=start_clip,
                         max_n=max_n, device=device)
        self.roll_window = 100                      # OutlierMahalanobis.py:30
        self.metric = [float("nan")] * 18           # L31
        self.label = None
        self._fb_state = None
        self._fb_metrics = None
        self._fb_restore = None
        self._last_pred = None
        self._last_score = None

    # ------------------------------------------------------------------ reference API
    def send_feedback(self, X, feature_names, reward, truth, routing=None):
        """OutlierMahalanobis.py:34-66: the labels `truth` belong to the rows of the last predict / transform_input."""
        import torch

        with self._lock:
            if self._flag_t is None:
                raise AttributeError("prediction")  # the reference reads self.prediction here
            dev = self._flag_t.device
            t = truth if isinstance(truth, torch.Tensor) else torch.as_tensor(np.asarray(truth).ravel())
            t = t.to(device=dev, dtype=torch.int32).contiguous().view(-1)
            if t.numel() != self._flag_t.numel():
                raise SeldonB200Error("send_feedback: %d labels for %d predictions" % (t.numel(), self._flag_t.numel()))
            L = _cabi.lib()
            stream = torch.cuda.current_stream(dev).cuda_stream
            with torch.cuda.device(dev):
                if self._fb_state is None:
                    nbytes = int(L.sb_feedback_state_bytes(int(self.roll_window)))
                    self._fb_state = torch.empty(nbytes // 8, dtype=torch.int64, device=dev)
                    self._fb_metrics = torch.empty(18, dtype=torch.float64, device=dev)
                    if self._fb_restore is not None:  # unpickled history
                        self._fb_state.copy_(torch.from_numpy(np.ascontiguousarray(self._fb_restore)))
                        self._fb_restore = None
                    else:
                        _cabi.check(L.sb_feedback_reset(self._fb_state.data_ptr(), int(self.roll_window), stream))
                _cabi.check(L.sb_feedback_update(self._flag_t.data_ptr(), t.data_ptr(), t.numel(),
                                                 self._fb_state.data_ptr(), self._fb_metrics.data_ptr(), stream))
            m = self._fb_metrics.cpu().numpy()
            # L60-64: native Python types; the counts are ints in the reference
            self.metric = [int(v) for v in m[:4]] + [float(v) for v in m[4:14]] + [int(v) for v in m[14:]]
            self.label = t[-1:].cpu().numpy() if t.numel() == 1 else t.cpu().numpy()
            self._last_pred = int(self._flag_t[-1].item())
            self._last_score = float(self._score_t[-1].item())
        return []

    def metrics(self):
        """OutlierMahalanobis.py:69-118: 23 gauges, printed one prediction late because labels arrive as feedback."""
        with self._lock:
            if self._score_t is None:
                raise AttributeError("score")
            if self._score_t.numel() > 1:
                raise ValueError("Metrics can only handle single observations.")  # L74-75
            if self.n == 1 or self._last_pred is None:
                pred = err = y_true = float("nan")  # L77-80
            else:
                pred, err, y_true = self._last_pred, self._last_score, int(np.ravel(self.label)[0])
            g = lambda key, value: {"type": "GAUGE", "key": key, "value": value}
            m = self.metric
            return [
                g("is_outlier", pred), g("outlier_score", err), g("observation", self.n - 1),
                g("threshold", self.threshold), g("label", y_true),
                g("accuracy_tot", m[4]), g("precision_tot", m[5]), g("recall_tot", m[6]), g("f1_tot", m[7]),
                g("f2_tot", m[8]),
                g("accuracy_roll", m[9]), g("precision_roll", m[10]), g("recall_roll", m[11]), g("f1_roll", m[12]),
                g("f2_roll", m[13]),
                g("true_negative", m[0]), g("false_positive", m[1]), g("false_negative", m[2]),
                g("true_positive", m[3]),
                g("nb_outliers_roll", m[14]), g("nb_labels_roll", m[15]), g("nb_outliers_tot", m[16]),
                g("nb_labels_tot", m[17]),
            ]

    # ------------------------------------------------------------------ pickling (persistence.py:84 of the wrapper)
    def __getstate__(self):
        state = super().__getstate__()
        state["fb_state"] = self._fb_restore if self._fb_state is None else self._fb_state.cpu().numpy()
        return state

    def __setstate__(self, state):
        fb = state.pop("fb_state", None)
        super().__setstate__(state)
        self._fb_state = self._fb_metrics = None
        self._fb_restore = fb  # uploaded by the first send_feedback after unpickling
        self._last_pred = self._last_score = None
