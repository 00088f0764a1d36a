"""fp64 CPU oracle for the online Mahalanobis outlier detector.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  Restates
``components/outlier-detection/mahalanobis/CoreMahalanobis.py`` of the reference:

* :class:`MahalanobisPort` follows ``CoreMahalanobis._get_preds`` (L68-151) step by
  step with the same numpy operations, including the (nb,p,p) outer-product
  tensor and the per-row inverse chain.  It is what ``bench.py`` times as the CPU
  baseline ("port") and what the lean form is validated against when the
  reference tree itself is not there.
* :class:`MahalanobisLean` is the memory-lean closed form of SURVEY.md App. A.3/A.4
  (Gram identity + two prefix sums + one direct solve per row).  It never forms
  an (nb,p,p) or (nb,k,k) array, so it runs the 16 Mi x 64 configuration the
  reference cannot hold (``CoreMahalanobis.py:99`` would need 550 GB).  The inner
  loop is in C (``oracle/maha_lean.c``) with a pure-numpy twin for cross-checking.

Pinning: both classes were run against the unmodified reference
(``oracle/ref_shim.py``) in ``oracle/make_golden.py``; the outputs are frozen
in ``tests/golden/maha_*.npz``.  The reference ships no test or golden vector
for this component, so those fixtures *are* the pin.
"""
import ctypes
import os

import numpy as np
import scipy.linalg

_EPSILON = 1e-8  # CoreMahalanobis.py:120
_RANK_TOL = 1e-12  # rank guard of the warm-up gate, see maha_lean.c / DESIGN.md


def gate_open(P):
    """|det P| > 1e-8 (CoreMahalanobis.py:124) on a numerically full-rank P (LU pivots, rank guard)."""
    lu, piv = scipy.linalg.lu_factor(P, check_finite=False)
    d = np.abs(np.diag(lu))
    if d.min() <= _RANK_TOL * d.max():
        return False
    return float(np.prod(d)) > _EPSILON


def _top_k_eigvecs(cov, k):
    """Top-k eigenvectors in ascending eigenvalue order (CoreMahalanobis.py:103)."""
    p = cov.shape[0]
    _, vecs = scipy.linalg.eigh(cov, subset_by_index=[p - k, p - 1])
    return vecs


class _State:
    """Constructor + public attributes of CoreMahalanobis.__init__ (L26-40)."""

    def __init__(self, threshold=25, n_components=3, n_stdev=3, start_clip=50, max_n=-1):
        self.threshold = threshold
        self.n_components = n_components
        self.max_n = max_n
        self.n_stdev = n_stdev
        self.start_clip = start_clip
        self.clip = None
        self.mean = 0
        self.C = 0
        self.n = 0
        self.nb_outliers = 0

    # API surface, CoreMahalanobis.py:42-65, 154-192
    def predict(self, X, feature_names=None):
        return self._get_preds(np.asarray(X, dtype=np.float64))

    def transform_input(self, X, feature_names=None):
        self.prediction_meta = self._get_preds(np.asarray(X, dtype=np.float64))
        return X

    def send_feedback(self, X, feature_names, reward, truth):
        return []

    def tags(self):
        if hasattr(self, "prediction_meta"):
            return {"outlier-predictions": self.prediction_meta.tolist()}
        return None

    def metrics(self):
        self.nb_outliers += np.sum(self.prediction)
        g = lambda k, v: {"type": "GAUGE", "key": k, "value": v}
        return [
            g("is_outlier", np.mean(self.prediction)),
            g("outlier_score", np.mean(self.score)),
            g("nb_outliers", int(self.nb_outliers)),
            g("fraction_outliers", int(self.nb_outliers) / self.n),
            g("observation", self.n),
            g("threshold", self.threshold),
        ]

    def _effective_n(self):
        # L79-82: the capped count drives every weight; the clip switch (L85) does not use it
        return min(self.n, self.max_n) if self.max_n > 0 else self.n


class MahalanobisPort(_State):
    """Same operations, same order, same temporaries as the reference."""

    def _get_preds(self, X):
        nb, p = X.shape
        k = min(self.n_components, p)
        n = self._effective_n()

        Xc = np.clip(X, self.clip[0], self.clip[1]) if self.n > self.start_clip else X  # L85-88

        idx = np.arange(nb)
        partial_means = Xc.cumsum(axis=0) / (idx + 1).reshape(nb, 1)  # L91
        blend = (idx + 1.0) / (idx + n + 1.0)  # L92
        means_after = self.mean + blend.reshape(nb, 1) * (partial_means - self.mean)  # L93
        means_before = np.empty_like(means_after)  # L94-96
        means_before[0] = self.mean
        means_before[1:] = means_after[:-1]

        dev = Xc - means_before
        w = ((n + idx) / (n + idx + 1.0)).reshape(nb, 1, 1)  # L98
        outer = w * np.matmul(dev[:, :, None], dev[:, None, :])  # L99, (nb,p,p)
        den = n + max(1, nb - 1.0)
        cov_batch = (n - 1.0) / den * self.C + 1.0 / den * outer.sum(axis=0)  # L100

        V = _top_k_eigvecs(cov_batch, k)  # L103

        proj_x = X @ V  # L106-108
        proj_xc = Xc @ V
        proj_m = means_before @ V
        if type(self.C) == int and self.C == 0:  # L109-112
            proj_cov = np.zeros((k, k))
        else:
            proj_cov = V.T @ (self.C @ V)

        pdev = proj_xc - proj_m
        wk = (1.0 / (n + idx + 1.0)).reshape(nb, 1, 1)  # L115
        Bk = wk * np.matmul(pdev[:, :, None], pdev[:, None, :])  # L116

        inv_all = np.zeros_like(Bk)  # L118
        cur = None
        for i in range(nb):  # L122-136
            if cur is None:
                if abs(np.linalg.det(proj_cov)) > _EPSILON:
                    cur = np.linalg.inv(proj_cov)
                    inv_all[i] = cur
                elif n + i != 0:
                    proj_cov = (n + i - 1.0) / (n + i) * proj_cov + Bk[i]
                continue
            cur = (n + i - 1.0) / float(n + i - 2.0) * inv_all[i - 1]
            t = Bk[i - 1] @ cur
            inv_all[i] = cur - (cur @ t) / (1.0 + np.trace(t))

        self.mean = means_after[-1]  # L139-144
        self.C = cov_batch
        self.n += nb
        if self.n > self.start_clip:
            sd = np.sqrt(np.diag(cov_batch))
            self.clip = [self.mean - self.n_stdev * sd, self.mean + self.n_stdev * sd]

        xd = proj_x - proj_m  # L147-149
        self.score = np.matmul(xd[:, None, :], np.matmul(inv_all, xd[:, :, None])).reshape(nb)
        self.prediction = (self.score > self.threshold).astype(int)
        return self.prediction


# --------------------------------------------------------------------------- lean

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def _clib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "_build", "libsb_oracle.so")
        if not os.path.isfile(path):
            from . import build as _b

            _b.build()
        lib = ctypes.CDLL(path)
        dp = ctypes.POINTER(ctypes.c_double)
        lib.sbo_maha_stats.argtypes = [dp, ctypes.c_int64, ctypes.c_int, dp, dp, dp, dp, dp]
        lib.sbo_maha_stats.restype = None
        lib.sbo_maha_score.argtypes = [
            dp, ctypes.c_int64, ctypes.c_int, ctypes.c_int, dp, dp, dp, dp, dp,
            ctypes.c_int, ctypes.c_int64, ctypes.c_int64, dp, dp, dp, ctypes.POINTER(ctypes.c_int64),
        ]
        lib.sbo_maha_score.restype = ctypes.c_int64
        _LIB = lib
    return _LIB


def _p(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double)) if a is not None else None


class MahalanobisLean(_State):
    """Closed form of SURVEY App. A.3: additive stats, one eigh, prefix sums, direct solve.

    ``backend`` = "c" (default, ``oracle/maha_lean.c``) or "numpy" (slow twin).
    ``row0``/``pre`` let a caller emulate a row shard of a larger request batch
    (used by the world_size-2 tests): see :meth:`shard_stats` / :meth:`score_shard`.
    """

    def __init__(self, *a, backend="c", **kw):
        super().__init__(*a, **kw)
        self.backend = backend

    # -- pass 1: additive statistics with shift a = m0 (App. A.2)
    def _stats(self, X, m0):
        nb, p = X.shape
        use_clip = self.n > self.start_clip
        if self.backend == "c":
            s = np.zeros(p)
            G = np.zeros((p, p))
            lo = np.ascontiguousarray(self.clip[0]) if use_clip else None
            hi = np.ascontiguousarray(self.clip[1]) if use_clip else None
            _clib().sbo_maha_stats(_p(X), nb, p, _p(lo), _p(hi), _p(m0), _p(s), _p(G))
            return s, G
        Xc = np.clip(X, self.clip[0], self.clip[1]) if use_clip else X
        D = Xc - m0
        return D.sum(axis=0), D.T @ D

    def _cov_update(self, n, nb, C0, s, G):
        # CoreMahalanobis.py:98-100 through the Gram identity (App. A.2)
        delta = s / nb
        M2 = G - np.outer(s, s) / nb
        sumB = M2 + (n * nb / (n + nb)) * np.outer(delta, delta) if n + nb > 0 else M2
        den = n + max(1, nb - 1.0)
        return (n - 1.0) / den * C0 + sumB / den, delta

    def _score(self, X, m0, V, n, PC0, Z0, Q0, row0, was_open):
        nb, p = X.shape
        k = V.shape[1]
        use_clip = self.n > self.start_clip
        score = np.zeros(nb)
        if self.backend == "c":
            lo = np.ascontiguousarray(self.clip[0]) if use_clip else None
            hi = np.ascontiguousarray(self.clip[1]) if use_clip else None
            Vc = np.ascontiguousarray(V)
            Z = np.array(Z0, dtype=np.float64)
            Q = np.array(Q0, dtype=np.float64)
            open_at = ctypes.c_int64(-1)
            _clib().sbo_maha_score(
                _p(X), nb, p, k, _p(Vc), _p(m0), _p(lo), _p(hi), _p(np.ascontiguousarray(PC0)),
                int(was_open), int(n), int(row0), _p(Z), _p(Q), _p(score), ctypes.byref(open_at),
            )
            return score, Z, Q, open_at.value
        Xc = np.clip(X, self.clip[0], self.clip[1]) if use_clip else X
        Y = (X - m0) @ V
        Zp = (Xc - m0) @ V
        S0 = (n - 1.0) * PC0
        Z = np.array(Z0, dtype=np.float64)
        Q = np.array(Q0, dtype=np.float64)
        is_open = bool(was_open)
        open_at = -1
        for r in range(nb):
            i = row0 + r
            ni = n + i
            pm = Z / ni if ni > 0 else np.zeros(k)
            S = S0 + Q - np.outer(Z, Z) / ni if ni > 0 else S0
            if not is_open:
                if i == 0:
                    P = PC0
                elif ni - 1 != 0:
                    P = S / (ni - 1.0)
                else:
                    P = None
                if P is not None and gate_open(P):
                    is_open = True
                    open_at = i
            if is_open:
                u = Y[r] - pm
                score[r] = (ni - 1.0) * (u @ np.linalg.solve(S, u))
            z = Zp[r]
            Z += z
            Q += np.outer(z, z)
        return score, Z, Q, open_at

    def _get_preds(self, X):
        X = np.ascontiguousarray(X, dtype=np.float64)
        nb, p = X.shape
        k = min(self.n_components, p)
        n = self._effective_n()
        first = type(self.C) == int and self.C == 0
        m0 = np.zeros(p) if np.isscalar(self.mean) else np.asarray(self.mean, dtype=np.float64)
        C0 = np.zeros((p, p)) if first else self.C

        s, G = self._stats(X, m0)
        C1, delta = self._cov_update(n, nb, C0, s, G)
        V = _top_k_eigvecs(C1, k)
        PC0 = V.T @ C0 @ V
        score, _, _, open_at = self._score(X, m0, V, n, PC0, np.zeros(k), np.zeros((k, k)), 0, False)

        self.mean = m0 + (nb / (n + nb)) * delta  # L93 at i = nb-1
        self.C = C1
        self.n += nb
        if self.n > self.start_clip:
            sd = np.sqrt(np.diag(C1))
            self.clip = [self.mean - self.n_stdev * sd, self.mean + self.n_stdev * sd]
        self.score = score
        self.prediction = (score > self.threshold).astype(int)
        self.V = V
        self.open_at = open_at
        return self.prediction
