"""The Mahalanobis oracle against the fixtures produced by the unmodified reference (oracle/make_golden.py)."""
import glob
import os

import numpy as np
import pytest

from oracle import ref_shim
from oracle.mahalanobis import MahalanobisLean, MahalanobisPort

from conftest import GOLDEN

CASES = sorted(os.path.basename(p)[5:-4] for p in glob.glob(os.path.join(GOLDEN, "maha_*.npz")))


def load_case(name):
    g = np.load(os.path.join(GOLDEN, "maha_%s.npz" % name))
    kw = {str(k): (int(v) if float(v).is_integer() else float(v)) for k, v in zip(g["kw_keys"], g["kw_vals"])}
    batches = [g["X%d" % i] for i in range(int(g["n_batches"]))]
    return g, kw, batches


def rel_err(a, ref):
    return float(np.max(np.abs(a - ref) / np.maximum(np.abs(ref), 1e-6)))


@pytest.mark.parametrize("name", CASES)
@pytest.mark.parametrize("impl", ["port", "lean_c", "lean_numpy"])
def test_oracle_matches_reference_fixture(name, impl):
    g, kw, batches = load_case(name)
    if impl == "lean_numpy" and sum(len(b) for b in batches) > 8000:
        pytest.skip("numpy twin is slow; covered by the C twin")
    if impl == "port" and batches[0].shape[1] > 64:
        pytest.skip("port materialises (nb,p,p)")
    m = {"port": MahalanobisPort, "lean_c": MahalanobisLean,
         "lean_numpy": lambda **k: MahalanobisLean(backend="numpy", **k)}[impl](**kw)
    for i, X in enumerate(batches):
        pred = m.predict(X.astype(np.float64), None)
        assert rel_err(m.score, g["score%d" % i]) < 1e-9          # fp64 vs fp64: far inside the 1e-4 budget
        assert np.array_equal(pred, g["pred%d" % i])               # flags exact
    assert m.n == int(g["n"])
    assert np.allclose(m.mean, g["mean"], rtol=1e-11, atol=1e-12)
    assert np.max(np.abs(m.C - g["C"])) <= 1e-11 * np.max(np.abs(g["C"]))
    if g["clip_lo"].size:
        assert np.allclose(m.clip[0], g["clip_lo"], rtol=1e-10, atol=1e-12)
        assert np.allclose(m.clip[1], g["clip_hi"], rtol=1e-10, atol=1e-12)


def test_batch_boundary_dependence():
    """one 800-row call differs from two 400-row calls - the request batch is the parity unit."""
    rs = np.random.RandomState(3)
    warm, X = rs.randn(500, 6), rs.randn(800, 6)
    a, b = MahalanobisLean(), MahalanobisLean()
    a.predict(warm, None), b.predict(warm, None)
    a.predict(X, None)
    one = a.score.copy()
    b.predict(X[:400], None)
    s1 = b.score.copy()
    b.predict(X[400:], None)
    two = np.concatenate([s1, b.score])
    assert rel_err(two, one) > 1e-3


def test_rank_guard_cold_start_gate():
    """d=64,k=16 cold start: rows 0..k score 0, the gate opens at row k+1 and never on a rank-deficient matrix."""
    rs = np.random.RandomState(100)  # a seed on which the reference's gate fires on noise (skipped by make_golden)
    X = rs.randn(400, 64) * np.linspace(0.2, 3, 64)
    m = MahalanobisLean(n_components=16)
    m.predict(X, None)
    assert m.open_at == 17
    assert np.all(m.score[:17] == 0.0) and np.all(m.score[17:] > 0.0)
    assert np.all(np.isfinite(m.score)) and m.score.max() < 1e6


def test_metrics_and_tags_contract():
    m = MahalanobisLean()
    X = np.random.RandomState(0).randn(60, 4)
    assert m.tags() is None
    out = m.transform_input(X, None)
    assert out is X
    assert m.tags() == {"outlier-predictions": m.prediction.tolist()}
    met = m.metrics()
    assert [d["key"] for d in met] == ["is_outlier", "outlier_score", "nb_outliers", "fraction_outliers",
                                       "observation", "threshold"]
    assert all(d["type"] == "GAUGE" for d in met)


@pytest.mark.skipif(not ref_shim.available(), reason="reference tree not present (GPU box)")
def test_live_reference_random_regimes():
    Ref = ref_shim.core_mahalanobis_class()
    rs = np.random.RandomState(42)
    for p, k, sizes in ((4, 3, (100, 1, 1, 40)), (7, 2, (60, 60)), (12, 12, (200, 100))):
        ref, lean = Ref(n_components=k), MahalanobisLean(n_components=k)
        for nb in sizes:
            X = rs.randn(nb, p).astype(np.float32).astype(np.float64)
            pr, pl = ref.predict(X, None), lean.predict(X, None)
            assert rel_err(lean.score, ref.score) < 1e-9
            assert np.array_equal(pr, pl)
