"""GPU parity of the Mahalanobis path (through the C ABI) against the reference fixtures and the fp64 oracle.

Tolerances (BASELINE.json north_star): scores 1e-4 relative (fp32 arithmetic on the device), covariance/mean 1e-4
norm-wise, projector 1e-4, flags exactly equal.  On the rows right after a cold start the matrix being inverted is
nearly singular (it has just reached full rank), the score is then as sensitive as cond(S) and even the reference
and its own port differ; those rows (the first `settle` rows after the gate opens) are checked at 5e-2.
"""
import glob
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle.mahalanobis import MahalanobisLean
from seldon_core_b200 import OutlierMahalanobisB200, _cabi, synth
from seldon_core_b200.errors import SeldonB200Error

from conftest import GOLDEN
from gpu_util import norm_err, projector_err, rel_err

CASES = sorted(os.path.basename(p)[5:-4] for p in glob.glob(os.path.join(GOLDEN, "maha_*.npz")))
TOL = 1e-4


def load_case(name):
    g = np.load(os.path.join(GOLDEN, "maha_%s.npz" % name))
    kw = {str(k): (int(v) if float(v).is_integer() else float(v)) for k, v in zip(g["kw_keys"], g["kw_vals"])}
    return g, kw, [g["X%d" % i] for i in range(int(g["n_batches"]))]


def check_batch(det, ref_score, ref_pred, n_before, k, label):
    """Compare one batch; returns the number of rows compared strictly."""
    sg = det.score
    pg = det.prediction
    nb = len(ref_score)
    err = rel_err(sg, ref_score)
    first_open = int(np.flatnonzero(ref_score)[0]) if np.any(ref_score) else nb
    # rows where the reference itself says "gate closed" must be exactly zero
    assert np.all(sg[:first_open] == 0.0), "%s: non-zero score before the gate opened" % label
    settle = first_open + (12 * k if n_before < 50 * k else 0)
    loose = slice(first_open, min(settle, nb))
    strict = slice(min(settle, nb), nb)
    if err[loose].size:
        assert err[loose].max() < 5e-2, "%s: early rows off by %g" % (label, err[loose].max())
    if err[strict].size:
        assert err[strict].max() < TOL, "%s: score rel err %g at row %d" % (
            label, err[strict].max(), settle + int(np.argmax(err[strict])))
    # flags: exact wherever the reference score is not within the 1e-4 guard band of the threshold
    thr = det.threshold
    safe = np.abs(ref_score - thr) > 2 * TOL * thr
    safe[loose] &= np.abs(ref_score[loose] - thr) > 0.1 * thr
    assert np.array_equal(pg[safe], ref_pred[safe]), "%s: flag mismatch" % label
    return int(err[strict].size)


@pytest.mark.parametrize("path", ["default", "generic", "fused", "legacy_mma"])
@pytest.mark.parametrize("name", CASES)
def test_reference_fixture(name, path):
    g, kw, batches = load_case(name)
    d = batches[0].shape[1]
    if path != "default" and d != 64 and not (path == "legacy_mma" and d == 256):
        pytest.skip("only d = 64 (and d = 256 for the Gram kernel) has more than one kernel path")
    det = OutlierMahalanobisB200(**kw)
    det.set_option(_cabi.SB_OPT_FORCE_GENERIC, int(path == "generic"))
    det.set_option(_cabi.SB_OPT_FUSED_PASS2, int(path == "fused"))
    det.set_option(_cabi.SB_OPT_LEGACY_MMA, int(path == "legacy_mma"))
    k = min(kw.get("n_components", 3), batches[0].shape[1])
    n = 0
    strict_rows = 0
    for i, X in enumerate(batches):
        pred = det.predict(X, None)
        assert pred.shape == (len(X),) and pred.dtype == int
        strict_rows += check_batch(det, g["score%d" % i], g["pred%d" % i], n, k, "%s batch %d" % (name, i))
        n += len(X)
    assert det.n == int(g["n"])
    assert norm_err(det.mean, g["mean"]) < TOL
    assert norm_err(det.C, g["C"]) < TOL
    if g["clip_lo"].size:
        assert norm_err(det.clip[0], g["clip_lo"]) < TOL and norm_err(det.clip[1], g["clip_hi"]) < TOL
    else:
        assert det.clip is None
    assert strict_rows > 0 or name in ("stream_nb1",)


def test_projector_matches_scipy_eigh():
    import scipy.linalg

    X1, _ = synth.maha_rows_numpy(5000, 64, 16, seed=3)
    det = OutlierMahalanobisB200(n_components=16)
    det.predict(X1, None)
    V = det.eigvecs()
    assert np.allclose(V.T @ V, np.eye(16), atol=1e-10)
    _, Vref = scipy.linalg.eigh(det.C, subset_by_index=[64 - 16, 63])
    assert projector_err(V, Vref) < TOL


@pytest.mark.parametrize("gram_terms,fused", [(3, 0), (1, 0), (3, 1)])
def test_c2_shape_primed_then_scored(gram_terms, fused):
    """BASELINE config 2 at a size the oracle finishes in seconds: prime with 40k rows, score 200k x 64, k = 16."""
    prime, _ = synth.maha_rows_numpy(40000, 64, 16, seed=11)
    X, labels = synth.maha_rows_numpy(200000, 64, 16, seed=12)
    det = OutlierMahalanobisB200(n_components=16)
    det.set_option(_cabi.SB_OPT_GRAM_TERMS, gram_terms)
    det.set_option(_cabi.SB_OPT_FUSED_PASS2, fused)
    ref = MahalanobisLean(n_components=16)
    det.predict(prime, None)
    ref.predict(prime.astype(np.float64), None)
    pg = det.predict(X, None)
    pr = ref.predict(X.astype(np.float64), None)
    err = rel_err(det.score, ref.score)
    assert err.max() < TOL, "max rel err %g (gram_terms=%d)" % (err.max(), gram_terms)
    assert np.array_equal(pg, pr)                      # flags bit-exact (score gap by construction)
    assert np.array_equal(pr, labels.astype(int))      # and equal to the planted labels
    assert norm_err(det.C, ref.C) < TOL and norm_err(det.mean, ref.mean) < TOL
    met = det.metrics()
    refm = ref.metrics()
    for a, b in zip(met, refm):
        assert a["key"] == b["key"] and a["type"] == "GAUGE"
        assert abs(a["value"] - b["value"]) <= 1e-4 * max(abs(b["value"]), 1e-9), a["key"]


@pytest.mark.parametrize("legacy", [0, 1])
@pytest.mark.parametrize("d", [128, 256])
def test_wide_feature_tensor_paths(d, legacy):
    """d = 128 / 256 (config 5 width): blocked Gram kernel + generalised projection kernel vs the fp64 oracle."""
    prime, _ = synth.maha_rows_numpy(6000, d, 16, seed=51)
    X, labels = synth.maha_rows_numpy(30001, d, 16, seed=52)
    det, ref = OutlierMahalanobisB200(n_components=16), MahalanobisLean(n_components=16)
    det.set_option(_cabi.SB_OPT_LEGACY_MMA, legacy)
    det.predict(prime, None), ref.predict(prime.astype(np.float64), None)
    pg, pr = det.predict(X, None), ref.predict(X.astype(np.float64), None)
    assert rel_err(det.score, ref.score).max() < TOL
    assert np.array_equal(pg, pr) and np.array_equal(pr, labels.astype(int))
    assert norm_err(det.C, ref.C) < TOL and norm_err(det.mean, ref.mean) < TOL


def test_unconstrained_data_guard_band():
    """Plain N(0, Sigma) rows: flags must agree outside the |score - thr| <= 1e-4 thr band."""
    prime, _ = synth.maha_rows_numpy(30000, 64, 16, seed=21, unconstrained=True)
    X, _ = synth.maha_rows_numpy(100000, 64, 16, seed=22, unconstrained=True)
    det, ref = OutlierMahalanobisB200(n_components=16), MahalanobisLean(n_components=16)
    det.predict(prime, None), ref.predict(prime.astype(np.float64), None)
    pg, pr = det.predict(X, None), ref.predict(X.astype(np.float64), None)
    assert rel_err(det.score, ref.score).max() < TOL
    band = np.abs(ref.score - 25.0) <= TOL * 25.0
    assert np.array_equal(pg[~band], pr[~band])
    assert pr.sum() > 1000  # ~7 % of chi^2_16 rows exceed 25


def test_two_shards_equal_one_call():
    """The multi-GPU exchange emulated on one device: two handles score the two halves of a batch from the stacked
    statistics records and must reproduce the single-handle result (the request batch is the parity unit)."""
    import torch

    prime, _ = synth.maha_rows_numpy(20000, 64, 16, seed=31)
    X, _ = synth.maha_rows_numpy(60001, 64, 16, seed=32)
    one = OutlierMahalanobisB200(n_components=16)
    one.predict(prime, None)
    one.predict(X, None)
    dets = [OutlierMahalanobisB200(n_components=16) for _ in range(2)]
    Xd = torch.from_numpy(X).cuda()
    bounds = [(0, 25000), (25000, 60001)]
    stats = torch.empty((2, 1 + 64 + 64 * 64), dtype=torch.float64, device="cuda")
    for r, d in enumerate(dets):
        d.predict(prime, None)
        d.pass1_device(Xd[bounds[r][0]:bounds[r][1]], stats[r])
    scores = []
    for r, d in enumerate(dets):
        s, f = d.finish_device(Xd[bounds[r][0]:bounds[r][1]], stats, 2, r, len(X), bounds[r][0])
        scores.append(s.cpu().numpy())
    both = np.concatenate(scores)
    assert rel_err(both, one.score).max() < 2e-5
    for d in dets:
        assert d.n == one.n and norm_err(d.C, one.C) < 1e-6 and norm_err(d.mean, one.mean) < 1e-6


def test_pickle_round_trip_continues_identically():
    import pickle

    X1, _ = synth.maha_rows_numpy(3000, 64, 16, seed=41)
    X2, _ = synth.maha_rows_numpy(3000, 64, 16, seed=42)
    a = OutlierMahalanobisB200(n_components=16)
    a.predict(X1, None)
    b = pickle.loads(pickle.dumps(a))
    pa, pb = a.predict(X2, None), b.predict(X2, None)
    assert np.array_equal(pa, pb) and np.allclose(a.score, b.score, rtol=1e-6)
    assert b.n == a.n == 6000


def test_transformer_mode_tags_and_errors():
    from seldon_core_b200.errors import SeldonB200Error

    X, _ = synth.maha_rows_numpy(500, 8, 3, seed=5)
    det = OutlierMahalanobisB200()
    assert det.transform_input(X, None) is X
    assert det.tags() == {"outlier-predictions": det.prediction.tolist()}
    with pytest.raises(SeldonB200Error):
        det.predict(np.zeros((4, 9), dtype=np.float32), None)   # feature count changed
    with pytest.raises(SeldonB200Error):
        det.predict(np.zeros(5), None)


def test_full_size_properties_c2():
    """BASELINE config 2 at full size (16 Mi x 64): size-independent properties instead of the oracle.
    (a) state after the batch equals the state after the same rows given the same way twice from a copy (determinism),
    (b) count/mean bookkeeping: n grows by nb and the new mean is the weighted mean of old mean and batch mean,
    (c) flags equal the planted labels (score gap), (d) the batch sums equal the sums of the outputs."""
    import torch

    nb = 1 << 24
    dev = torch.device("cuda")
    prime, _ = synth.maha_rows_torch(1 << 20, 64, 16, 101, dev)
    X, labels = synth.maha_rows_torch(nb, 64, 16, 102, dev)
    det = OutlierMahalanobisB200(n_components=16)
    det.score_device(prime)
    m0, n0 = det.mean.copy(), det.n
    s1, f1 = det.score_device(X)
    assert det.n == n0 + nb
    batch_mean = X.double().mean(dim=0).cpu().numpy()
    expect = m0 + nb / (n0 + nb) * (batch_mean - m0)
    assert np.max(np.abs(det.mean - expect)) < 5e-3      # clipping moves the mean by less than this
    assert torch.equal(f1.to(torch.int8), labels)
    nflag, ssum = int(f1.sum().item()), float(s1.double().sum().item())
    _, nf, ss = det._last_sums()
    assert nf == nflag and abs(ss - ssum) <= 1e-9 * abs(ssum)
    assert torch.isfinite(s1).all() and float(s1.min()) >= 0.0


@pytest.mark.parametrize("d", [64, 256])
def test_fused_standardise_equals_the_two_step_graph(d):
    """standardise -> detector (SURVEY config 5) with the transformer fused into the detector's loads: the same
    arithmetic per value, so scores agree with the two-component graph to the detector's own run-to-run spread (a few
    1e-7 relative: the order of its shared-memory atomics) and flags are equal; shapes without the tensor-core path
    refuse the fusion."""
    import pickle

    import torch

    from seldon_core_b200 import StandardiseB200

    dev = torch.device("cuda")
    g = torch.Generator(device=dev).manual_seed(5)
    mu = torch.randn(d, generator=g, device=dev).double().cpu().numpy()
    sigma = (0.5 + torch.rand(d, generator=g, device=dev)).double().cpu().numpy()
    std = StandardiseB200(mu, sigma)
    two, one = OutlierMahalanobisB200(n_components=16), OutlierMahalanobisB200(n_components=16)
    one.fuse_standardise(std)
    for i, rows in enumerate((40000, 70001, 128, 300000)):  # cold start (fp64 warm-up path), ragged, tiny, steady
        Xs, _ = synth.maha_rows_torch(rows, d, 16, 400 + i, dev)
        Xraw = (Xs * (torch.tensor(sigma, dtype=torch.float32, device=dev) + 1e-10)
                + torch.tensor(mu, dtype=torch.float32, device=dev)).contiguous()
        s2, f2 = two.score_device(std.transform_device(Xraw))
        s1, f1 = one.score_device(Xraw)
        assert float(((s1 - s2).abs() / s2.abs().clamp_min(1e-3)).max()) <= 2e-6 and torch.equal(f1, f2), i
        if i == 1:
            one = pickle.loads(pickle.dumps(one))
    assert np.allclose(one.mean, two.mean, rtol=1e-6, atol=1e-9) and one.n == two.n
    assert norm_err(one.C, two.C) <= 1e-6
    odd = OutlierMahalanobisB200(n_components=3)
    odd.fuse_standardise(StandardiseB200(mu[:18], sigma[:18]))
    with pytest.raises(SeldonB200Error):
        odd.score_device(torch.randn(500, 18, device=dev))


def _counters(det):
    import ctypes

    import torch

    out = (ctypes.c_int64 * 8)()
    _cabi.check(_cabi.lib().sb_maha_debug_counters(det._h, out, torch.cuda.current_stream().cuda_stream))
    return list(out)


@pytest.mark.parametrize("d", [128, 256])
def test_subspace_iteration_equals_the_jacobi_solver(d):
    """d = 128 / 256, batches after the first: the top-k eigenvectors come from the warm-started block subspace
    iteration (csrc/maha_subspace.cu).  On a spectrum with a gap it must be accepted and give the Jacobi solver's
    projector, scores and flags; the oracle (scipy eigh) is the third witness."""
    a, b, ref = (OutlierMahalanobisB200(n_components=16), OutlierMahalanobisB200(n_components=16),
                 MahalanobisLean(n_components=16))
    b.set_option(_cabi.SB_OPT_NO_SUBSPACE, 1)
    for i, rows in enumerate((6000, 30001, 20000, 64, 9000)):
        X, _ = synth.maha_rows_numpy(rows, d, 16, seed=700 + i)
        pa, pb = a.predict(X, None), b.predict(X, None)
        pr = ref.predict(X.astype(np.float64), None)
        ca, cb = _counters(a), _counters(b)
        assert cb[7] == 0
        if i == 0:
            assert ca[7] == 0                      # no previous basis yet: the Jacobi solver
        else:
            assert 1 <= ca[7] <= 8, "subspace iteration not accepted on a gapped spectrum: %d" % ca[7]
        assert projector_err(a.eigvecs(), b.eigvecs()) < 3e-5   # acceptance bound of the iteration: 1e-5 * sqrt(2)
        assert projector_err(a.eigvecs(), ref.V) < TOL
        if i > 0:
            assert rel_err(a.score, ref.score).max() < TOL and np.array_equal(pa, pr) and np.array_equal(pa, pb)


def test_subspace_iteration_falls_back_without_a_gap():
    """Isotropic data has no spectral gap at k: the iteration cannot prove that its block holds the dominant subspace,
    reports failure, and the Jacobi solver runs - results equal to those of a handle with the iteration switched off."""
    rs = np.random.RandomState(9)
    a, b = OutlierMahalanobisB200(n_components=16), OutlierMahalanobisB200(n_components=16)
    b.set_option(_cabi.SB_OPT_NO_SUBSPACE, 1)
    for i in range(3):
        X = rs.randn(20000, 128).astype(np.float32)
        pa, pb = a.predict(X, None), b.predict(X, None)
        if i > 0:
            assert _counters(a)[7] < 0
        assert np.array_equal(pa, pb) and rel_err(a.score, b.score).max() < 1e-5


def test_cold_start_relaxed_region_report(capsys):
    """How much of a cold batch is held to the relaxed tolerance, and how bad it gets there: the
    rows between the opening of the warm-up gate and 12 k rows later, on a batch that starts from an empty state.
    Printed for the record (DESIGN.md section 2 quotes it) and bounded."""
    k, d, rows = 16, 64, 200000
    worst_loose, worst_strict, n_loose, gates = 0.0, 0.0, 0, []
    for seed in (101, 102, 103):
        X, _ = synth.maha_rows_numpy(rows, d, k, seed=seed)
        det, ref = OutlierMahalanobisB200(n_components=k), MahalanobisLean(n_components=k)
        det.predict(X, None), ref.predict(X.astype(np.float64), None)
        err = rel_err(det.score, ref.score)
        first_open = int(np.flatnonzero(ref.score)[0])
        settle = first_open + 12 * k
        assert np.all(det.score[:first_open] == 0.0)
        gates.append(first_open)
        n_loose += settle - first_open
        worst_loose = max(worst_loose, float(err[first_open:settle].max()))
        worst_strict = max(worst_strict, float(err[settle:].max()))
        thr = det.threshold
        safe = np.abs(ref.score - thr) > 2 * TOL * thr
        safe[first_open:settle] &= np.abs(ref.score[first_open:settle] - thr) > 0.1 * thr
        assert np.array_equal(det.prediction[safe], ref.prediction[safe])
    with capsys.disabled():
        print("cold start, d=%d k=%d, %d-row batches: gate opens at rows %s; %d of %d rows per batch (%.4f %%) are in the "
              "relaxed region; worst relative score error there %.2e (bound 5e-2), after it %.2e (bound 1e-4)"
              % (d, k, rows, gates, n_loose // 3, rows, 100.0 * n_loose / 3 / rows, worst_loose, worst_strict))
    assert worst_loose < 5e-2 and worst_strict < TOL
