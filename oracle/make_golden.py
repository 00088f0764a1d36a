"""Generate tests/golden/*.npz by running the UNMODIFIED reference (build container only).

    python -m oracle.make_golden            # rewrites every fixture

TEST INFRASTRUCTURE.  Inputs are stored as float32 (what the device path consumes), reference
outputs as float64 / int64.  A Mahalanobis case is only kept when the reference agrees with its own
line-by-line port (``MahalanobisPort``) to 1e-9: on cold starts with many components the reference's
warm-up gate can fire on determinant rounding noise (|det| of a rank-deficient matrix against 1e-8),
after which its scores are garbage that no other implementation - including its own port with a
different operation order - reproduces.  Such seeds are skipped and listed in the fixture.
"""
import logging
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from oracle import ref_shim  # noqa: E402
from oracle.mahalanobis import MahalanobisPort  # noqa: E402
from seldon_core_b200 import synth  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def f32(a):
    return np.asarray(a, dtype=np.float32)


def maha_case(name, make_batches, kw, seeds):
    Ref = ref_shim.core_mahalanobis_class()
    skipped = []
    for seed in seeds:
        batches = [f32(b) for b in make_batches(np.random.RandomState(seed))]
        ref, port = Ref(**kw), MahalanobisPort(**kw)
        scores, preds, worst = [], [], 0.0
        for X in batches:
            X64 = X.astype(np.float64)
            ref.predict(X64, None)
            port.predict(X64, None)
            worst = max(worst, float(np.max(np.abs(port.score - ref.score) / np.maximum(np.abs(ref.score), 1e-6))))
            scores.append(ref.score.copy())
            preds.append(np.asarray(ref.prediction, dtype=np.int64))
        if worst > 1e-9:
            skipped.append(seed)
            continue
        out = {"seed": seed, "skipped_seeds": np.array(skipped, dtype=np.int64), "n_batches": len(batches),
               "kw_keys": np.array(sorted(kw)), "kw_vals": np.array([float(kw[k]) for k in sorted(kw)]),
               "mean": np.asarray(ref.mean), "C": np.asarray(ref.C), "n": ref.n,
               "clip_lo": ref.clip[0] if ref.clip is not None else np.zeros(0),
               "clip_hi": ref.clip[1] if ref.clip is not None else np.zeros(0)}
        for i, (X, s, p) in enumerate(zip(batches, scores, preds)):
            out["X%d" % i], out["score%d" % i], out["pred%d" % i] = X, s, p
        np.savez_compressed(os.path.join(OUT, "maha_%s.npz" % name), **out)
        print("maha_%s: seed %d kept (ref vs port %.1e), skipped %s, outliers in last batch %d"
              % (name, seed, worst, skipped, int(preds[-1].sum())))
        return
    raise RuntimeError("no well-posed seed for " + name)


def gap_batches(d, k, sizes):
    def make(rs):
        return [synth.maha_rows_numpy(m, d, k, seed=int(rs.randint(1 << 30)))[0] for m in sizes]
    return make


def router_case(name, kind, n, seed, T, eps=0.1, rows_mod=1):
    logging.disable(logging.CRITICAL)
    p, u = synth.router_rewards(T, n)
    rows = np.array([1 + t % rows_mod for t in range(T)], dtype=np.int64)

    class Feat:
        def __init__(self, r):
            self.shape = (r, 4)

    if kind == 0:
        r = ref_shim.epsilon_greedy_class()(n_branches=n, epsilon=eps, seed=seed)
        init_best = int(r.best_branch)
    else:
        r = ref_shim.thompson_sampling_class()(n_branches=n, seed=seed)
        init_best = -1
    dec = np.zeros(T, dtype=np.int32)
    for t in range(T):
        a = int(r.route(None, None))
        dec[t] = a
        r.send_feedback(Feat(int(rows[t])), None, a, 1.0 if u[t] < p[a] else 0.0, None)
    np.savez_compressed(os.path.join(OUT, "router_%s.npz" % name), kind=kind, n_branches=n, seed=seed, epsilon=eps,
                        T=T, rows=rows, arm_p=p, u=u, decisions=dec, init_best=init_best,
                        success=np.array(r.branch_success, dtype=np.int64),
                        tries=np.array(r.branch_tries, dtype=np.int64),
                        final_best=int(getattr(r, "best_branch", -1)))
    print("router_%s: %d calls, %d distinct arms" % (name, T, len(set(dec.tolist()))))


def combiner_case():
    Comb = ref_shim.mnist_combiner_class()
    import contextlib
    import io

    a = synth.softmax_rows_numpy(64, 10, 1)
    b = synth.softmax_rows_numpy(64, 10, 2)
    with contextlib.redirect_stdout(io.StringIO()):
        out = Comb().aggregate([a.astype(np.float64), b.astype(np.float64)], [])
    np.savez_compressed(os.path.join(OUT, "combiner_mnist.npz"), a=a, b=b, out=out)
    print("combiner_mnist: ok")


def feedback_case():
    """Feedback metrics: the reference's own performance / outlier_stats (utils.py:106-142, imported unmodified) on a
    stream of feedback batches; one row of 18 numbers per batch."""
    import importlib.util
    import warnings

    spec = importlib.util.spec_from_file_location(
        "ref_maha_utils", "/root/reference/components/outlier-detection/mahalanobis/utils.py")
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    rs = np.random.RandomState(11)
    sizes = [1, 1, 1, 3, 7, 50, 120, 1, 260, 5]
    preds, labels, rows = [], [], []
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")  # undefined precision / recall on the first all-negative batches
        for i, nb in enumerate(sizes):
            y = (rs.rand(nb) < (0.0 if i < 2 else 0.2)).astype(np.int64)
            p = np.where(rs.rand(nb) < 0.8, y, 1 - y)
            labels.append(y)
            preds.append(p)
            yy, pp = np.concatenate(labels), np.concatenate(preds)
            rows.append([float(v) for v in ref.performance(yy, pp, roll_window=100)] +
                        [float(v) for v in ref.outlier_stats(yy, pp, roll_window=100)])
    np.savez_compressed(os.path.join(OUT, "feedback_stream.npz"), sizes=np.array(sizes), pred=np.concatenate(preds),
                        truth=np.concatenate(labels), metrics=np.array(rows))
    print("feedback_stream", len(sizes), "batches")


def main():
    os.makedirs(OUT, exist_ok=True)
    if not ref_shim.available():
        raise SystemExit("reference tree not found at %s" % ref_shim.REFERENCE_ROOT)
    A18 = np.random.RandomState(5).randn(18, 18)
    sc64 = np.linspace(0.2, 3, 64)
    maha_case("c1_10kx4", lambda rs: [rs.randn(10000, 4), rs.randn(10000, 4)], {}, [0])
    maha_case("d18_k4", lambda rs: [rs.randn(300, 18) @ A18, rs.randn(500, 18) @ A18, rs.randn(50, 18) @ A18 + 3],
              {"n_components": 4}, [1, 2, 3])
    maha_case("k_eq_p", lambda rs: [rs.randn(200, 5), rs.randn(200, 5)], {"n_components": 5}, [1, 2])
    maha_case("max_n", lambda rs: [rs.randn(300, 6) for _ in range(3)], {"n_components": 3, "max_n": 100}, [1, 2])
    maha_case("stream_nb1", lambda rs: [rs.randn(1, 4) for _ in range(150)], {"n_components": 3}, [1, 2])
    maha_case("mixed_nb", lambda rs: [rs.randn(m, 4) for m in (2, 1, 30, 100, 1, 1, 57)], {"n_components": 2}, [1, 2])
    maha_case("noclip", lambda rs: [rs.randn(300, 8), rs.randn(300, 8)], {"n_components": 3, "start_clip": 10 ** 9},
              [1, 2])
    maha_case("d64_k16_cold", lambda rs: [rs.randn(3000, 64) * sc64, rs.randn(4000, 64) * sc64],
              {"n_components": 16}, list(range(100, 120)))
    maha_case("d64_k16_gap", gap_batches(64, 16, (3000, 4000, 4000)), {"n_components": 16}, list(range(200, 230)))
    maha_case("d256_k16_gap", gap_batches(256, 16, (700, 700)), {"n_components": 16}, list(range(300, 330)))
    router_case("eps_32", 0, 32, 1, 20000)
    router_case("eps_3_rows", 0, 3, 5, 5000, eps=0.3, rows_mod=5)
    router_case("thompson_32", 1, 32, 1, 3000)
    router_case("thompson_2_rows", 1, 2, 3, 3000, rows_mod=7)
    combiner_case()
    feedback_case()


if __name__ == "__main__":
    main()
