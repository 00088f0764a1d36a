"""Row-sharded request batches against the ORACLE: the per-rank exchange is emulated on one device - one
handle per rank, the statistics records stacked in rank order exactly as the all-gather of seldon_core_b200/dist.py
delivers them - and the concatenated shard outputs must equal the fp64 oracle's single call on the whole batch (the
request batch is the parity unit)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle.mahalanobis import MahalanobisLean
from seldon_core_b200 import OutlierMahalanobisB200, synth

from gpu_util import norm_err, rel_err

TOL = 1e-4


def run_sharded(dets, Xd, bounds, d):
    import torch

    R = len(dets)
    stats = torch.empty((R, 1 + d + d * d), dtype=torch.float64, device="cuda")
    for r, det in enumerate(dets):
        det.pass1_device(Xd[bounds[r]:bounds[r + 1]], stats[r])
    scores, flags = [], []
    for r, det in enumerate(dets):
        s, f = det.finish_device(Xd[bounds[r]:bounds[r + 1]], stats, R, r, int(bounds[-1]), int(bounds[r]))
        scores.append(s.cpu().numpy())
        flags.append(f.cpu().numpy())
    return np.concatenate(scores).astype(np.float64), np.concatenate(flags).astype(int)


@pytest.mark.parametrize("d,bounds", [
    (256, [0, 5000, 5001, 5001, 17000, 17100, 30000, 52000, 60001]),   # 8 uneven shards, one of them empty
    (64, [0, 1, 40000, 40000, 90001]),                                  # a one-row and an empty shard
    (64, [0, 0, 0, 3]),                                                 # fewer rows than ranks
])
def test_uneven_shards_against_the_oracle(d, bounds):
    import torch

    prime, _ = synth.maha_rows_numpy(20000, d, 16, seed=61)
    X, labels = synth.maha_rows_numpy(bounds[-1], d, 16, seed=62)
    ref = MahalanobisLean(n_components=16)
    ref.predict(prime.astype(np.float64), None)
    pr = ref.predict(X.astype(np.float64), None)
    dets = [OutlierMahalanobisB200(n_components=16) for _ in range(len(bounds) - 1)]
    for det in dets:
        det.predict(prime, None)
    sg, pg = run_sharded(dets, torch.from_numpy(X).cuda(), bounds, d)
    assert rel_err(sg, ref.score).max() < TOL
    assert np.array_equal(pg, pr)
    if bounds[-1] > 1000:
        assert np.array_equal(pr, labels.astype(int))
    for det in dets:   # every rank ends in the oracle's state, the empty ones included
        assert det.n == ref.n and norm_err(det.C, ref.C) < TOL and norm_err(det.mean, ref.mean) < TOL
    # and the next request batch continues from there
    X2, _ = synth.maha_rows_numpy(3000, d, 16, seed=63)
    p2 = ref.predict(X2.astype(np.float64), None)
    for det in (dets[0], dets[-2]):
        assert np.array_equal(det.predict(X2, None), p2) and rel_err(det.score, ref.score).max() < TOL


def test_cold_start_batch_split_over_ranks():
    """The FIRST request batch of a deployment, split over three ranks (no priming): rank 0 finds the warm-up gate in its
    rows, the later ranks start from the projected prefix of the ranks before them and are open from their first row -
    scores equal to the oracle's single call (the rows right after the gate opening at the cold-start tolerance)."""
    import torch

    k, d = 16, 64
    bounds = [0, 20000, 20001, 70000]
    X, _ = synth.maha_rows_numpy(bounds[-1], d, k, seed=64)
    ref = MahalanobisLean(n_components=k)
    pr = ref.predict(X.astype(np.float64), None)
    dets = [OutlierMahalanobisB200(n_components=k) for _ in range(len(bounds) - 1)]
    sg, pg = run_sharded(dets, torch.from_numpy(X).cuda(), bounds, d)
    first_open = int(np.flatnonzero(ref.score)[0])
    settle = first_open + 12 * k
    assert np.all(sg[:first_open] == 0.0)
    err = rel_err(sg, ref.score)
    assert err[first_open:settle].max() < 5e-2 and err[settle:].max() < TOL
    safe = np.abs(ref.score - 25.0) > 2 * TOL * 25.0
    safe[first_open:settle] &= np.abs(ref.score[first_open:settle] - 25.0) > 2.5
    assert np.array_equal(pg[safe], pr[safe])
    for det in dets:
        assert det.n == ref.n and norm_err(det.C, ref.C) < TOL
