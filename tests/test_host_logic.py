"""Host-side logic that needs no GPU: drop-in constructors, argument conventions, sharding algebra (gloo, 2 ranks)."""
import os
import pickle
import socket

import numpy as np
import pytest

import seldon_core_b200 as s
from seldon_core_b200 import dist as sdist
from seldon_core_b200.errors import SeldonB200Error


def test_constructor_signatures_match_reference():
    import inspect

    sig = inspect.signature(s.OutlierMahalanobisB200.__init__)
    assert list(sig.parameters)[1:6] == ["threshold", "n_components", "n_stdev", "start_clip", "max_n"]
    d = {k: v.default for k, v in sig.parameters.items()}
    assert (d["threshold"], d["n_components"], d["n_stdev"], d["start_clip"], d["max_n"]) == (25, 3, 3, 50, -1)
    sig = inspect.signature(s.EpsilonGreedyB200.__init__)
    assert list(sig.parameters)[1:8] == ["n_branches", "epsilon", "best_branch", "verbose", "seed", "history",
                                         "branch_names"]
    sig = inspect.signature(s.ThompsonSamplingB200.__init__)
    assert list(sig.parameters)[1:6] == ["n_branches", "verbose", "seed", "history", "branch_names"]
    for cls, meths in ((s.OutlierMahalanobisB200, ("predict", "transform_input", "send_feedback", "tags", "metrics")),
                       (s.EpsilonGreedyB200, ("route", "send_feedback", "n_success_failures")),
                       (s.MeanCombinerB200, ("aggregate",)), (s.StandardiseB200, ("transform_input",))):
        for m in meths:
            assert callable(getattr(cls, m))


def test_router_constructor_errors_mirror_reference():
    # components/routers/epsilon-greedy/test_EpsilonGreedy.py:6-12
    with pytest.raises(TypeError):
        s.EpsilonGreedyB200()
    with pytest.raises(AssertionError):
        s.EpsilonGreedyB200(n_branches=-1)
    eg = s.EpsilonGreedyB200(n_branches=3, epsilon=0.15, seed=1, branch_names="a:b:c")
    assert eg.n_branches == 3 and eg.epsilon == 0.15 and eg.branch_names == ["a", "b", "c"]


def test_fresh_detector_attributes_and_tags():
    det = s.OutlierMahalanobisB200()
    assert det.n == 0 and det.mean == 0 and det.C == 0 and det.clip is None and det.nb_outliers == 0
    assert det.tags() is None
    assert det.send_feedback(None, None, 1.0, None) == []
    det2 = pickle.loads(pickle.dumps(det))
    assert det2.n == 0 and det2.threshold == 25


def test_combiner_shape_errors_without_gpu():
    c = s.MeanCombinerB200()
    with pytest.raises(SeldonB200Error, match="no inputs"):
        c.aggregate([], [])
    with pytest.raises(SeldonB200Error, match="not 2 dimensional"):
        c.aggregate([np.zeros(3)], [])
    with pytest.raises(SeldonB200Error, match="Expected batch length"):
        c.aggregate([np.zeros((1, 2)), np.zeros((1, 3))], [])


def test_error_type_is_wrapper_compatible():
    e = SeldonB200Error("boom")
    assert e.status_code == 400 and e.to_dict()["status"]["info"] == "boom"
    assert e.to_dict()["status"]["reason"] == "MICROSERVICE_BAD_DATA"


def test_shard_bounds_cover_the_batch():
    for total, world in ((10, 3), (16, 8), (7, 8), (1 << 24, 8)):
        edges = [sdist.shard_bounds(total, world, r) for r in range(world)]
        assert edges[0][0] == 0 and edges[-1][1] == total
        assert all(a[1] == b[0] for a, b in zip(edges, edges[1:]))


def _free_port():
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        return sk.getsockname()[1]


def _gloo_worker(rank, world, port, total_rows, d, out):
    import torch
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle.mahalanobis import MahalanobisLean

        X = np.random.RandomState(0).randn(total_rows, d)
        m0 = np.zeros(d)
        r0, r1 = sdist.shard_bounds(total_rows, world, rank)
        lean = MahalanobisLean(n_components=2)
        sv, G = lean._stats(np.ascontiguousarray(X[r0:r1]), m0)
        local = torch.tensor(np.concatenate([[r1 - r0], sv, G.ravel()]), dtype=torch.float64)
        allst = sdist.gather_stats(local)
        total, prefix = sdist.combine_stats_reference(allst.numpy(), rank)
        out.put((rank, total, prefix))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("total_rows", [101, 1])
def test_gloo_two_rank_stats_exchange_matches_single_call(total_rows):
    """World size 2 on CPU: the gathered per-shard (count, sum, Gram) records add up to the single-call statistics
    and rank 1's prefix is rank 0's record (the only collective on the path).  With fewer rows than ranks the empty
    rank contributes an all-zero record and still joins the collective (it used to fail before it, leaving the other
    rank blocked in the all-gather)."""
    import torch.multiprocessing as mp

    from oracle.mahalanobis import MahalanobisLean

    d, world = 5, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_worker, args=(r, world, port, total_rows, d, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = {}
    for _ in range(world):
        rank, total, prefix = q.get(timeout=120)
        res[rank] = (total, prefix)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    X = np.random.RandomState(0).randn(total_rows, d)
    sv, G = MahalanobisLean(n_components=2)._stats(X, np.zeros(d))
    whole = np.concatenate([[total_rows], sv, G.ravel()])
    assert np.allclose(res[0][0], whole, rtol=1e-12, atol=1e-12) and np.allclose(res[1][0], whole, rtol=1e-12)
    assert np.all(res[0][1] == 0)
    r0, r1 = sdist.shard_bounds(total_rows, world, 0)
    s0, G0 = MahalanobisLean(n_components=2)._stats(np.ascontiguousarray(X[r0:r1]), np.zeros(d))
    assert np.allclose(res[1][1], np.concatenate([[r1 - r0], s0, G0.ravel()]), rtol=1e-12, atol=1e-12)


def test_json_ndarray_span_locates_the_value_without_parsing_numbers():
    """Host side of the REST JSON decode: where the "ndarray" value starts and ends (utils.py:499-505 payload shape)."""
    import json

    import pytest

    from seldon_core_b200 import codec
    from seldon_core_b200.errors import SeldonB200Error

    body = json.dumps({"meta": {"tags": {"a": "]]"}}, "data": {"ndarray": [[1, 2.5], [3, 4]], "names": ["x", "y"]}}).encode()
    a, b, nd = codec.json_ndarray_span(body)
    assert nd == 2 and json.loads(body[a:b]) == [[1, 2.5], [3, 4]]
    body = b'{"data": {"names": ["t:0"], "ndarray" :\n [ [1.0 , 2.0] ,\n [3.0, 4.0]\n ] }}'
    a, b, nd = codec.json_ndarray_span(body)
    assert nd == 2 and json.loads(body[a:b]) == [[1.0, 2.0], [3.0, 4.0]]
    a, b, nd = codec.json_ndarray_span(b'{"data":{"ndarray":[1,2,3]}}')
    assert nd == 1 and a == 19 and b == 26
    for bad in (b'{"data":{"tensor":{"shape":[1],"values":[1]}}}', b'{"data":{"ndarray":[[1,2],[3,4]'):
        with pytest.raises(SeldonB200Error):
            codec.json_ndarray_span(bad)


def test_tensor_header_parser_rejects_truncated_and_lying_payloads():
    """shape [1000, 1000] with 16 bytes of values used to give a negative values offset."""
    import numpy as np

    from oracle import codec as ocodec
    from seldon_core_b200 import codec
    from seldon_core_b200.errors import SeldonB200Error

    good = ocodec.encode_tensor(np.arange(6.0).reshape(2, 3))
    shape, off = codec.parse_tensor_header(good)
    assert shape == [2, 3] and off == len(good) - 48 == ocodec.tensor_values_offset((2, 3))
    lying = b"\x0a\x04\xe8\x07\xe8\x07" + b"\x12\x10" + b"\x00" * 16
    for bad in (lying, good[:-8], good + b"\x00" * 8, good[:6], good[:3], b"\x0a\x02\x02\x03"):
        with pytest.raises(SeldonB200Error):
            codec.parse_tensor_header(bad)


def _gloo_router_worker(rank, world, port, out):
    import torch
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import routers as R
        from seldon_core_b200 import synth

        T = 1000
        p, u = synth.router_rewards(T, 8)
        counts = [400, 0, 600] if world == 3 else [123, 877]
        lo = sum(counts[:rank])

        def play(total, rewards):        # rank 0 only: the C oracle stands in for the device router
            dec, _ = R.run_batched(0, 8, 0.1, 3, 0, total, reward=rewards.numpy(), reward_mode=R.REWARD_BERNOULLI, arm_p=p)
            return dec

        sr = sdist.ShardedRouter(play)
        mine = sr.play(torch.from_numpy(u[lo:lo + counts[rank]]))
        out.put((rank, mine.numpy().tolist()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_gloo_router_reward_gather_and_decision_broadcast(world):
    """SURVEY 8(e), last bullet: rewards gathered to rank 0, the sequential bandit played there over all ranks'
    requests in rank order, decisions broadcast; a rank without requests still takes part."""
    import torch.multiprocessing as mp

    from oracle import routers as R
    from seldon_core_b200 import synth

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_router_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    p_arm, u = synth.router_rewards(1000, 8)
    want, _ = R.run_batched(0, 8, 0.1, 3, 0, 1000, reward=u, reward_mode=R.REWARD_BERNOULLI, arm_p=p_arm)
    got = sum((res[r] for r in range(world)), [])
    assert got == want.tolist()
