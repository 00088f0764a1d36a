"""GPU parity of the routers: decisions bit-exact against the fixtures produced by the unmodified reference classes."""
import glob
import os
import pickle

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle import routers as R
from seldon_core_b200 import EpsilonGreedyB200, ThompsonSamplingB200, _cabi, synth

from conftest import GOLDEN

CASES = sorted(os.path.basename(p)[7:-4] for p in glob.glob(os.path.join(GOLDEN, "router_*.npz")))


def make(g):
    kind, n, seed = int(g["kind"]), int(g["n_branches"]), int(g["seed"])
    if kind == 0:
        return EpsilonGreedyB200(n_branches=n, epsilon=float(g["epsilon"]), seed=seed)
    return ThompsonSamplingB200(n_branches=n, seed=seed)


@pytest.mark.parametrize("name", CASES)
def test_batched_decisions_match_reference(name):
    g = np.load(os.path.join(GOLDEN, "router_%s.npz" % name))
    r = make(g)
    n, T = int(g["n_branches"]), int(g["T"])
    if int(g["kind"]) == 0:
        assert r.best_branch == int(g["init_best"])
    dec = r.run_batched(T, g["u"], reward_mode=_cabi.SB_REWARD_BERNOULLI, arm_p=g["arm_p"], rows_per_call=g["rows"])
    assert np.array_equal(dec.cpu().numpy(), g["decisions"])
    assert r.branch_success == g["success"].tolist() and r.branch_tries == g["tries"].tolist()
    if int(g["kind"]) == 0:
        assert r.best_branch == int(g["final_best"])


@pytest.mark.parametrize("name", ["eps_3_rows", "thompson_2_rows"])
def test_call_by_call_api_and_both_feedback_orders(name):
    g = np.load(os.path.join(GOLDEN, "router_%s.npz" % name))
    r = make(g)

    class Feat:
        def __init__(self, rows):
            self.shape = (rows, 4)

    for t in range(300):
        a = r.route(None, None)
        assert type(a) is int and a == int(g["decisions"][t])          # seldon_methods.py:231 wants a Python int
        rew = 1.0 if g["u"][t] < g["arm_p"][a] else 0.0
        if t % 2:
            r.send_feedback(Feat(int(g["rows"][t])), None, a, rew, None)              # component order
        else:
            r.send_feedback(Feat(int(g["rows"][t])), None, rew, None, routing=a)      # wrapper order
    # persistence: a restored router continues the identical sequence
    r2 = pickle.loads(pickle.dumps(r))
    for t in range(300, 400):
        a, b = r.route(None, None), r2.route(None, None)
        assert a == b == int(g["decisions"][t])
        rew = 1.0 if g["u"][t] < g["arm_p"][a] else 0.0
        r.send_feedback(Feat(int(g["rows"][t])), None, a, rew, None)
        r2.send_feedback(Feat(int(g["rows"][t])), None, a, rew, None)


def test_c4_shape_against_c_oracle():
    """BASELINE config 4 shape (32 arms, seed 1, interleaved route/feedback) on 1 M eps-greedy and 100 k Thompson calls."""
    for kind, T in ((0, 1000000), (1, 100000)):
        p, u = synth.router_rewards(T, 32)
        ref, st = R.run_batched(kind, 32, 0.1, 1, 0, T, reward=u, reward_mode=R.REWARD_BERNOULLI, arm_p=p)
        r = EpsilonGreedyB200(n_branches=32, epsilon=0.1, seed=1) if kind == 0 else ThompsonSamplingB200(n_branches=32, seed=1)
        dec = r.run_batched(T, u, reward_mode=_cabi.SB_REWARD_BERNOULLI, arm_p=p).cpu().numpy()
        assert np.array_equal(dec, ref), "first mismatch at call %d" % int(np.flatnonzero(dec != ref)[0])
        assert r.branch_success == st[:32].tolist() and r.branch_tries == st[32:64].tolist()


def test_statistics_like_reference_test():
    """test_EpsilonGreedy.py:21-25: without feedback the best branch is chosen ~ (1 - eps) of the time."""
    r = EpsilonGreedyB200(n_branches=3, epsilon=0.1, seed=1)
    first = r.best_branch
    import torch

    dec = torch.empty(100000, dtype=torch.int32, device="cuda")
    _cabi.check(_cabi.lib().sb_router_run(r._handle(), 100000, None, None, None, -1, dec.data_ptr(), r._stream()))
    frac = float((dec == first).float().mean().item())
    assert np.isclose(frac, 0.9, atol=0, rtol=0.01)


@pytest.mark.parametrize("n,T,chunks", [(200, 600, (600,)), (32, 5000, (1, 7, 992, 4000)), (5, 3000, (1500, 1, 1499)),
                                        (64, 3000, (3000,)), (128, 1500, (700, 800))])
def test_thompson_window_paths_against_c_oracle(n, T, chunks):
    """The look-ahead window of the Thompson kernel: more arms than lanes, more than the tabulated 128 (draw-by-draw
    through the window), and launches that end anywhere inside a window (the committed generator state must be the one
    a draw-by-draw generator would have)."""
    p, u = synth.router_rewards(T, n)
    ref, st = R.run_batched(1, n, 0.1, 3, 0, T, reward=u, reward_mode=R.REWARD_BERNOULLI, arm_p=p)
    r = ThompsonSamplingB200(n_branches=n, seed=3)
    got, off = [], 0
    for c in chunks:
        got.append(r.run_batched(c, u[off:off + c], reward_mode=_cabi.SB_REWARD_BERNOULLI, arm_p=p).cpu().numpy())
        off += c
        if off == chunks[0]:
            r = pickle.loads(pickle.dumps(r))  # state round trip in the middle of the stream
    dec = np.concatenate(got)
    assert np.array_equal(dec, ref), "first mismatch at call %d" % int(np.flatnonzero(dec != ref)[0])
    assert r.branch_success == st[:n].tolist() and r.branch_tries == st[n:2 * n].tolist()


@pytest.mark.parametrize("n,period", [(32, 1), (32, 7), (5, 3), (200, 11), (64, 64)])
def test_thompson_draw_by_draw_fallback_keeps_the_decisions(n, period):
    """A beta draw that outruns the look-ahead window is taken draw by draw from the committed stream position
    (probability ~1e-20 per draw, so the test forces it on every `period`-th draw): decisions, bandit state and the
    generator position afterwards must be those of the C oracle."""
    T = 1500
    p, u = synth.router_rewards(T, n)
    ref, st = R.run_batched(1, n, 0.1, 5, 0, T, reward=u, reward_mode=R.REWARD_BERNOULLI, arm_p=p)
    r = ThompsonSamplingB200(n_branches=n, seed=5)
    _cabi.check(_cabi.lib().sb_router_set_option(r._handle(), _cabi.SB_ROUTER_OPT_DEBUG_FALLBACK_PERIOD, period))
    a = r.run_batched(1000, u[:1000], reward_mode=_cabi.SB_REWARD_BERNOULLI, arm_p=p).cpu().numpy()
    _cabi.check(_cabi.lib().sb_router_set_option(r._handle(), _cabi.SB_ROUTER_OPT_DEBUG_FALLBACK_PERIOD, 0))
    b = r.run_batched(500, u[1000:], reward_mode=_cabi.SB_REWARD_BERNOULLI, arm_p=p).cpu().numpy()  # windowed again
    dec = np.concatenate([a, b])
    assert np.array_equal(dec, ref), "first mismatch at call %d" % int(np.flatnonzero(dec != ref)[0])
    assert r.branch_success == st[:n].tolist() and r.branch_tries == st[n:2 * n].tolist()
    c = r.counters()
    assert c["fallbacks"] == 1000 * n // period
    assert (dec >= 0).all()


def test_history_and_branch_values_like_the_reference_classes():
    """EpsilonGreedy.py:99-101 / ThompsonSampling.py:81-90: with history=True both branch_history and value_history grow
    on every route; Thompson's branch_values are the beta samples of the last route (NumPy legacy stream)."""
    class Feat:
        shape = (1, 4)

    t = ThompsonSamplingB200(n_branches=6, seed=9, history=True)
    assert t.branch_values == [0.0] * 6
    rs = np.random.RandomState(9)
    params = [[1, 1] for _ in range(6)]
    for step in range(40):
        want = [rs.beta(a, b) for a, b in params]
        arm = t.route(None, None)
        assert arm == int(np.argmax(want))
        # the samples carry device log() results, which are within 1 ulp of glibc's, not bit-equal to them (observed:
        # 0.6934328758713832 against 0.6934328758713834) - the DECISIONS are what must be exact
        assert np.allclose(t.branch_values, want, rtol=1e-14, atol=0)
        assert t.value_history[-1] == t.branch_values and t.branch_history[-1] == arm
        rew = 1.0 if step % 3 else 0.0
        t.send_feedback(Feat, None, arm, rew, None)
        params[arm][0] += int(rew)
        params[arm][1] += 1 - int(rew)
    assert t.models_beta_params == params
    e = EpsilonGreedyB200(n_branches=4, epsilon=0.3, seed=2, history=True)
    for step in range(20):
        arm = e.route(None, None)
        assert e.value_history[-1] == e.branch_values and len(e.branch_history) == step + 1
        e.send_feedback(Feat, None, arm, float(step % 2), None)
