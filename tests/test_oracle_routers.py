"""Router oracle: generators word for word against this interpreter, decisions against the reference fixtures."""
import glob
import os
import random

import numpy as np
import pytest

from oracle import routers as R

from conftest import GOLDEN

CASES = sorted(os.path.basename(p)[7:-4] for p in glob.glob(os.path.join(GOLDEN, "router_*.npz")))


@pytest.mark.parametrize("seed", [1, 7, 2 ** 33 + 5])
def test_python_stream_bit_exact(seed):
    p = R.PyRandom()
    p.seed(seed)
    random.seed(seed)
    for i in range(1500):
        assert p.random() == random.random()
        n = 1 + (i * 7) % 37
        assert p.randbelow(n) == random.choice(range(n))


@pytest.mark.parametrize("seed", [1, 12345])
def test_numpy_legacy_stream_bit_exact(seed):
    g = R.NpRandom()
    g.seed(seed)
    np.random.seed(seed)
    rs = np.random.RandomState(99)
    for it in range(6000):
        a = float(rs.choice([0.3, 0.7, 1, 1, 2, 3, 17, 250, 2000]))
        b = float(rs.choice([0.5, 1, 1, 2, 5, 40, 1000]))
        assert g.beta(a, b) == np.random.beta(a, b)
        if it % 3 == 0:
            m = int(rs.randint(1, 33))
            assert g.choice_index(m) == np.random.choice(np.arange(m))


@pytest.mark.parametrize("name", CASES)
def test_decisions_match_reference_fixture(name):
    g = np.load(os.path.join(GOLDEN, "router_%s.npz" % name))
    kind, n, seed, T = int(g["kind"]), int(g["n_branches"]), int(g["seed"]), int(g["T"])
    dec, state = R.run_batched(kind, n, float(g["epsilon"]), seed, 0, T, rows_per_call=g["rows"], reward=g["u"],
                               reward_mode=R.REWARD_BERNOULLI, arm_p=g["arm_p"])
    assert np.array_equal(dec, g["decisions"])
    assert np.array_equal(state[:n], g["success"]) and np.array_equal(state[n:2 * n], g["tries"])
    if kind == 0:
        assert int(state[2 * n]) == int(g["final_best"])
    # pure-Python twin on a prefix
    m = min(T, 1500)
    o = R.EpsilonGreedyOracle(n, float(g["epsilon"]), None, seed) if kind == 0 else R.ThompsonSamplingOracle(n, seed)
    if kind == 0:
        assert o.best_branch == int(g["init_best"])
    for t in range(m):
        a = o.route()
        assert a == int(g["decisions"][t])
        o.send_feedback(int(g["rows"][t]), a, 1.0 if g["u"][t] < g["arm_p"][a] else 0.0)


def test_reference_statistical_test_restated():
    """components/routers/epsilon-greedy/test_EpsilonGreedy.py:21-25 - P(best) ~ 1-eps within 1 %."""
    T = 100000
    dec, _ = R.run_batched(R.KIND_EPSILON_GREEDY, 3, 0.1, 1, 0, T, reward=np.zeros(T, np.float32))
    # with zero rewards every value ties at 0 and the tie-break redraws best each call; restate the reference
    # setting instead: no feedback at all => best_branch stays the initial draw
    o = R.EpsilonGreedyOracle(3, 0.1, None, 1)
    first = o.best_branch
    hits = sum(o.route() == first for _ in range(T))
    assert np.isclose(hits / T, 0.9, atol=0, rtol=0.01)
    assert dec.min() >= 0 and dec.max() <= 2


def test_constructor_errors_mirror_reference():
    """test_EpsilonGreedy.py:6-12."""
    with pytest.raises(TypeError):
        R.EpsilonGreedyOracle(seed=1)
    with pytest.raises(AssertionError):
        R.EpsilonGreedyOracle(n_branches=-1, seed=1)
