"""BASELINE config 4 at its configured size: 10 M strictly interleaved route / send_feedback calls, 32 arms, seed 1,
for BOTH routers, every decision compared with the C oracle (oracle/routers.c, the restatement of
EpsilonGreedy.py:87-130 and ThompsonSampling.py:78-104 with the CPython / NumPy MT19937 streams).

Collected late (file name) because it is the longest GPU test: ~2.5 minutes of device time."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle import routers as R
from seldon_core_b200 import EpsilonGreedyB200, ThompsonSamplingB200, _cabi, synth

T_FULL = 10_000_000
CHUNK = 1_000_000  # launches of 1 M calls: the committed generator state carries over (tested at small sizes too)


@pytest.mark.parametrize("kind", [0, 1], ids=["epsilon_greedy", "thompson"])
def test_c4_ten_million_calls_bit_exact(kind):
    p, u = synth.router_rewards(T_FULL, 32)
    ref, st = R.run_batched(kind, 32, 0.1, 1, 0, T_FULL, reward=u, reward_mode=R.REWARD_BERNOULLI, arm_p=p)
    r = EpsilonGreedyB200(n_branches=32, epsilon=0.1, seed=1) if kind == 0 else ThompsonSamplingB200(n_branches=32, seed=1)
    got = []
    for lo in range(0, T_FULL, CHUNK):
        got.append(r.run_batched(CHUNK, u[lo:lo + CHUNK], reward_mode=_cabi.SB_REWARD_BERNOULLI, arm_p=p).cpu().numpy())
    dec = np.concatenate(got)
    assert dec.dtype == np.int32 and dec.shape == (T_FULL,)
    bad = np.flatnonzero(dec != ref)
    assert bad.size == 0, "first mismatch at call %d of %d" % (int(bad[0]), T_FULL)
    assert r.branch_success == st[:32].tolist() and r.branch_tries == st[32:64].tolist()
    c = r.counters()
    print("C4 %s: 10 M decisions equal; libm guard-band hits %d, draw-by-draw fallbacks %d"
          % ("thompson" if kind else "epsilon-greedy", c["guard_hits"], c["fallbacks"]))
    assert c["fallbacks"] == 0  # the look-ahead window was never outrun
