"""CPU oracle for the epsilon-greedy and Thompson-sampling routers.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  Restates
``components/routers/epsilon-greedy/EpsilonGreedy.py`` (L35-136) and
``components/routers/thompson-sampling/ThompsonSampling.py`` (L35-115) of the
reference *including the two MT19937 streams they consume*:

* CPython ``random`` (``random.seed`` / ``random.random`` / ``random.choice``):
  ``init_by_array`` seeding, 53-bit doubles from two words, ``_randbelow`` by
  ``getrandbits`` rejection.  Third-party to the reference (CPython stdlib,
  ``Modules/_randommodule.c`` + ``Lib/random.py``, algorithm unchanged 3.6-3.12).
* NumPy legacy ``RandomState`` (``np.random.seed`` / ``choice`` / ``beta``):
  ``init_genrand`` seeding, masked-rejection bounded integers, legacy beta
  (Johnk, or two Marsaglia-Tsang gammas fed by the cached polar Gaussian).
  Third-party (numpy ``random/src/legacy/legacy-distributions.c``,
  ``random/src/distributions/distributions.c``; the legacy stream is frozen by
  NumPy policy, requirement of the reference: ``numpy>=1.15.1``).

Pinning: ``tests/test_oracle_routers.py`` checks the generators word for word
against the live ``random`` / ``numpy.random`` of this interpreter and the
decision sequences against ``tests/golden/router_*.npz`` produced by running the
unmodified reference classes (``oracle/make_golden.py``).  The reference's own
test (``test_EpsilonGreedy.py:21-25``) is statistical only and is restated in
the same test file.

A C twin (``oracle/routers.c``) runs the 10 M-call configuration.
"""
import ctypes
import math
import os

import numpy as np

_N, _M = 624, 397


class MT19937:
    """Mersenne Twister core shared by both streams."""

    def __init__(self):
        self.mt = [0] * _N
        self.pos = _N

    def init_genrand(self, s):
        mt = self.mt
        mt[0] = s & 0xFFFFFFFF
        for i in range(1, _N):
            mt[i] = (1812433253 * (mt[i - 1] ^ (mt[i - 1] >> 30)) + i) & 0xFFFFFFFF
        self.pos = _N

    def init_by_array(self, key):
        self.init_genrand(19650218)
        mt = self.mt
        i, j = 1, 0
        for _ in range(max(_N, len(key))):
            mt[i] = ((mt[i] ^ ((mt[i - 1] ^ (mt[i - 1] >> 30)) * 1664525)) + key[j] + j) & 0xFFFFFFFF
            i += 1
            j += 1
            if i >= _N:
                mt[0] = mt[_N - 1]
                i = 1
            if j >= len(key):
                j = 0
        for _ in range(_N - 1):
            mt[i] = ((mt[i] ^ ((mt[i - 1] ^ (mt[i - 1] >> 30)) * 1566083941)) - i) & 0xFFFFFFFF
            i += 1
            if i >= _N:
                mt[0] = mt[_N - 1]
                i = 1
        mt[0] = 0x80000000
        self.pos = _N

    def _twist(self):
        mt = self.mt
        for k in range(_N):
            y = (mt[k] & 0x80000000) | (mt[(k + 1) % _N] & 0x7FFFFFFF)
            mt[k] = mt[(k + _M) % _N] ^ (y >> 1) ^ (0x9908B0DF if y & 1 else 0)
        self.pos = 0

    def u32(self):
        if self.pos >= _N:
            self._twist()
        y = self.mt[self.pos]
        self.pos += 1
        y ^= y >> 11
        y ^= (y << 7) & 0x9D2C5680
        y ^= (y << 15) & 0xEFC60000
        y ^= y >> 18
        return y & 0xFFFFFFFF

    def double53(self):
        a = self.u32() >> 5
        b = self.u32() >> 6
        return (a * 67108864.0 + b) / 9007199254740992.0


class PyRandom(MT19937):
    """CPython ``random`` module stream (EpsilonGreedy.py:48,68,90,93)."""

    def seed(self, a):
        a = abs(int(a))
        key = []
        while a:
            key.append(a & 0xFFFFFFFF)
            a >>= 32
        self.init_by_array(key or [0])

    def random(self):
        return self.double53()

    def randbelow(self, n):
        k = n.bit_length()
        r = self.u32() >> (32 - k)
        while r >= n:
            r = self.u32() >> (32 - k)
        return r


class NpRandom(MT19937):
    """NumPy legacy global RandomState stream (EpsilonGreedy.py:49,124; ThompsonSampling.py:49,81)."""

    def __init__(self):
        super().__init__()
        self.has_gauss = False
        self.gauss_cache = 0.0

    def seed(self, s):
        self.init_genrand(int(s))
        self.has_gauss = False
        self.gauss_cache = 0.0

    def choice_index(self, m):
        """index drawn by ``np.random.choice(arr)`` for len(arr) == m (masked rejection, 32-bit words)."""
        rng = m - 1
        if rng == 0:
            return 0
        mask = rng
        for s in (1, 2, 4, 8, 16):
            mask |= mask >> s
        while True:
            v = self.u32() & mask
            if v <= rng:
                return v

    def gauss(self):
        if self.has_gauss:
            self.has_gauss = False
            t, self.gauss_cache = self.gauss_cache, 0.0
            return t
        while True:
            x1 = 2.0 * self.double53() - 1.0
            x2 = 2.0 * self.double53() - 1.0
            r2 = x1 * x1 + x2 * x2
            if r2 < 1.0 and r2 != 0.0:
                break
        f = math.sqrt(-2.0 * math.log(r2) / r2)
        self.gauss_cache = f * x1
        self.has_gauss = True
        return f * x2

    def std_exponential(self):
        return -math.log(1.0 - self.double53())

    def std_gamma(self, shape):
        if shape == 1.0:
            return self.std_exponential()
        if shape == 0.0:
            return 0.0
        if shape < 1.0:
            while True:
                u = self.double53()
                v = self.std_exponential()
                if u <= 1.0 - shape:
                    x = math.pow(u, 1.0 / shape)
                    if x <= v:
                        return x
                else:
                    y = -math.log((1 - u) / shape)
                    x = math.pow(1.0 - shape + shape * y, 1.0 / shape)
                    if x <= v + y:
                        return x
        b = shape - 1.0 / 3.0
        c = 1.0 / math.sqrt(9 * b)
        while True:
            while True:
                x = self.gauss()
                v = 1.0 + c * x
                if v > 0.0:
                    break
            v = v * v * v
            u = self.double53()
            if u < 1.0 - 0.0331 * (x * x) * (x * x):
                return b * v
            if math.log(u) < 0.5 * x * x + b * (1.0 - v + math.log(v)):
                return b * v

    def beta(self, a, b):
        a = float(a)
        b = float(b)
        if a <= 1.0 and b <= 1.0:
            while True:
                u = self.double53()
                v = self.double53()
                x = math.pow(u, 1.0 / a)
                y = math.pow(v, 1.0 / b)
                xpy = x + y
                if xpy <= 1.0 and u + v > 0.0:
                    if xpy > 0:
                        return x / xpy
                    lx = math.log(u) / a
                    ly = math.log(v) / b
                    lm = max(lx, ly)
                    lx -= lm
                    ly -= lm
                    return math.exp(lx - math.log(math.exp(lx) + math.exp(ly)))
        ga = self.std_gamma(a)
        gb = self.std_gamma(b)
        return ga / (ga + gb)


def n_success_failures(rows, reward):
    """EpsilonGreedy.py:132-136 / ThompsonSampling.py:111-115 (int() truncates toward zero)."""
    n_success = int(reward * rows)
    return n_success, rows - n_success


class EpsilonGreedyOracle:
    """Decision-for-decision restatement with private stream objects seeded like the globals."""

    def __init__(self, n_branches=None, epsilon=0.1, best_branch=None, seed=None):
        self.py = PyRandom()
        self.np = NpRandom()
        if seed:  # L46-49 (a falsy seed leaves the global generators untouched)
            self.py.seed(seed)
            self.np.seed(seed)
        else:
            raise ValueError("the oracle needs a seed to be comparable")
        n_branches = int(n_branches)
        assert n_branches > 0
        self.n_branches = n_branches
        self.epsilon = epsilon
        self.best_branch = best_branch if best_branch else self.py.randbelow(n_branches)  # L64-69
        self.branch_success = [0] * n_branches
        self.branch_tries = [0] * n_branches
        self.branch_values = [0] * n_branches

    def route(self, features=None, feature_names=None):
        x = self.py.random()  # L90
        if x > self.epsilon:
            return self.best_branch
        r = self.py.randbelow(self.n_branches - 1)  # L91-94: index into the list without best_branch
        return r if r < self.best_branch else r + 1

    def send_feedback(self, rows, routing, reward):
        ns, nf = n_success_failures(rows, reward)
        self.branch_success[routing] += ns
        self.branch_tries[routing] += ns + nf
        self.branch_values[routing] = self.branch_success[routing] / self.branch_tries[routing]
        top = max(self.branch_values)
        ties = [i for i, v in enumerate(self.branch_values) if v == top]
        self.best_branch = ties[self.np.choice_index(len(ties))]  # L124-125


class ThompsonSamplingOracle:
    def __init__(self, n_branches=None, seed=None):
        self.np = NpRandom()
        if seed:
            self.np.seed(seed)
        else:
            raise ValueError("the oracle needs a seed to be comparable")
        self.n_branches = int(n_branches)
        self.params = [[1, 1] for _ in range(self.n_branches)]
        self.branch_success = [0] * self.n_branches
        self.branch_tries = [0] * self.n_branches

    def route(self, features=None, feature_names=None):
        self.branch_values = [self.np.beta(a, b) for a, b in self.params]  # L81-82, arm order
        best, arg = self.branch_values[0], 0
        for j in range(1, self.n_branches):  # np.argmax: first maximum
            if self.branch_values[j] > best:
                best, arg = self.branch_values[j], j
        return arg

    def send_feedback(self, rows, routing, reward):
        ns, nf = n_success_failures(rows, reward)
        self.params[routing][0] += ns
        self.params[routing][1] += nf
        self.branch_success[routing] += ns
        self.branch_tries[routing] += ns + nf


# ------------------------------------------------------------------ C twin for the long runs

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
KIND_EPSILON_GREEDY, KIND_THOMPSON = 0, 1
REWARD_PER_CALL, REWARD_TABLE, REWARD_BERNOULLI = 0, 1, 2


def _clib():
    global _LIB
    if _LIB is None:
        from . import build as _b

        lib = ctypes.CDLL(_b.build())
        lib.sbo_router_run.restype = ctypes.c_int
        lib.sbo_router_run.argtypes = [
            ctypes.c_int, ctypes.c_int, ctypes.c_double, ctypes.c_int64, ctypes.c_int, ctypes.c_int64,
            ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
        ]
        _LIB = lib
    return _LIB


def run_batched(kind, n_branches, epsilon, seed, best_branch, T, rows_per_call=None, reward=None, reward_mode=0,
                arm_p=None):
    """T interleaved route/send_feedback calls.  Returns (decisions int32[T], state int64[2*n_branches+1]).

    reward_mode 0: reward[t] (float32) whatever the arm; 1: reward[t, arm]; 2: reward[t] is a uniform u_t and
    the reward is 1.0 if u_t < arm_p[arm] else 0.0.  rows_per_call None = 1 row per call.
    state = [success[0..n), tries[0..n), best_branch].
    """
    dec = np.zeros(T, dtype=np.int32)
    state = np.zeros(2 * n_branches + 1, dtype=np.int64)
    rows = None if rows_per_call is None else np.ascontiguousarray(rows_per_call, dtype=np.int64)
    rew = np.ascontiguousarray(reward, dtype=np.float32)
    ap = None if arm_p is None else np.ascontiguousarray(arm_p, dtype=np.float32)
    vp = lambda a: None if a is None else a.ctypes.data_as(ctypes.c_void_p)
    rc = _clib().sbo_router_run(kind, n_branches, float(epsilon), int(seed), int(best_branch or 0), T, vp(rows),
                                vp(rew), vp(ap), reward_mode, vp(dec), vp(state))
    if rc != 0:
        raise RuntimeError("sbo_router_run failed: %d" % rc)
    return dec, state
