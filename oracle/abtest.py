"""CPU oracle of the engine's RANDOM_ABTEST router.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  Restates
``engine/src/main/java/io/seldon/engine/predictors/RandomABTestUnit.java:29-59``: ``new Random(1337)`` (L31), one
``rand.nextFloat()`` per route (L45), branch 0 iff ``comparator <= ratioA`` (L52).

``java.util.Random`` is third-party to the reference (the JDK; there is no JVM in this image): its algorithm is the
one its specification publishes - seed scrambling ``(seed ^ 0x5DEECE66D) & (2^48 - 1)``, step
``s = (s * 0x5DEECE66D + 0xB) & (2^48 - 1)``, ``next(bits) = (int)(s >>> (48 - bits))``,
``nextFloat() = next(24) / (float)(1 << 24)``.  Pinned by the reference's own known answers
(``engine/src/test/java/io/seldon/engine/predictors/RandomABTestUnitInternalTest.java:59-74``: seed 1337, ratioA 0.5
-> branches 1, 0, 1) and by the widely published ``new Random(42).nextInt() == -1170105035``.
"""
import numpy as np

MUL, ADD, MASK = 0x5DEECE66D, 0xB, (1 << 48) - 1


class JavaRandom(object):
    def __init__(self, seed):
        self.s = (int(seed) ^ MUL) & MASK

    def next(self, bits):
        self.s = (self.s * MUL + ADD) & MASK
        v = self.s >> (48 - bits)
        return v - (1 << 32) if bits == 32 and v >= (1 << 31) else v

    def next_int(self):
        return self.next(32)

    def next_float(self):
        return np.float32(self.next(24)) / np.float32(1 << 24)


class RandomABTestOracle(object):
    def __init__(self, ratio_a, seed=1337):
        self.ratio_a = np.float32(ratio_a)
        self.rand = JavaRandom(seed)

    def route(self, features=None, feature_names=None):
        return 0 if self.rand.next_float() <= self.ratio_a else 1

    def run(self, T):
        return np.array([self.route() for _ in range(T)], dtype=np.int32)
