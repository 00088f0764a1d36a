"""Oracle pins for SURVEY 8(f) row 3: java.util.Random / RANDOM_ABTEST against the reference's own known answers."""
import numpy as np

from oracle.abtest import JavaRandom, RandomABTestOracle


def test_java_random_known_answers():
    assert JavaRandom(42).next_int() == -1170105035            # the JDK's documented generator, widely published value
    r = JavaRandom(0)
    assert [r.next_int() for _ in range(3)] == [-1155484576, -723955400, 1033096058]
    f = JavaRandom(1337).next_float()
    assert isinstance(f, np.float32) and 0.0 <= f < 1.0 and float(f) * (1 << 24) == int(float(f) * (1 << 24))


def test_random_abtest_reference_known_answers():
    # engine/src/test/java/io/seldon/engine/predictors/RandomABTestUnitInternalTest.java:59-74: seed 1337, ratioA 0.5
    assert RandomABTestOracle(0.5).run(3).tolist() == [1, 0, 1]
    # ratio 0 routes everything to B unless nextFloat() is exactly 0; ratio 1 everything to A (comparator <= ratioA)
    assert RandomABTestOracle(1.0).run(1000).sum() == 0
    assert RandomABTestOracle(0.0).run(1000).sum() == 1000
    frac_a = 1.0 - RandomABTestOracle(0.3).run(200000).mean()
    assert abs(frac_a - 0.3) < 0.005
