"""CPU checks behind the REST JSON decode: the committed power-of-five table is the generator's output, and the
Eisel-Lemire restatement of the device conversion agrees with CPython's float() (what json.loads applies)."""
import os
import random

import numpy as np

from oracle import jsoncodec

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_committed_table_matches_generator():
    from gen_pow5_table import table

    words = []
    with open(os.path.join(ROOT, "seldon_core_b200", "csrc", "pow5_table.inc")) as f:
        for line in f:
            if line.startswith("0x"):
                hi, lo = (int(t.strip().rstrip("ul,"), 16) for t in line.split(",")[:2])
                words.append((hi, lo))
    assert words == table() and len(words) == 651


def test_restatement_matches_float_on_random_decimals():
    rnd = random.Random(7)
    for _ in range(40000):
        nd = rnd.randint(1, 19)
        w = rnd.randint(10 ** (nd - 1), 10 ** nd - 1)
        q = rnd.randint(-345, 310 - nd)
        assert jsoncodec.eisel_lemire(q, w) == float("%de%d" % (w, q)), (w, q)


def test_restatement_on_float_reprs_and_boundaries():
    xs = np.random.RandomState(0).randn(20000).astype(np.float32).astype(np.float64)
    cases = [repr(float(abs(x))) for x in xs if x != 0] + [
        "5e-324", "2.4703282292062327e-324", "2.4703282292062328e-324", "1.7976931348623157e308",
        "1.7976931348623159e308", "2.2250738585072014e-308", "2.2250738585072011e-308", "9007199254740993",
        "9007199254740992.5", "0.1", "1e23", "8.5e-46"]
    for s in cases:
        mant, _, ex = s.lower().partition("e")
        ip, _, fp = mant.partition(".")
        w, q = int(ip + fp), (int(ex) if ex else 0) - len(fp)
        assert jsoncodec.eisel_lemire(q, w) == float(s), s


def test_shortest_decimal_restatement_equals_repr():
    """The Schubfach restatement (oracle/jsoncodec.py, twin of csrc/jsonencode.cu) against CPython's repr(): random bit
    patterns, fp32-representable values, powers of two and their neighbours (the irregular interval), the extremes."""
    import math
    import random
    import struct

    from oracle import jsoncodec as J

    random.seed(5)
    vals = [0.0, -0.0, 1.0, 0.1, 1e16, 1e15, 9999999999999998.0, 1e22, 1e23, 5e-324, 2.2250738585072014e-308,
            1.7976931348623157e308, 1e-4, 1e-5, 123456.789, float(2 ** 53), 4.35, 2 / 3, 1e21, 1e-7, 100.0]
    vals += [struct.unpack("<d", struct.pack("<Q", random.getrandbits(64)))[0] for _ in range(60000)]
    vals += [float(np.float32(random.uniform(-1, 1) * 10 ** random.randint(-10, 10))) for _ in range(40000)]
    for e in range(-1074, 1024, 7):
        vals += [2.0 ** e, math.nextafter(2.0 ** e, 0), math.nextafter(2.0 ** e, math.inf)]
    n = 0
    for x in vals:
        if math.isfinite(x):
            assert J.format_repr(x) == repr(x), x
            assert J.format_repr(-x) == repr(-x), x
            n += 1
    assert n > 100000
    assert J.format_repr(float("nan")) == "NaN" and J.format_repr(float("-inf")) == "-Infinity"
    assert J.encode(np.array([[1.5, 2.0], [0.1, -3.0]])) == b"[[1.5, 2.0], [0.1, -3.0]]"


def test_pow10_table_file_is_the_generator_output():
    import os

    sys_path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools")
    import sys

    sys.path.insert(0, sys_path)
    from gen_pow10_table import K_MAX, K_MIN, table

    inc = open(os.path.join(os.path.dirname(sys_path), "seldon_core_b200", "csrc", "pow10_table.inc")).read()
    rows = [ln for ln in inc.splitlines() if ln.startswith("0x")]
    t = table()
    assert len(rows) == len(t) == K_MAX - K_MIN + 1
    for ln, (g1, g0) in zip(rows, t):
        assert ln == "0x%016xull, 0x%016xull," % (g1, g0)
