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
