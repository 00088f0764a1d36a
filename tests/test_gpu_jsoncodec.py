"""REST JSON ndarray decode on the device against the reference path (json.loads + np.array, utils.py:499-505)."""
import json

import numpy as np
import pytest

from seldon_core_b200 import codec
from seldon_core_b200.errors import SeldonB200Error

pytestmark = pytest.mark.gpu


def ref_decode(payload):  # = oracle.jsoncodec.decode
    """extract_request_parts_json, ndarray branch: utils.py:499-505, narrowed like every other decode path."""
    with np.errstate(over="ignore"):  # doubles beyond the fp32 range narrow to inf, as on the device
        return np.array(json.loads(payload)["data"]["ndarray"]).astype(np.float32)


def body(arr, **kw):
    return json.dumps({"meta": {}, "data": {"names": ["a", "b"], "ndarray": arr}}, **kw).encode()


@pytest.mark.parametrize("shape", [(1, 1), (3, 4), (1000, 7), (257, 64), (5000, 18)])
def test_round_trip_of_float_reprs(shape):
    a = np.random.RandomState(shape[0]).randn(*shape).astype(np.float32).astype(np.float64)
    a.flat[0] = 0.0
    p = body(a.tolist())
    got = codec.decode_json_ndarray(p).cpu().numpy()
    assert got.shape == shape and np.array_equal(got, ref_decode(p))


def test_full_double_precision_texts_and_extremes():
    rs = np.random.RandomState(1)
    vals = np.concatenate([rs.randn(20000) * 10.0 ** rs.randint(-300, 300, 20000),        # 17 significant digits
                           [5e-324, 2.2250738585072014e-308, 1.7976931348623157e308, 9007199254740993.0, 0.1, -0.0, 1e23,
                            8.5e-46, 1.401298464324817e-45, 3.4028234663852886e38, 3.4028235677973366e38, 123456789.0]])
    p = body(vals.reshape(-1, 4).tolist())
    got = codec.decode_json_ndarray(p).cpu().numpy()
    want = ref_decode(p)
    assert np.array_equal(got, want) and np.array_equal(np.signbit(got), np.signbit(want))


def test_number_grammar_variants_whitespace_and_specials():
    p = b'{"data": {"ndarray": [ [1, -2.5e3 ,3E-2,4.000], [\n 0.0, 1e0, -0, 12345678901234567890 ] ,[NaN, Infinity, -Infinity, 7]\t] , "names": ["x"]}}'
    got = codec.decode_json_ndarray(p).cpu().numpy()
    want = np.array([[1, -2500.0, 0.03, 4.0], [0.0, 1.0, -0.0, 12345678901234567890.0],
                     [np.nan, np.inf, -np.inf, 7]], dtype=np.float32)
    assert got.shape == (3, 4) and np.array_equal(got, want, equal_nan=True)
    flat = codec.decode_json_ndarray(b'{"data":{"ndarray":[1,2,3.5]}}').cpu().numpy()
    assert flat.shape == (3,) and np.array_equal(flat, np.array([1, 2, 3.5], dtype=np.float32))
    # indent=2 output of json.dumps: one number per line
    a = np.arange(600, dtype=np.float64).reshape(100, 6) / 7
    p = body(a.tolist(), indent=2)
    assert np.array_equal(codec.decode_json_ndarray(p).cpu().numpy(), ref_decode(p))


def test_long_digit_strings():
    # more than 19 significant digits: decided from the first 19 when w and w + 1 round alike (almost always)
    vals = ["3.14159265358979323846264338327950288", "0.000000000000000000000000000001234567890123456789012345",
            "123456789012345678901234567890", "1.00000000000000000000000000000000001"]
    p = ('{"data":{"ndarray":[[' + ",".join(vals) + ']]}}').encode()
    got = codec.decode_json_ndarray(p).cpu().numpy()
    assert np.array_equal(got, ref_decode(p))


@pytest.mark.parametrize("bad", [b'{"data":{"ndarray":[[1,2],[3]]}}', b'{"data":{"ndarray":[[1,"a"],[3,4]]}}',
                                 b'{"data":{"ndarray":[[1,true],[3,4]]}}', b'{"data":{"ndarray":[[1,2],[3,4.]]}}',
                                 b'{"data":{"ndarray":[[1,2],[3,-]]}}', b'{"data":{"ndarray":[[1,2],[3,1e]]}}',
                                 b'{"data":{"ndarray":[[1,2],[3,4],[5,6,7],[8]]}}', b'{"data":{"tensor":{}}}'])
def test_rejects_non_numeric_or_ragged(bad):
    with pytest.raises(SeldonB200Error):
        codec.decode_json_ndarray(bad)


def test_chunk_boundaries_large():
    # many chunks of 4096 bytes; tokens straddle chunk ends at every alignment
    a = (np.random.RandomState(3).randn(40000, 16) * 1e3).astype(np.float32).astype(np.float64)
    p = body(a.tolist())
    assert len(p) > 100 * 4096
    assert np.array_equal(codec.decode_json_ndarray(p).cpu().numpy(), ref_decode(p))
