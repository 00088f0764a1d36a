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


def _corpus(n, seed):
    """doubles that are exactly representable in fp32, spread over magnitudes, with the layout edge cases of repr()."""
    rs = np.random.RandomState(seed)
    a = (rs.randn(n) * 10.0 ** rs.randint(-12, 13, size=n)).astype(np.float32)
    special = np.array([0.0, -0.0, 1.0, -1.0, 0.1, 0.5, 1e16, 1e15, 9999999827968.0, 1e22, 1e-5, 1e-4, 123456.789, 16777216.0,
                        3.4028234663852886e38, 1.401298464324817e-45, 1.1754943508222875e-38, 1e-7, 100.0, 2.5e-5,
                        np.inf, -np.inf, np.nan, 0.30000001192092896, 1e21, 123456792.0], dtype=np.float32)
    a[:len(special)] = special
    return a


@pytest.mark.parametrize("shape", [(1,), (7,), (1, 1), (3, 4), (1000, 17), (4096, 64), (1, 5000)])
def test_encode_json_text_equals_json_dumps(shape):
    """SURVEY 8(f) row 1, encode direction: the device text is byte for byte what json.dumps(array.tolist()) writes
    (utils.py:345-384), for floats (shortest repr) and ints."""
    import torch

    n = int(np.prod(shape))
    a = _corpus(max(n, 32), 7)[:n].reshape(shape)
    got = codec.encode_json_ndarray(torch.from_numpy(a).cuda())
    with np.errstate(all="ignore"):
        want = json.dumps(a.astype(np.float64).tolist()).encode()
    assert got == want
    ints = np.random.RandomState(1).randint(-2 ** 31, 2 ** 31 - 1, size=shape).astype(np.int32)
    ints.flat[0] = -2 ** 31
    assert codec.encode_json_ndarray(torch.from_numpy(ints).cuda()) == json.dumps(ints.tolist()).encode()
    assert codec.encode_json_ndarray(torch.empty(0, dtype=torch.float32, device="cuda")) == b"[]"


def test_encode_json_corpus_500k_and_round_trip():
    """500 k values (the size of the decode table's check corpus): every number's text equals repr(), and decoding the
    text on the device gives the values back bit for bit."""
    import torch

    a = _corpus(500000, 11)
    a = a[np.isfinite(a)]
    a = a[:len(a) // 4 * 4].reshape(-1, 4)
    t = torch.from_numpy(a).cuda()
    text = codec.encode_json_ndarray(t)
    assert text == json.dumps(a.astype(np.float64).tolist()).encode()
    back = codec.decode_json_ndarray(b'{"data":{"ndarray":' + text + b'}}')
    assert torch.equal(back.view(torch.int32), t.view(torch.int32))


def test_predict_raw_on_rest_text():
    """The REST body as text through predict_raw / transform_input_raw: numbers decoded and encoded on the device; the
    response parses to what construct_response_json builds from the oracle's flags."""
    from oracle.mahalanobis import MahalanobisLean
    from seldon_core_b200 import OutlierMahalanobisB200, synth

    prime, _ = synth.maha_rows_numpy(3000, 18, 4, seed=83)
    X, _ = synth.maha_rows_numpy(400, 18, 4, seed=84)
    det, ref = OutlierMahalanobisB200(n_components=4), MahalanobisLean(n_components=4)
    det.predict(prime, None), ref.predict(prime.astype(np.float64), None)
    names = ["c%d" % i for i in range(18)]
    body = json.dumps({"meta": {"puid": "r1"}, "data": {"names": names, "ndarray": X.astype(np.float64).tolist()}}).encode()
    out = det.predict_raw(body)
    want = ref.predict(X.astype(np.float64), names)
    resp = json.loads(out)
    assert resp["data"] == {"ndarray": want.tolist(), "names": []} and resp["meta"]["puid"] == "r1"
    assert [m["key"] for m in resp["meta"]["metrics"]] == [m["key"] for m in ref.metrics()]
    assert b'"ndarray": ' + json.dumps(want.tolist()).encode() in out        # the device text, spliced in as it is
    X2, _ = synth.maha_rows_numpy(77, 18, 4, seed=85)
    body = json.dumps({"data": {"tensor": {"shape": [77, 18], "values": X2.astype(np.float64).ravel().tolist()}}}).encode()
    out = det.transform_input_raw(body)
    want2 = ref.predict(X2.astype(np.float64), None)
    resp = json.loads(out)
    assert resp["data"]["tensor"]["shape"] == [77, 18]
    assert resp["data"]["tensor"]["values"] == X2.astype(np.float64).ravel().tolist()
    assert resp["meta"]["tags"] == {"outlier-predictions": want2.tolist()}


@pytest.mark.parametrize("text", [b"[[1 2],[3 4]]", b"[[1,,2]]", b"[[1,2,],[3,4,5]]", b"[[1,2],[3,4],]", b"[[01,2]]", b"[[1,2][3,4]]",
                                  b"[,1,2]", b"[[1,2],,[3,4]]", b"[[+1,2]]", b"[[1.,2]]", b"[[.5,2]]",
                                  b"[[-NaN,1]]", b"[[1,2]" + b" " * 40 + b"[3,4]]"])
def test_decode_rejects_what_json_loads_rejects(text):
    """the token splitter alone accepted separators json.loads refuses.  Every case here is invalid JSON (or
    not a rectangular numeric array) and must raise."""
    with pytest.raises(Exception):
        json.loads(b'{"data":{"ndarray":' + text + b'}}')
        raise ValueError("json accepts it: not rectangular")  # reached only by the ragged / nested cases
    with pytest.raises(SeldonB200Error):
        codec.decode_json_ndarray(b'{"data":{"ndarray":' + text + b'}}')


@pytest.mark.parametrize("text", [b"[[1,2],[3,4]]", b"[ [ 1 , 2 ] ,\n\t[ 3 , 4 ] ]", b"[[0,0.5],[-0,0e1]]", b"[1,2,3]",
                                  b"[[1,2]," + b" " * 70 + b"[3,4]]", b"[[1," + b"\n" * 40 + b"2],[3,4]]", b"[[1e5,2E-3],[NaN,-Infinity]]"])
def test_decode_accepts_what_json_loads_accepts(text):
    want = np.array(json.loads(text)).astype(np.float32)
    got = codec.decode_json_ndarray(b'{"data":{"ndarray":' + text + b'}}').cpu().numpy()
    assert got.shape == want.shape and np.array_equal(got, want, equal_nan=True)
