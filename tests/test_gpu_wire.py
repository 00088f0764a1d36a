"""Bytes in -> bytes out through the `*_raw` hooks (SURVEY 8 rows a1/a2/a17/a18 as ONE request path): the response of
the device path must be the bytes the protobuf runtime produces from the reference's construct_response
(utils.py:410-469, restated in oracle/seldon_pb.py) applied to the ORACLE's flags / means."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle import seldon_pb
from oracle.combiner import mean_combiner
from oracle.mahalanobis import MahalanobisLean
from seldon_core_b200 import MeanCombinerB200, OutlierMahalanobisB200, synth, wire
from seldon_core_b200.errors import SeldonB200Error

C = seldon_pb.classes()


def metrics_of(resp_msg):
    names = {0: "COUNTER", 1: "GAUGE", 2: "TIMER"}
    return [{"type": names[m.type], "key": m.key, "value": m.value} for m in resp_msg.meta.metrics]


@pytest.mark.parametrize("kind", ["tensor", "ndarray"])
@pytest.mark.parametrize("chunk_bytes", [128 << 20, 40000])
def test_predict_raw_bytes_equal_the_reference_response(kind, chunk_bytes):
    prime, _ = synth.maha_rows_numpy(4000, 64, 16, seed=81)
    X, _ = synth.maha_rows_numpy(5003, 64, 16, seed=82)
    det, ref = OutlierMahalanobisB200(n_components=16), MahalanobisLean(n_components=16)
    det.predict(prime, None), ref.predict(prime.astype(np.float64), None)
    det._wire = wire.WireStage(det._pick_device(), chunk_bytes=chunk_bytes)   # small ring: many chunks, ragged tail
    names = ["f%d" % i for i in range(64)]
    req = seldon_pb.request_message(X, kind, names, puid="req-17")
    resp = det.predict_raw(req.SerializeToString())
    want_flags = ref.predict(X.astype(np.float64), names)
    got = C["SeldonMessage"].FromString(resp.tobytes())
    # payload: same kind as the request, the oracle's flags, no names for a 1-D result, puid echoed
    assert got.data.WhichOneof("data_oneof") == kind and list(got.data.names) == [] and got.meta.puid == "req-17"
    if kind == "tensor":
        assert list(got.data.tensor.shape) == [5003] and np.array_equal(np.array(got.data.tensor.values), want_flags)
    else:
        assert np.array_equal(np.array([v.number_value for v in got.data.ndarray.values]), want_flags)
    # the six gauges, values within the score tolerance (they are float32 on the wire)
    mine = metrics_of(got)
    for a, b in zip(mine, ref.metrics()):
        assert a["key"] == b["key"] and a["type"] == "GAUGE"
        assert abs(a["value"] - b["value"]) <= 1e-4 * max(abs(b["value"]), 1e-9)
    # and the framing: byte for byte what the protobuf runtime writes for construct_response with these metrics
    want = seldon_pb.construct_response(False, req, want_flags, tags=None, metrics=mine)
    assert resp.tobytes() == want.SerializeToString()
    assert np.array_equal(wire.parse_tensor_response(resp), want_flags) if kind == "tensor" else True


@pytest.mark.parametrize("kind", ["tensor", "ndarray"])
def test_transform_input_raw_echoes_the_data_and_tags_the_flags(kind):
    prime, _ = synth.maha_rows_numpy(3000, 18, 4, seed=83)
    X, _ = synth.maha_rows_numpy(700, 18, 4, seed=84)
    det, ref = OutlierMahalanobisB200(n_components=4), MahalanobisLean(n_components=4)
    det.predict(prime, None), ref.predict(prime.astype(np.float64), None)
    names = ["c%d" % i for i in range(18)]
    req = seldon_pb.request_message(X, kind, names)
    resp = det.transform_input_raw(req.SerializeToString())
    want_flags = ref.predict(X.astype(np.float64), names)
    got = C["SeldonMessage"].FromString(resp.tobytes())
    assert got.data == req.data                                           # X unchanged, names = the request's
    assert [v.number_value for v in got.meta.tags["outlier-predictions"].list_value.values] == want_flags.tolist()
    assert det.tags() == {"outlier-predictions": want_flags.tolist()}
    want = seldon_pb.construct_response(True, req, X.astype(np.float64), tags={"outlier-predictions": want_flags.tolist()},
                                        metrics=metrics_of(got))
    assert resp.tobytes() == want.SerializeToString()


def test_predict_raw_with_message_objects_and_rest_dicts():
    X, _ = synth.maha_rows_numpy(600, 8, 3, seed=85)
    det, ref = OutlierMahalanobisB200(), MahalanobisLean()
    req = seldon_pb.request_message(X, "tensor")
    out = det.predict_raw(req)                                            # gRPC: a message object in, one out
    want = ref.predict(X.astype(np.float64), None)
    assert type(out) is type(req) and np.array_equal(np.array(out.data.tensor.values), want)
    X2, _ = synth.maha_rows_numpy(50, 8, 3, seed=86)
    body = {"meta": {"puid": "p1"}, "data": {"names": ["a"] * 8, "ndarray": X2.astype(np.float64).tolist()}}
    out = det.predict_raw(body)                                           # REST: the parsed dict in, a dict out
    want2 = ref.predict(X2.astype(np.float64), None)
    assert out["data"]["ndarray"] == want2.tolist() and out["data"]["names"] == [] and out["meta"]["puid"] == "p1"
    assert [m["key"] for m in out["meta"]["metrics"]][:2] == ["is_outlier", "outlier_score"]
    body = {"data": {"tensor": {"shape": [50, 8], "values": X2.astype(np.float64).ravel().tolist()}}}
    out = det.transform_input_raw(body)
    assert out["data"]["tensor"]["shape"] == (50, 8) and "outlier-predictions" in out["meta"]["tags"]


def test_raw_rejects_truncated_and_non_numeric_messages():
    X, _ = synth.maha_rows_numpy(100, 8, 3, seed=87)
    det = OutlierMahalanobisB200()
    good = seldon_pb.request_message(X, "tensor").SerializeToString()
    for bad in (good[:-5], good[:20], b"\x2a\x03abc", C["SeldonMessage"](strData="hello").SerializeToString(), b""):
        with pytest.raises(SeldonB200Error):
            det.predict_raw(bad)
    nd = bytearray(seldon_pb.request_message(X, "ndarray").SerializeToString())
    nd[-9] ^= 0x40                                                        # the tag byte of the last number record
    with pytest.raises(SeldonB200Error):
        det.predict_raw(bytes(nd))
    assert det.n == 0 or det.n == 100  # nothing half-committed: a later good request still works
    assert det.predict_raw(good).nbytes > 0


@pytest.mark.parametrize("kind", ["tensor", "ndarray"])
def test_aggregate_raw_bytes_equal_the_reference_response(kind):
    rs = np.random.RandomState(5)
    # dyadic values: sums and the division by the child count are exact in fp32 and fp64 -> identical bytes
    kids = [rs.randint(-64, 64, size=(333, 10)) / 8.0 for _ in range(4)]
    reqs = [seldon_pb.request_message(k, kind, puid="agg" if i == 0 else None) for i, k in enumerate(kids)]
    comb = MeanCombinerB200()
    want = seldon_pb.construct_response(False, reqs[0], mean_combiner(kids)).SerializeToString()
    assert comb.aggregate_raw([r.SerializeToString() for r in reqs]).tobytes() == want
    lst = C["SeldonMessageList"](seldonMessages=reqs)
    assert comb.aggregate_raw(lst.SerializeToString()).tobytes() == want       # one buffer holding the list
    out = comb.aggregate_raw(lst)                                               # the object the gRPC servicer passes
    assert type(out) is type(reqs[0]) and out.SerializeToString() == want
    assert list(out.data.names) == ["t:%d" % i for i in range(10)]
    # softmax outputs: within the combiner's tolerance
    kids = [synth.softmax_rows_numpy(129, 10, s).astype(np.float64) for s in range(8)]
    resp = comb.aggregate_raw([seldon_pb.request_message(k, "tensor").SerializeToString() for k in kids])
    assert np.allclose(wire.parse_tensor_response(resp), mean_combiner(kids), rtol=1e-6, atol=0)
    with pytest.raises(SeldonB200Error, match="Expected batch length"):
        comb.aggregate_raw([seldon_pb.request_message(kids[0], "tensor").SerializeToString(),
                            seldon_pb.request_message(kids[1][:, :9], "tensor").SerializeToString()])
    with pytest.raises(SeldonB200Error, match="no inputs"):
        comb.aggregate_raw([])


def test_pinned_request_helper_round_trip():
    import torch

    X = torch.randn((1000, 64), device="cuda")
    req = wire.make_tensor_request_pinned(X, names=["a", "b"], puid="z")
    msg = C["SeldonMessage"].FromString(req.tobytes())
    assert msg.meta.puid == "z" and list(msg.data.names) == ["a", "b"] and list(msg.data.tensor.shape) == [1000, 64]
    assert np.array_equal(np.array(msg.data.tensor.values).reshape(1000, 64), X.double().cpu().numpy())
