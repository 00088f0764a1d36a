"""Codec oracle against the protobuf runtime and the reference's known answers."""
import numpy as np
import pytest

from oracle import codec


def tensor_message_class():
    """Tensor{repeated int32 shape=1; repeated double values=2} built at run time (proto/prediction.proto:33-38)."""
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

    fd = descriptor_pb2.FileDescriptorProto(name="sb200_tensor_test.proto", package="sb200test", syntax="proto3")
    m = fd.message_type.add(name="Tensor")
    m.field.add(name="shape", number=1, type=descriptor_pb2.FieldDescriptorProto.TYPE_INT32,
                label=descriptor_pb2.FieldDescriptorProto.LABEL_REPEATED)
    m.field.add(name="values", number=2, type=descriptor_pb2.FieldDescriptorProto.TYPE_DOUBLE,
                label=descriptor_pb2.FieldDescriptorProto.LABEL_REPEATED)
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    return message_factory.GetMessageClass(pool.FindMessageTypeByName("sb200test.Tensor"))


def test_tensor_bytes_match_protobuf():
    Tensor = tensor_message_class()
    rs = np.random.RandomState(0)
    for shape in ((2, 3), (1, 1), (300, 7), (5,), (129, 200)):
        a = rs.randn(*shape)
        ref = Tensor(shape=list(shape), values=a.ravel().tolist()).SerializeToString()
        mine = codec.encode_tensor(a)
        assert mine == ref
        assert codec.tensor_values_offset(shape) == len(ref) - 8 * a.size
        assert np.array_equal(codec.decode_tensor(ref), a)


def test_tensor_known_answer_from_survey():
    a = np.array([[1.5, 2.5, 3.5], [4.5, 5.5, 6.5]])
    b = codec.encode_tensor(a)
    assert b[:6].hex() == "0a0202031230"
    assert b[6:14].hex() == "000000000000f83f" and b[-8:].hex() == "0000000000001a40"


def test_listvalue_bytes_match_protobuf():
    from google.protobuf.struct_pb2 import ListValue

    rs = np.random.RandomState(1)
    for shape in ((3, 4), (1, 1), (2, 64), (5, 256), (7,), (3, 12)):
        a = rs.randn(*shape)
        a.flat[0] = 0.0  # zero still takes 9 bytes (oneof presence)
        lv = ListValue()
        if a.ndim == 1:
            lv.extend(a.tolist())
        else:
            for row in a:
                lv.add_list().extend(row.tolist())
        ref = lv.SerializeToString()
        assert codec.encode_listvalue(a) == ref
        assert np.array_equal(codec.decode_listvalue(ref), a)
        if a.ndim == 2:
            row_bytes, hdr = codec.listvalue_row_layout(shape[1])
            assert row_bytes * shape[0] == len(ref)
    assert codec.listvalue_row_layout(4)[0] == 48
    assert codec.listvalue_row_layout(64)[0] == 710
    assert codec.listvalue_row_layout(256)[0] == 2822


def test_reference_known_answers():
    # python/tests/test_utils.py:53-71 - identity model on [[1,2,3]] as tensor
    req = {"data": {"tensor": {"shape": [1, 3], "values": [1, 2, 3]}}}
    X, kind = codec.extract_json(req)
    assert kind == "tensor" and X.shape == (1, 3)
    out = codec.response_datadef_json(req, X)
    assert out["tensor"]["values"] == [1, 2, 3] and tuple(out["tensor"]["shape"]) == (1, 3)
    # test_utils.py:342-380 - shape (2,1), values [1,2]
    X2, _ = codec.extract_json({"data": {"tensor": {"shape": [2, 1], "values": [1, 2]}}})
    assert X2.shape == (2, 1) and X2[1, 0] == 2
    # ndarray request is answered as ndarray; no data -> tensor (utils.py:356-384)
    req_nd = {"data": {"ndarray": [[1, 2], [3, 4]]}}
    Xn, kind = codec.extract_json(req_nd)
    assert kind == "ndarray" and codec.response_datadef_json(req_nd, Xn) == {"ndarray": [[1, 2], [3, 4]]}
    assert "tensor" in codec.response_datadef_json({"strData": "x"}, Xn)
    # python/tests/test_model_microservice.py:143-155 - names t:0,t:1 for a 2-D result, [] for 1-D
    assert codec.class_names(np.zeros((1, 2))) == ["t:0", "t:1"]
    assert codec.class_names(np.zeros(5)) == []
    assert codec.class_names(np.zeros((1, 2)), ["a", "b"]) == ["a", "b"]


def test_response_framing_of_the_wire_module_equals_the_protobuf_runtime():
    """Host side of seldon_core_b200/wire.py (no GPU): message parsing, metric and tag-value encoding against the
    run-time built SeldonMessage classes of oracle/seldon_pb.py (proto/prediction.proto:14-82)."""
    import numpy as np

    from oracle import seldon_pb
    from seldon_core_b200 import wire
    from seldon_core_b200.errors import SeldonB200Error

    C = seldon_pb.classes()
    X = np.arange(12.0).reshape(3, 4)
    for kind in ("tensor", "ndarray"):
        req = seldon_pb.request_message(X, kind, ["a", "b", "c", "d"], puid="p")
        raw = req.SerializeToString()
        info = wire.parse_seldon_message(raw)
        assert info["puid"] == "p" and info["names"] == ["a", "b", "c", "d"] and info["kind"] == kind
        off, ln = info["payload"]
        sub = getattr(req.data, kind).SerializeToString()
        assert raw[off:off + ln] == sub
        off, ln = info["data"]
        assert raw[off:off + ln] == req.data.SerializeToString()
    metrics = [{"type": "GAUGE", "key": "is_outlier", "value": np.float64(0.25)}, {"type": "COUNTER", "key": "c", "value": 3},
               {"type": "TIMER", "key": "t", "value": 0.0}, {"type": "GAUGE", "key": "", "value": -0.0}]
    tags = {"outlier-predictions": [0, 1, 1, 0]}
    want = seldon_pb.construct_response(False, seldon_pb.request_message(X), np.array([0, 1, 0]), tags=tags, metrics=metrics)
    meta = want.meta.SerializeToString()
    kb, vb = b"outlier-predictions", wire.encode_value_host(tags["outlier-predictions"])
    entry = b"\x0a" + bytes([len(kb)]) + kb + b"\x12" + bytes([len(vb)]) + vb
    assert meta == b"\x12" + bytes([len(entry)]) + entry + wire.encode_metrics_host(metrics)
    for v in (1.5, "txt", True, [1, "a", [2.0]], None):
        from google.protobuf import struct_pb2

        ref = struct_pb2.Value()
        if v is None:
            ref.null_value = 0
        elif isinstance(v, list):
            ref.list_value.extend(v)
        elif isinstance(v, bool):
            ref.bool_value = v
        elif isinstance(v, str):
            ref.string_value = v
        else:
            ref.number_value = v
        assert wire.encode_value_host(v) == ref.SerializeToString(), v
    lst = C["SeldonMessageList"](seldonMessages=[seldon_pb.request_message(X), seldon_pb.request_message(X + 1)])
    raw = lst.SerializeToString()
    parts = wire.split_message_list(raw)
    assert [raw[o:o + n] for o, n in parts] == [m.SerializeToString() for m in lst.seldonMessages]
    for bad in (b"\x1a\x05\x12\x03", b"\x1a\xff", C["SeldonMessage"](binData=b"x").SerializeToString()):
        with pytest.raises(SeldonB200Error):
            wire.parse_seldon_message(bad)
