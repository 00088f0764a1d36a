"""SeldonMessage and friends as protobuf classes built at run time, and the reference's response construction.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  ``seldon_core.proto.prediction_pb2`` of the reference cannot be
imported here (it imports tensorflow's tensor_pb2), so the messages of ``proto/prediction.proto`` L14-82 are declared
to the protobuf runtime from their field tables - every field except ``DefaultData.tftensor`` (number 4, a
tensorflow type).  The protobuf runtime then is the oracle for the wire bytes.

``construct_response`` restates ``python/seldon_core/utils.py:410-469`` (meta from tags / metrics / puid through
``json_format.ParseDict`` as in L431-443, payload kind mirrored from the request, names from ``client_class_names`` /
``client_feature_names``) with ``array_to_grpc_datadef`` (L235-278).
"""
import numpy as np

_CLASSES = None


def classes():
    """dict name -> message class: SeldonMessage, DefaultData, Tensor, Meta, Metric, Status, SeldonMessageList, Feedback."""
    global _CLASSES
    if _CLASSES is not None:
        return _CLASSES
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory, struct_pb2  # noqa: F401

    F = descriptor_pb2.FieldDescriptorProto
    fd = descriptor_pb2.FileDescriptorProto(name="sb200_oracle_prediction.proto", package="sb200oracle", syntax="proto3")
    fd.dependency.append("google/protobuf/struct.proto")

    def msg(name):
        return fd.message_type.add(name=name)

    def fld(m, name, number, ftype, label=F.LABEL_OPTIONAL, type_name=None, oneof=None):
        f = m.field.add(name=name, number=number, type=ftype, label=label)
        if type_name:
            f.type_name = type_name
        if oneof is not None:
            f.oneof_index = oneof
        return f

    def map_field(m, name, number, vtype, vtype_name=None):
        entry = m.nested_type.add(name="".join(p.capitalize() for p in name.split("_")) + "Entry")
        entry.options.map_entry = True
        fld(entry, "key", 1, F.TYPE_STRING)
        fld(entry, "value", 2, vtype, type_name=vtype_name)
        fld(m, name, number, F.TYPE_MESSAGE, F.LABEL_REPEATED, ".sb200oracle.%s.%s" % (m.name, entry.name))

    t = msg("Tensor")                                   # prediction.proto:33-36
    fld(t, "shape", 1, F.TYPE_INT32, F.LABEL_REPEATED)
    fld(t, "values", 2, F.TYPE_DOUBLE, F.LABEL_REPEATED)
    d = msg("DefaultData")                              # L24-31
    d.oneof_decl.add(name="data_oneof")
    fld(d, "names", 1, F.TYPE_STRING, F.LABEL_REPEATED)
    fld(d, "tensor", 2, F.TYPE_MESSAGE, type_name=".sb200oracle.Tensor", oneof=0)
    fld(d, "ndarray", 3, F.TYPE_MESSAGE, type_name=".google.protobuf.ListValue", oneof=0)
    mt = msg("Metric")                                  # L46-56
    e = mt.enum_type.add(name="MetricType")
    for i, n in enumerate(("COUNTER", "GAUGE", "TIMER")):
        e.value.add(name=n, number=i)
    fld(mt, "key", 1, F.TYPE_STRING)
    fld(mt, "type", 2, F.TYPE_ENUM, type_name=".sb200oracle.Metric.MetricType")
    fld(mt, "value", 3, F.TYPE_FLOAT)
    map_field(mt, "tags", 4, F.TYPE_STRING)
    me = msg("Meta")                                    # L38-44
    fld(me, "puid", 1, F.TYPE_STRING)
    map_field(me, "tags", 2, F.TYPE_MESSAGE, ".google.protobuf.Value")
    map_field(me, "routing", 3, F.TYPE_INT32)
    map_field(me, "requestPath", 4, F.TYPE_STRING)
    fld(me, "metrics", 5, F.TYPE_MESSAGE, F.LABEL_REPEATED, ".sb200oracle.Metric")
    s = msg("Status")                                   # L64-75
    e = s.enum_type.add(name="StatusFlag")
    e.value.add(name="SUCCESS", number=0)
    e.value.add(name="FAILURE", number=1)
    fld(s, "code", 1, F.TYPE_INT32)
    fld(s, "info", 2, F.TYPE_STRING)
    fld(s, "reason", 3, F.TYPE_STRING)
    fld(s, "status", 4, F.TYPE_ENUM, type_name=".sb200oracle.Status.StatusFlag")
    sm = msg("SeldonMessage")                           # L14-22
    sm.oneof_decl.add(name="data_oneof")
    fld(sm, "status", 1, F.TYPE_MESSAGE, type_name=".sb200oracle.Status")
    fld(sm, "meta", 2, F.TYPE_MESSAGE, type_name=".sb200oracle.Meta")
    fld(sm, "data", 3, F.TYPE_MESSAGE, type_name=".sb200oracle.DefaultData", oneof=0)
    fld(sm, "binData", 4, F.TYPE_BYTES, oneof=0)
    fld(sm, "strData", 5, F.TYPE_STRING, oneof=0)
    fld(sm, "jsonData", 6, F.TYPE_MESSAGE, type_name=".google.protobuf.Value", oneof=0)
    ml = msg("SeldonMessageList")                       # L60-62
    fld(ml, "seldonMessages", 1, F.TYPE_MESSAGE, F.LABEL_REPEATED, ".sb200oracle.SeldonMessage")
    fb = msg("Feedback")                                # L77-82
    fld(fb, "request", 1, F.TYPE_MESSAGE, type_name=".sb200oracle.SeldonMessage")
    fld(fb, "response", 2, F.TYPE_MESSAGE, type_name=".sb200oracle.SeldonMessage")
    fld(fb, "reward", 3, F.TYPE_FLOAT)
    fld(fb, "truth", 4, F.TYPE_MESSAGE, type_name=".sb200oracle.SeldonMessage")

    pool = descriptor_pool.DescriptorPool()
    pool.AddSerializedFile(struct_pb2.DESCRIPTOR.serialized_pb)
    pool.Add(fd) if hasattr(pool, "Add") else pool.AddSerializedFile(fd.SerializeToString())
    _CLASSES = {n: message_factory.GetMessageClass(pool.FindMessageTypeByName("sb200oracle." + n))
                for n in ("SeldonMessage", "DefaultData", "Tensor", "Meta", "Metric", "Status", "SeldonMessageList",
                          "Feedback")}
    return _CLASSES


def array_to_list_value(array, lv=None):
    """utils.py:281-304."""
    from google.protobuf.struct_pb2 import ListValue

    if lv is None:
        lv = ListValue()
    if len(array.shape) == 1:
        lv.extend(array.tolist())  # the runtime of this image wants Python numbers; the doubles written are the same
    else:
        for sub_array in array:
            array_to_list_value(sub_array, lv.add_list())
    return lv


def array_to_grpc_datadef(data_type, array, names=()):
    """utils.py:235-278, tensor and ndarray branches."""
    C = classes()
    if data_type == "tensor":
        return C["DefaultData"](names=list(names), tensor=C["Tensor"](shape=array.shape, values=array.ravel().tolist()))
    return C["DefaultData"](names=list(names), ndarray=array_to_list_value(array))


def request_message(X, kind="tensor", names=(), puid=None):
    C = classes()
    m = C["SeldonMessage"](data=array_to_grpc_datadef(kind, np.asarray(X, dtype=np.float64), names))
    if puid:
        m.meta.puid = puid
    return m


def construct_response(is_request, request, raw_response, tags=None, metrics=None, user_class_names=None,
                       user_feature_names=None):
    """utils.py:410-469 for array results.  `tags` / `metrics` are what client_custom_tags / client_custom_metrics
    returned (L431-436); names per client_feature_names (request flow) or client_class_names (response flow)."""
    from google.protobuf import json_format

    C = classes()
    meta = C["Meta"]()
    meta_json = {}
    if tags:
        meta_json["tags"] = tags
    if metrics:
        meta_json["metrics"] = metrics
    if request.meta.puid:
        meta_json["puid"] = request.meta.puid
    json_format.ParseDict(meta_json, meta)
    raw = np.array(raw_response)
    if is_request:
        names = user_feature_names if user_feature_names is not None else list(request.data.names)
    elif raw.ndim > 1:
        names = user_class_names if user_class_names is not None else ["t:{}".format(i) for i in range(raw.shape[1])]
    else:
        names = []
    if request.WhichOneof("data_oneof") == "data":
        kind = request.data.WhichOneof("data_oneof") if np.issubdtype(raw.dtype, np.number) else "ndarray"
    else:
        kind = "tensor" if np.issubdtype(raw.dtype, np.number) else "ndarray"
    return C["SeldonMessage"](data=array_to_grpc_datadef(kind, raw, names), meta=meta)
