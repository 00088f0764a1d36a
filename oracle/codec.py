"""CPU oracle for the SeldonMessage numeric payload codec (tensor / ndarray) and response naming.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  Restates, from
``python/seldon_core/utils.py`` of the reference:

* ``grpc_datadef_to_array`` tensor branch L161-169 - the packed doubles are the
  tail ``sz*8`` bytes of the serialised ``Tensor``; ndarray branch L174-176;
* ``array_to_grpc_datadef`` L235-278 / ``array_to_list_value`` L281-304 - encode;
* ``extract_request_parts_json`` L499-505 and ``construct_response_json``
  L345-384 - the REST (dict) twins;
* ``client_class_names`` naming rule, ``python/seldon_core/user_model.py:L103-131``.

``seldon_core.utils`` itself cannot be imported here (hard imports of tensorflow
and flask, ``utils.py:6,9``), so the wire bytes are pinned against the protobuf
runtime instead: ``google.protobuf.struct_pb2.ListValue`` for ndarray and a
``Tensor`` message built at run time from ``proto/prediction.proto:L33-38``
(``tests/test_oracle_codec.py``), plus the known answers of
``python/tests/test_utils.py`` (L53-71, L110-165, L342-380).

Wire layout (proto3, little-endian):
  Tensor    : 0a <varint n> <shape varints...>  12 <varint 8*count> <count x f64>
  ListValue : per row  0a <varint L1> 32 <varint L2> ( 0a 09 11 <f64> ) x cols,
              L2 = 11*cols, L1 = 1 + len(varint L2) + L2;   1-D: ( 0a 09 11 <f64> ) x n
"""
import numpy as np


def varint(n):
    out = bytearray()
    n = int(n)
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def read_varint(buf, pos):
    shift = 0
    val = 0
    while True:
        b = buf[pos]
        pos += 1
        val |= (b & 0x7F) << shift
        if not b & 0x80:
            return val, pos
        shift += 7


# ----------------------------------------------------------------------------- Tensor

def encode_tensor(arr):
    """bytes of ``Tensor(shape=arr.shape, values=arr.ravel().tolist())`` (utils.py:255-259)."""
    a = np.ascontiguousarray(arr, dtype="<f8")
    shape = b"".join(varint(s) for s in a.shape)
    head = b"\x0a" + varint(len(shape)) + shape if a.ndim else b""
    body = a.tobytes()
    if len(body) == 0:
        return head
    return head + b"\x12" + varint(len(body)) + body


def tensor_values_offset(shape):
    """Byte offset of the first packed double inside ``encode_tensor`` output."""
    shape_bytes = sum(len(varint(s)) for s in shape)
    count = int(np.prod(shape))
    return 1 + len(varint(shape_bytes)) + shape_bytes + 1 + len(varint(8 * count))


def decode_tensor(buf):
    """utils.py:161-169: parse ``shape``, then view the LAST sz*8 bytes as float64."""
    buf = bytes(buf)
    pos = 0
    shape = []
    if buf[:1] == b"\x0a":
        n, pos = read_varint(buf, 1)
        end = pos + n
        while pos < end:
            v, pos = read_varint(buf, pos)
            shape.append(v)
    sz = int(np.prod(shape)) if shape else 0
    vals = np.frombuffer(buf[len(buf) - sz * 8:], dtype="<f8", count=sz) if sz else np.zeros(0)
    return vals.reshape(shape)


# ----------------------------------------------------------------------------- ListValue (ndarray)

def listvalue_row_layout(cols):
    """(row_bytes, header_bytes) of one inner row of a 2-D ndarray payload."""
    l2 = 11 * cols
    l1 = 1 + len(varint(l2)) + l2
    hdr = 1 + len(varint(l1)) + 1 + len(varint(l2))
    return hdr + l2, hdr


def encode_listvalue(arr):
    """bytes of ``array_to_list_value(arr)`` (utils.py:281-304) for 1-D / 2-D numeric arrays."""
    a = np.ascontiguousarray(arr, dtype="<f8")
    if a.ndim == 1:
        rec = np.empty((a.shape[0], 11), dtype=np.uint8)
        rec[:, 0], rec[:, 1], rec[:, 2] = 0x0A, 0x09, 0x11
        rec[:, 3:] = a.view(np.uint8).reshape(-1, 8)
        return rec.tobytes()
    rows, cols = a.shape
    row_bytes, hdr = listvalue_row_layout(cols)
    l2 = 11 * cols
    l1 = 1 + len(varint(l2)) + l2
    header = np.frombuffer(b"\x0a" + varint(l1) + b"\x32" + varint(l2), dtype=np.uint8)
    out = np.empty((rows, row_bytes), dtype=np.uint8)
    out[:, :hdr] = header
    rec = out[:, hdr:].reshape(rows, cols, 11)
    rec[:, :, 0], rec[:, :, 1], rec[:, :, 2] = 0x0A, 0x09, 0x11
    rec[:, :, 3:] = a.view(np.uint8).reshape(rows, cols, 8)
    return out.tobytes()


def decode_listvalue(buf):
    """General (slow) parser of a numeric ListValue: the ``MessageToDict`` + ``np.array`` of utils.py:174-176."""
    buf = bytes(buf)

    def parse_list(lo, hi):
        items = []
        pos = lo
        while pos < hi:
            assert buf[pos] == 0x0A, "ListValue.values tag"
            n, pos = read_varint(buf, pos + 1)
            end = pos + n
            if buf[pos] == 0x11:  # number_value
                items.append(float(np.frombuffer(buf[pos + 1:pos + 9], dtype="<f8")[0]))
            elif buf[pos] == 0x32:  # list_value
                m, p2 = read_varint(buf, pos + 1)
                items.append(parse_list(p2, p2 + m))
            else:
                raise ValueError("unsupported Value kind 0x%02x" % buf[pos])
            pos = end
        return items

    return np.array(parse_list(0, len(buf)))


# ----------------------------------------------------------------------------- REST dict twins + naming

def extract_json(request):
    """extract_request_parts_json, numeric branches (utils.py:495-505)."""
    d = request["data"]
    if "tensor" in d:
        return np.array(d["tensor"]["values"]).reshape(d["tensor"]["shape"]), "tensor"
    if "ndarray" in d:
        return np.array(d["ndarray"]), "ndarray"
    return np.array([]), None


def response_datadef_json(request, result):
    """construct_response_json data part for numeric results (utils.py:345-384): mirror the request type."""
    result = np.asarray(result)
    if "data" in request and "tensor" not in request["data"]:
        return {"ndarray": result.tolist()}
    return {"tensor": {"values": result.ravel().tolist(), "shape": result.shape}}


def class_names(result, user_names=None):
    """user_model.py:117-131: user class_names() if any, else t:i for 2-D, [] for 1-D."""
    result = np.asarray(result)
    if result.ndim > 1:
        return list(user_names) if user_names is not None else ["t:%d" % i for i in range(result.shape[1])]
    return []
