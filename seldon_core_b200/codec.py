"""SeldonMessage numeric payloads <-> device tensors.

Host-side mirror of the numeric branches of ``python/seldon_core/utils.py``: ``grpc_datadef_to_array`` (L147-181),
``array_to_grpc_datadef`` / ``array_to_list_value`` (L235-304) and the naming rule of ``client_class_names``
(``python/seldon_core/user_model.py:103-131``).  Only the small varint headers are handled on the host; the bulk
(the packed doubles / the 11-byte ListValue records) is converted by libsb200.so on the GPU.
"""
import re
import threading

import numpy as np

from . import _cabi
from .errors import SeldonB200Error


def _varint(n):
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _read_varint(buf, pos):
    shift = val = 0
    while True:
        b = buf[pos]
        pos += 1
        val |= (b & 0x7F) << shift
        if not b & 0x80:
            return val, pos
        shift += 7


def _device(device):
    import torch

    if not torch.cuda.is_available():
        raise SeldonB200Error("no CUDA device: seldon_core_b200 has no CPU fallback", status_code=500)
    return torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())


_STAGING = {}  # device -> [pinned uint8 tensor, event of the last copy out of it]
_STAGING_LOCK = threading.Lock()


def _upload(wire, dev, offset_to_align):
    """Copy wire bytes to the device so that byte `offset_to_align` lands on a 16-byte boundary.  The bytes go through
    one pinned staging buffer per device that grows to the largest payload seen and is reused (pinning memory per
    request is a device-synchronising cudaHostAlloc); the copy out of it is awaited before the next payload is staged."""
    import torch

    pad = (-offset_to_align) % 16
    n = pad + len(wire)
    with _STAGING_LOCK:
        ent = _STAGING.get(dev)
        if ent is None or ent[0].numel() < n:
            if ent is not None:
                ent[1].synchronize()
            ent = [torch.empty(max(n, 1 << 16) * 5 // 4, dtype=torch.uint8).pin_memory(), torch.cuda.Event()]
            _STAGING[dev] = ent
        else:
            ent[1].synchronize()
        ent[0][pad:n].numpy()[:] = np.frombuffer(wire, dtype=np.uint8)
        with torch.cuda.device(dev):
            out = ent[0][:n].to(dev, non_blocking=True)
            ent[1].record(torch.cuda.current_stream(dev))
    return out, pad


def parse_tensor_header(wire):
    """Shape and offset of the packed doubles of a serialised Tensor (proto/prediction.proto:33-38):
    `0a <len> <varint dims>` then `12 <len> <little-endian doubles>`, the values last (utils.py:164-169 reads the
    tail).  A payload whose values field is absent, is not the tail, or does not hold prod(shape) doubles raises -
    the reference fails on the reshape (utils.py:169)."""
    pos, shape, n = 0, [], len(wire)
    try:
        if wire[:1] == b"\x0a":
            ln, pos = _read_varint(wire, 1)
            end = pos + ln
            if end > n:
                raise IndexError
            while pos < end:
                v, pos = _read_varint(wire, pos)
                shape.append(v)
        count = 1
        for v in shape:
            count *= int(v)
        if not shape:
            count = 0
        if pos == n and count == 0:
            return shape, n
        if wire[pos:pos + 1] != b"\x12":
            raise SeldonB200Error("tensor payload: no packed `values` field after the shape")
        vlen, pos = _read_varint(wire, pos + 1)
    except IndexError:
        raise SeldonB200Error("tensor payload is truncated")
    if vlen != 8 * count or pos + vlen != n:
        raise SeldonB200Error("tensor payload: shape %s needs %d bytes of values, the message carries %d"
                              % (shape, 8 * count, n - pos))
    return shape, pos


def decode_tensor(wire, device=None, mu=None, inv_sigma=None):
    """Serialised Tensor -> CUDA float32 (rows, cols); optional fused standardise (device float32 vectors)."""
    import torch

    dev = _device(device)
    shape, off = parse_tensor_header(wire)
    if len(shape) != 2:
        raise SeldonB200Error("tensor payload must be 2-D, got shape %s" % (shape,))
    rows, cols = shape
    buf, pad = _upload(wire, dev, off)
    out = torch.empty((rows, cols), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        _cabi.check(_cabi.lib().sb_unpack_tensor_f64(buf.data_ptr(), buf.numel(), pad + off, rows, cols,
                                                     None if mu is None else mu.data_ptr(),
                                                     None if inv_sigma is None else inv_sigma.data_ptr(),
                                                     out.data_ptr(), cols, torch.cuda.current_stream().cuda_stream))
    return out


def decode_ndarray(wire, cols, device=None):
    """Serialised 2-D numeric ListValue -> CUDA float32 (rows, cols).  `cols` comes from the first row header or
    from the request's `names`; a ragged or non-numeric payload raises."""
    import torch

    dev = _device(device)
    L = _cabi.lib()
    row_bytes = L.sb_listvalue_nbytes(1, cols)
    if row_bytes == 0 or len(wire) % row_bytes:
        raise SeldonB200Error("ndarray payload of %d bytes is not a whole number of %d-column rows" % (len(wire), cols))
    rows = len(wire) // row_bytes
    buf, pad = _upload(wire, dev, 0)
    out = torch.empty((rows, cols), dtype=torch.float32, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        _cabi.check(L.sb_unpack_listvalue(buf.data_ptr() + pad, len(wire), rows, cols, out.data_ptr(), cols,
                                          status.data_ptr(), torch.cuda.current_stream().cuda_stream))
    if int(status.item()) != 0:
        raise SeldonB200Error("malformed ndarray payload (not a rows x %d list of numbers)" % cols)
    return out


def ndarray_cols(wire):
    """Number of columns announced by the first row header of a 2-D ListValue."""
    if wire[:1] != b"\x0a":
        raise SeldonB200Error("not a ListValue")
    _, pos = _read_varint(wire, 1)
    if wire[pos:pos + 1] != b"\x32":
        raise SeldonB200Error("not a 2-D ListValue")
    l2, _ = _read_varint(wire, pos + 1)
    if l2 % 11:
        raise SeldonB200Error("not a numeric ListValue")
    return l2 // 11


def _dtype_code(t):
    import torch

    if t.dtype == torch.float32:
        return _cabi.SB_F32
    if t.dtype == torch.int32:
        return _cabi.SB_I32
    raise SeldonB200Error("encode expects float32 or int32 device tensors")


def encode_tensor(t):
    """CUDA float32/int32 tensor -> bytes of Tensor(shape, values) (utils.py:255-259)."""
    import torch

    t = t.contiguous()
    shape = b"".join(_varint(int(s)) for s in t.shape)
    head = b"\x0a" + _varint(len(shape)) + shape + b"\x12" + _varint(8 * t.numel())
    vals = torch.empty(8 * t.numel(), dtype=torch.uint8, device=t.device)
    with torch.cuda.device(t.device):
        _cabi.check(_cabi.lib().sb_pack_tensor_f64(t.data_ptr(), _dtype_code(t), t.numel(), vals.data_ptr(),
                                                   torch.cuda.current_stream().cuda_stream))
    return head + vals.cpu().numpy().tobytes()


def encode_ndarray(t):
    """CUDA float32/int32 tensor (1-D or 2-D) -> bytes of the ListValue of array_to_list_value (utils.py:281-304)."""
    import torch

    t = t.contiguous()
    if t.dim() == 1:
        rows, cols = t.shape[0], 0
    elif t.dim() == 2:
        rows, cols = t.shape
    else:
        raise SeldonB200Error("encode_ndarray supports 1-D and 2-D tensors")
    L = _cabi.lib()
    nbytes = L.sb_listvalue_nbytes(rows, cols)
    out = torch.empty(int(nbytes), dtype=torch.uint8, device=t.device)
    with torch.cuda.device(t.device):
        _cabi.check(L.sb_pack_listvalue(t.data_ptr(), _dtype_code(t), rows, cols, out.data_ptr(),
                                        torch.cuda.current_stream().cuda_stream))
    return out.cpu().numpy().tobytes()


def class_names(result_ndim, n_targets, user_names=None):
    """user_model.py:117-131."""
    if result_ndim > 1:
        return list(user_names) if user_names is not None else ["t:{}".format(i) for i in range(n_targets)]
    return []


# ---------------------------------------------------------------------------------------------- REST JSON
_NDARRAY_KEY = re.compile(rb'"ndarray"\s*:\s*\[')
_VALUES_KEY = re.compile(rb'"values"\s*:\s*\[')
_END_2D = re.compile(rb'\]\s*\]')


def json_ndarray_span(payload, key="ndarray"):
    """(start, end, ndim) of the value of "ndarray" (or of the tensor's "values") in a REST request body
    (`{"data": {"ndarray": [[..], ..]}}`, utils.py:499-505), found without parsing the numbers: start at its opening
    bracket, end one past its closing one."""
    m = (_NDARRAY_KEY if key == "ndarray" else _VALUES_KEY).search(payload)
    if m is None:
        raise SeldonB200Error('no "%s" member in the payload' % key)
    start = m.end() - 1
    i = start + 1
    while i < len(payload) and payload[i:i + 1] in (b" ", b"\n", b"\r", b"\t"):
        i += 1
    if payload[i:i + 1] == b"[":
        e = _END_2D.search(payload, i)
        if e is None:
            raise SeldonB200Error('unterminated "ndarray" value')
        return start, e.end(), 2
    e = payload.find(b"]", start)
    if e < 0:
        raise SeldonB200Error('unterminated "ndarray" value')
    return start, e + 1, 1


def decode_json_ndarray(payload, device=None, key="ndarray"):
    """REST JSON body (bytes) -> CUDA float32 array holding data.ndarray (or, with key="values", the flat values of
    data.tensor), shape (rows, cols) or (n,) for a flat list.

    Replaces `json.loads` + `np.array(datadef["ndarray"])` (utils.py:499-505): every number gets the correctly rounded
    binary64 value CPython's float() gives it, narrowed to fp32.  Non-numeric or ragged arrays raise."""
    import torch

    if isinstance(payload, str):
        payload = payload.encode("utf-8")
    start, end, ndim = json_ndarray_span(payload, key)
    text = payload[start:end]
    dev = _device(device)
    L = _cabi.lib()
    buf, _pad = _upload(text, dev, 0)
    ws = torch.empty((int(L.sb_json_workspace_bytes(len(text))) + 7) // 8, dtype=torch.int64, device=dev)
    counts = torch.zeros(4, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream().cuda_stream
        _cabi.check(L.sb_json_ndarray_count(buf.data_ptr(), len(text), ws.data_ptr(), ws.numel() * 8, counts.data_ptr(),
                                            stream))
        numbers, nopen, nclose, nbad = (int(v) for v in counts.cpu())
        if nbad or nopen != nclose or nopen != (1 if ndim == 1 else nopen) or nopen < ndim:
            raise SeldonB200Error('"ndarray" is not an array of numbers')
        rows = 1 if ndim == 1 else nopen - 1
        if rows == 0 or numbers % max(rows, 1):
            if numbers == 0:
                return torch.empty((0,) if ndim == 1 or rows == 0 else (rows, 0), dtype=torch.float32, device=dev)
            raise SeldonB200Error('"ndarray" rows have different lengths')
        cols = numbers // rows
        out = torch.empty((rows, cols), dtype=torch.float32, device=dev)
        status = torch.zeros(1, dtype=torch.int32, device=dev)
        _cabi.check(L.sb_unpack_json_ndarray(buf.data_ptr(), len(text), rows, cols, ws.data_ptr(), out.data_ptr(), cols,
                                             status.data_ptr(), stream))
        st = int(status.item())
    if st == 2:
        raise SeldonB200Error('"ndarray" holds a number with more than 19 significant digits that needs arbitrary '
                              'precision to round; not supported')
    if st != 0:
        raise SeldonB200Error('"ndarray" is not a rectangular array of JSON numbers')
    return out.view(-1) if ndim == 1 else out


def encode_json_ndarray_device(t):
    """CUDA float32 / int32 tensor (1-D or 2-D) -> (CUDA uint8 tensor holding the text json.dumps(t.tolist()) writes,
    its length as a CUDA int64 scalar).  No host synchronisation."""
    import torch

    t = t.contiguous()
    if t.dim() == 1:
        rows, cols = t.shape[0], 0
    elif t.dim() == 2:
        rows, cols = t.shape
        if rows and not cols:
            raise SeldonB200Error("rows of zero columns are not supported")
    else:
        raise SeldonB200Error("encode_json_ndarray supports 1-D and 2-D tensors")
    L = _cabi.lib()
    text = torch.empty(int(L.sb_json_text_capacity(rows, cols)), dtype=torch.uint8, device=t.device)
    ws = torch.empty((int(L.sb_json_encode_workspace_bytes(rows, cols)) + 7) // 8, dtype=torch.int64, device=t.device)
    nbytes = torch.zeros(1, dtype=torch.int64, device=t.device)
    with torch.cuda.device(t.device):
        _cabi.check(L.sb_pack_json_ndarray(t.data_ptr(), _dtype_code(t), rows, cols, ws.data_ptr(), ws.numel() * 8,
                                           text.data_ptr(), nbytes.data_ptr(), torch.cuda.current_stream().cuda_stream))
    return text, nbytes


def encode_json_ndarray(t):
    """CUDA float32 / int32 tensor -> bytes: the value of "ndarray" (or of tensor "values" for a flat tensor) exactly
    as `json.dumps(array.tolist())` of the reference's response path writes it (utils.py:345-384)."""
    text, nbytes = encode_json_ndarray_device(t)
    n = int(nbytes.item())
    return text[:n].cpu().numpy().tobytes()
