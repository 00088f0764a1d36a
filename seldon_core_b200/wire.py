"""Serialised SeldonMessage in -> serialised SeldonMessage out, with the payload decoded and encoded on the GPU.

This is what the `*_raw` hooks of the drop-in classes run (`predict_raw`, `transform_input_raw`, `aggregate_raw`):
the reference's wrapper offers every request to `user_model.<method>_raw(request)` before it extracts anything
(python/seldon_core/seldon_methods.py:43-47, 124-128, 268-272; python/seldon_core/user_model.py:36-78), and what it
would otherwise do around the user method is restated here for numeric `data.tensor` / `data.ndarray` payloads:

* request  : `extract_request_parts` + `grpc_datadef_to_array` (utils.py:147-181, 472-525)  -> `parse_seldon_message`
             finds the payload inside the message; `WireStage.decode_matrix` moves it to the device in chunks (pinned
             host -> a two-slot device ring on a copy stream) and unpacks each chunk with `sb_unpack_tensor_f64` /
             `sb_unpack_listvalue` while the next one is in flight;
* response : `construct_response` + `array_to_grpc_datadef` (utils.py:235-304, 410-469): same payload kind as the
             request, `names` per `client_class_names` / `client_feature_names` (user_model.py:103-131, 230-255), `meta`
             with puid, tags and metrics; the values are packed on the device (`sb_pack_tensor_f64`,
             `sb_pack_listvalue`) and land directly behind the host-written headers in a pinned response buffer.

Only the few header bytes of the protobuf framing are handled on the host (SURVEY App. C gives the layouts).
The buffers (pinned request staging, device ring, pinned response) belong to the component instance and are reused
from call to call: nothing is pinned or allocated per request once they have grown to the request size.
"""
import struct

import numpy as np

from . import _cabi
from .codec import _read_varint, _varint
from .errors import SeldonB200Error

_METRIC_TYPES = {"COUNTER": 0, "GAUGE": 1, "TIMER": 2}


class HostBuf(object):
    """A byte string in pinned host memory (a view of a reusable buffer: copy it with tobytes() to keep it)."""

    __slots__ = ("tensor", "offset", "nbytes")

    def __init__(self, tensor, offset, nbytes):
        self.tensor, self.offset, self.nbytes = tensor, int(offset), int(nbytes)

    def view(self):
        return self.tensor[self.offset:self.offset + self.nbytes]

    def numpy(self):
        return self.view().numpy()

    def tobytes(self):
        return self.numpy().tobytes()

    def __len__(self):
        return self.nbytes


def _mv(buf):
    if isinstance(buf, HostBuf):
        return memoryview(buf.numpy())
    return memoryview(buf).cast("B") if not isinstance(buf, memoryview) else buf.cast("B")


def _fields(mv, lo, hi):
    """(field, wire type, value or (offset, length)) of every field in mv[lo:hi]; raises on truncation."""
    pos = lo
    try:
        while pos < hi:
            tag, pos = _read_varint(mv, pos)
            field, wt = tag >> 3, tag & 7
            if wt == 2:
                ln, pos = _read_varint(mv, pos)
                if pos + ln > hi:
                    raise IndexError
                yield field, wt, (pos, ln)
                pos += ln
            elif wt == 0:
                v, pos = _read_varint(mv, pos)
                yield field, wt, v
            elif wt == 1:
                yield field, wt, (pos, 8)
                pos += 8
            elif wt == 5:
                yield field, wt, (pos, 4)
                pos += 4
            else:
                raise SeldonB200Error("unsupported protobuf wire type %d" % wt)
        if pos != hi:
            raise IndexError
    except IndexError:
        raise SeldonB200Error("truncated SeldonMessage")


def parse_seldon_message(buf, lo=0, hi=None):
    """Where things are inside a serialised SeldonMessage (proto/prediction.proto:14-38): puid, names, the kind of
    numeric payload (`tensor` | `ndarray`) and the byte range of that payload and of the whole `data` field."""
    mv = _mv(buf)
    hi = len(mv) if hi is None else hi
    info = {"puid": None, "names": [], "kind": None, "payload": None, "data": None}
    for field, wt, v in _fields(mv, lo, hi):
        if field == 2 and wt == 2:  # Meta
            for f2, w2, v2 in _fields(mv, v[0], v[0] + v[1]):
                if f2 == 1 and w2 == 2:
                    info["puid"] = bytes(mv[v2[0]:v2[0] + v2[1]]).decode("utf-8")
        elif field == 3 and wt == 2:  # DefaultData
            info["data"] = v
            for f2, w2, v2 in _fields(mv, v[0], v[0] + v[1]):
                if f2 == 1 and w2 == 2:
                    info["names"].append(bytes(mv[v2[0]:v2[0] + v2[1]]).decode("utf-8"))
                elif f2 == 2 and w2 == 2:
                    info["kind"], info["payload"] = "tensor", v2
                elif f2 == 3 and w2 == 2:
                    info["kind"], info["payload"] = "ndarray", v2
                elif f2 == 4:
                    raise SeldonB200Error("tftensor payloads are not handled by the device codec")
        elif field in (4, 5, 6):
            raise SeldonB200Error("binData / strData / jsonData payloads are not handled by the device codec")
    if info["kind"] is None:
        raise SeldonB200Error("the message carries no data.tensor / data.ndarray payload")
    return info


def split_message_list(buf):
    """Byte ranges of the messages of a serialised SeldonMessageList (proto/prediction.proto:60-62)."""
    mv = _mv(buf)
    return [v for field, wt, v in _fields(mv, 0, len(mv)) if field == 1 and wt == 2]


def _tensor_layout(mv, off, ln):
    """Shape and absolute offset of the packed doubles of the Tensor at mv[off:off+ln] (codec.parse_tensor_header)."""
    from .codec import parse_tensor_header

    head = bytes(mv[off:off + min(ln, 64)])
    # the header parser needs the total length, not the bytes of the values: give it a stand-in of the right size
    shape, rel = parse_tensor_header(_Sized(head, ln))
    return shape, off + rel


class _Sized(object):
    """The first bytes of a long byte string that reports the full length (header parsing without copying)."""

    def __init__(self, head, n):
        self.head, self.n = head, n

    def __len__(self):
        return self.n

    def __getitem__(self, i):
        return self.head[i]


# ------------------------------------------------------------------------------------------------ response framing
def _ld(tag, body_len):
    return bytes([tag]) + _varint(body_len)


def encode_value_host(v):
    """google.protobuf.Value for the tag values the components produce (numbers, strings, bools, lists)."""
    if isinstance(v, bool):
        return b"\x20" + (b"\x01" if v else b"\x00")
    if isinstance(v, (int, float, np.integer, np.floating)):
        return b"\x11" + struct.pack("<d", float(v))
    if isinstance(v, str):
        s = v.encode("utf-8")
        return _ld(0x1A, len(s)) + s
    if isinstance(v, (list, tuple, np.ndarray)):
        try:
            a = np.asarray(v)
        except ValueError:  # ragged / mixed list
            a = np.empty(0, dtype=object)
        if a.ndim == 1 and a.dtype.kind in "iuf" and not isinstance(v, np.ndarray) and any(isinstance(x, bool) for x in v):
            a = np.empty(0, dtype=object)
        if a.ndim == 1 and a.dtype.kind in "iuf":  # the common case: one record of 11 bytes per number
            rec = np.empty((a.shape[0], 11), dtype=np.uint8)
            rec[:, 0], rec[:, 1], rec[:, 2] = 0x0A, 0x09, 0x11
            rec[:, 3:] = np.ascontiguousarray(a, dtype="<f8").view(np.uint8).reshape(-1, 8)
            body = rec.tobytes()
        else:
            body = b"".join(_ld(0x0A, len(e)) + e for e in (encode_value_host(x) for x in v))
        return _ld(0x32, len(body)) + body
    if v is None:
        return b"\x08\x00"
    raise SeldonB200Error("cannot encode a tag value of type %s" % type(v).__name__)


def encode_metrics_host(metrics):
    """repeated Metric metrics = 5 of Meta (proto/prediction.proto:40-58): what json_format.ParseDict makes of the
    dicts returned by metrics() (utils.py:438-442); `value` is a float32 on the wire."""
    out = []
    for m in metrics or []:
        body = b""
        key = str(m.get("key", "")).encode("utf-8")
        if key:
            body += _ld(0x0A, len(key)) + key
        t = _METRIC_TYPES[m["type"]]
        if t:
            body += b"\x10" + bytes([t])
        val = struct.pack("<f", np.float32(m["value"]))
        if val != b"\x00\x00\x00\x00":
            body += b"\x1d" + val
        for k, v in (m.get("tags") or {}).items():
            kb, vb = str(k).encode("utf-8"), str(v).encode("utf-8")
            e = _ld(0x0A, len(kb)) + kb + _ld(0x12, len(vb)) + vb
            body += _ld(0x22, len(e)) + e
        out.append(_ld(0x2A, len(body)) + body)
    return b"".join(out)


class Segments(object):
    """A byte string under construction: host byte strings and device byte tensors, in order, with nested
    length-delimited fields closed from the inside out."""

    def __init__(self):
        self.parts = []  # bytes | (device uint8 tensor)
        self.n = 0

    def add(self, part):
        """part: bytes / bytearray / numpy uint8 array (host) or a CUDA uint8 tensor (device)"""
        n = len(part) if isinstance(part, (bytes, bytearray, np.ndarray)) else int(part.numel())
        if n:
            self.parts.append(part)
            self.n += n
        return self

    def extend(self, other):
        self.parts.extend(other.parts)
        self.n += other.n
        return self

    def wrapped(self, tag):
        """This string as the body of a length-delimited field `tag`."""
        out = Segments().add(_ld(tag, self.n))
        return out.extend(self)


class WireStage(object):
    """Reusable buffers and the copy stream of one component instance."""

    def __init__(self, device, chunk_bytes=128 << 20):
        self.dev = device
        self.chunk_bytes = int(chunk_bytes)
        self.ring = None
        self.copy_stream = None
        self.ev_copied = None
        self.ev_free = None
        self.pinned_in = None
        self.pinned_out = None

    # -- buffers
    def _setup(self):
        import torch

        if self.ring is None:
            with torch.cuda.device(self.dev):
                self.ring = [torch.empty(self.chunk_bytes + 64, dtype=torch.uint8, device=self.dev) for _ in range(2)]
                self.copy_stream = torch.cuda.Stream(device=self.dev)
                self.ev_copied = [torch.cuda.Event() for _ in range(2)]
                self.ev_free = [torch.cuda.Event() for _ in range(2)]
                for e in self.ev_free:
                    e.record(torch.cuda.current_stream(self.dev))

    @staticmethod
    def _pinned(n):
        import torch

        return torch.empty(int(n), dtype=torch.uint8).pin_memory()

    def host_bytes(self, msg):
        """The message as a uint8 CPU tensor, pinned when this stage had to copy it (bytes objects are pageable)."""
        import torch

        if isinstance(msg, HostBuf):
            return msg.view()
        if isinstance(msg, torch.Tensor):
            if msg.dtype != torch.uint8 or msg.is_cuda or msg.dim() != 1:
                raise SeldonB200Error("a raw message tensor must be a 1-D uint8 CPU tensor")
            return msg
        if hasattr(msg, "SerializeToString"):
            msg = msg.SerializeToString()
        a = np.frombuffer(msg, dtype=np.uint8)
        if self.copy_stream is not None:
            self.copy_stream.synchronize()  # an earlier message may still be on its way out of the staging buffer
        if self.pinned_in is None or self.pinned_in.numel() < a.size:
            self.pinned_in = self._pinned(max(a.size, 1 << 16) * 5 // 4)
        self.pinned_in[:a.size].numpy()[:] = a
        return self.pinned_in[:a.size]

    # -- request payload -> device fp32 matrix
    def decode_matrix(self, host, info, mu=None, inv_sigma=None):
        """host: uint8 CPU tensor holding the message; info: parse_seldon_message(host).  Returns a CUDA float32
        (rows, cols) tensor; the H2D copies run on the stage's copy stream, chunk by chunk through a two-slot ring, the
        unpack kernels on the current stream."""
        import torch

        self._setup()
        L = _cabi.lib()
        mv = memoryview(host.numpy())
        off, ln = info["payload"]
        if info["kind"] == "tensor":
            shape, voff = _tensor_layout(mv, off, ln)
            if len(shape) != 2:
                raise SeldonB200Error("tensor payload must be 2-D, got shape %s" % (shape,))
            rows, cols = int(shape[0]), int(shape[1])
            row_bytes, align_rows = 8 * cols, 1
        else:
            from .codec import ndarray_cols

            cols = ndarray_cols(bytes(mv[off:off + min(ln, 32)]))
            row_bytes = int(L.sb_listvalue_nbytes(1, cols))
            if cols < 1 or ln % row_bytes:
                raise SeldonB200Error("ndarray payload of %d bytes is not a whole number of %d-column rows" % (ln, cols))
            rows, voff, align_rows = ln // row_bytes, off, 16
        out = torch.empty((rows, cols), dtype=torch.float32, device=self.dev)
        if rows == 0:
            return out
        rpc = max(align_rows, self.chunk_bytes // row_bytes // align_rows * align_rows)
        if row_bytes * align_rows > self.chunk_bytes:
            raise SeldonB200Error("a row of %d bytes does not fit the staging ring" % row_bytes)
        status = torch.zeros(1, dtype=torch.int32, device=self.dev) if info["kind"] == "ndarray" else None
        bad = torch.zeros(1, dtype=torch.int32, device=self.dev) if status is not None else None
        with torch.cuda.device(self.dev):
            cur = torch.cuda.current_stream(self.dev)
            for i, r0 in enumerate(range(0, rows, rpc)):
                slot, nr = i & 1, min(rpc, rows - r0)
                nb = nr * row_bytes
                src = host[voff + r0 * row_bytes: voff + r0 * row_bytes + nb]
                self.copy_stream.wait_event(self.ev_free[slot])
                with torch.cuda.stream(self.copy_stream):
                    self.ring[slot][:nb].copy_(src, non_blocking=True)
                    self.ev_copied[slot].record(self.copy_stream)
                cur.wait_event(self.ev_copied[slot])
                dst = out.data_ptr() + 4 * r0 * cols
                if info["kind"] == "tensor":
                    _cabi.check(L.sb_unpack_tensor_f64(self.ring[slot].data_ptr(), nb, 0, nr, cols,
                                                       None if mu is None else mu.data_ptr(),
                                                       None if inv_sigma is None else inv_sigma.data_ptr(),
                                                       dst, cols, cur.cuda_stream))
                else:
                    _cabi.check(L.sb_unpack_listvalue(self.ring[slot].data_ptr(), nb, nr, cols, dst, cols,
                                                      status.data_ptr(), cur.cuda_stream))
                    bad.add_(status)
                self.ev_free[slot].record(cur)
        if bad is not None and int(bad.item()) != 0:
            raise SeldonB200Error("malformed ndarray payload (not a rows x %d list of numbers)" % cols)
        return out

    # -- result -> response payload segments (values packed on the device)
    def pack_payload(self, kind, t):
        """Segments of a Tensor (kind 'tensor') or ListValue ('ndarray') holding the device tensor t (fp32 / int32,
        1-D or 2-D), as array_to_grpc_datadef writes them (utils.py:255-278)."""
        import torch

        L = _cabi.lib()
        t = t.contiguous()
        code = _cabi.SB_F32 if t.dtype == torch.float32 else _cabi.SB_I32
        if t.dtype not in (torch.float32, torch.int32):
            raise SeldonB200Error("device results must be float32 or int32")
        with torch.cuda.device(self.dev):
            st = torch.cuda.current_stream(self.dev).cuda_stream
            if kind == "tensor":
                shape = b"".join(_varint(int(s)) for s in t.shape)
                vals = torch.empty(8 * t.numel(), dtype=torch.uint8, device=self.dev)
                if t.numel():
                    _cabi.check(L.sb_pack_tensor_f64(t.data_ptr(), code, t.numel(), vals.data_ptr(), st))
                seg = Segments().add(_ld(0x0A, len(shape)) + shape)
                if t.numel():
                    seg.add(_ld(0x12, 8 * t.numel())).add(vals)
                return seg
            rows, cols = (t.shape[0], 0) if t.dim() == 1 else (t.shape[0], t.shape[1])
            nbytes = int(L.sb_listvalue_nbytes(rows, cols))
            body = torch.empty(nbytes, dtype=torch.uint8, device=self.dev)
            if nbytes:
                _cabi.check(L.sb_pack_listvalue(t.data_ptr(), code, rows, cols, body.data_ptr(), st))
            return Segments().add(body)

    def finish(self, seg):
        """Lay the segments out in the pinned response buffer (host parts by memcpy, device parts by D2H on the current
        stream), wait for the copies, return the HostBuf."""
        import torch

        if self.pinned_out is None or self.pinned_out.numel() < seg.n:
            self.pinned_out = self._pinned(max(seg.n, 1 << 16) * 5 // 4)
        out_np = self.pinned_out.numpy()
        pos = 0
        with torch.cuda.device(self.dev):
            for p in seg.parts:
                if isinstance(p, (bytes, bytearray, np.ndarray)):
                    out_np[pos:pos + len(p)] = p if isinstance(p, np.ndarray) else np.frombuffer(p, dtype=np.uint8)
                    pos += len(p)
                else:
                    self.pinned_out[pos:pos + p.numel()].copy_(p, non_blocking=True)
                    pos += p.numel()
            torch.cuda.current_stream(self.dev).synchronize()
        return HostBuf(self.pinned_out, 0, seg.n)


def build_message(stage, kind, names, result_dev, puid=None, tags=None, tags_dev=None, metrics=None, data_raw=None):
    """construct_response (utils.py:410-469): SeldonMessage{meta{puid, tags, metrics}, data{names, <kind>}}.
    tags: host values; tags_dev: {key: 1-D device tensor} packed on the device as number lists; data_raw: the bytes of
    a `data` field to echo verbatim instead of encoding result_dev (transform_input returns X unchanged)."""
    meta = Segments()
    if puid:
        pb = puid.encode("utf-8")
        meta.add(_ld(0x0A, len(pb)) + pb)
    for k, v in (tags or {}).items():
        kb, vb = str(k).encode("utf-8"), encode_value_host(v)
        e = _ld(0x0A, len(kb)) + kb + _ld(0x12, len(vb)) + vb
        meta.add(_ld(0x12, len(e)) + e)
    for k, t in (tags_dev or {}).items():
        kb = str(k).encode("utf-8")
        lst = stage.pack_payload("ndarray", t)                       # ListValue body: 11 bytes per number
        val = lst.wrapped(0x32)                                       # Value{list_value}
        entry = Segments().add(_ld(0x0A, len(kb)) + kb).extend(val.wrapped(0x12))
        meta.extend(entry.wrapped(0x12))
    meta.add(encode_metrics_host(metrics))
    msg = meta.wrapped(0x12)                                          # meta is always present (set explicitly, L468)
    if data_raw is not None:
        msg.add(_ld(0x1A, len(data_raw))).add(data_raw)
        return stage.finish(msg)
    data = Segments()
    for nm in names or []:
        nb = str(nm).encode("utf-8")
        data.add(_ld(0x0A, len(nb)) + nb)
    data.extend(stage.pack_payload(kind, result_dev).wrapped(0x12 if kind == "tensor" else 0x1A))
    msg.extend(data.wrapped(0x1A))
    return stage.finish(msg)


# ------------------------------------------------------------------------------------------------ helpers for callers
def make_tensor_request_pinned(X, names=None, puid=None):
    """A serialised SeldonMessage{data{names, tensor{shape, values = X as doubles}}} in pinned host memory, built from
    a CUDA/CPU float tensor or numpy array (benchmark and test helper: what a gRPC client would send)."""
    import torch

    if isinstance(X, torch.Tensor):
        rows, cols = X.shape
    else:
        X = np.asarray(X)
        rows, cols = X.shape
    shape = _varint(rows) + _varint(cols)
    tensor_head = _ld(0x0A, len(shape)) + shape + _ld(0x12, 8 * rows * cols)
    names_b = b"".join(_ld(0x0A, len(n.encode())) + n.encode() for n in (names or []))
    data_len = len(names_b) + 1 + len(_varint(len(tensor_head) + 8 * rows * cols)) + len(tensor_head) + 8 * rows * cols
    head = b""
    if puid:
        pb = puid.encode("utf-8")
        m = _ld(0x0A, len(pb)) + pb
        head += _ld(0x12, len(m)) + m
    head += _ld(0x1A, data_len) + names_b + _ld(0x12, len(tensor_head) + 8 * rows * cols) + tensor_head
    pad = (-len(head)) % 16                      # the doubles start on a 16-byte boundary of the pinned buffer
    buf = torch.empty(pad + len(head) + 8 * rows * cols, dtype=torch.uint8).pin_memory()
    buf[pad:pad + len(head)] = torch.frombuffer(bytearray(head), dtype=torch.uint8)
    vals = buf[pad + len(head):].view(torch.float64).view(rows, cols)
    chunk = max(1, (64 << 20) // max(8 * cols, 1))
    for lo in range(0, rows, chunk):
        hi = min(rows, lo + chunk)
        if isinstance(X, torch.Tensor):
            vals[lo:hi].copy_(X[lo:hi].to(torch.float64))
        else:
            vals[lo:hi] = torch.from_numpy(np.ascontiguousarray(X[lo:hi], dtype=np.float64))
    return HostBuf(buf, pad, len(head) + 8 * rows * cols)


def parse_tensor_response(resp):
    """numpy view (float64) of the values of a response's data.tensor, reshaped (test / benchmark helper)."""
    info = parse_seldon_message(resp)
    mv = _mv(resp)
    if info["kind"] != "tensor":
        raise SeldonB200Error("not a tensor response")
    shape, voff = _tensor_layout(mv, info["payload"][0], info["payload"][1])
    n = int(np.prod(shape)) if shape else 0
    return np.frombuffer(mv[voff:voff + 8 * n], dtype="<f8").reshape(shape)
