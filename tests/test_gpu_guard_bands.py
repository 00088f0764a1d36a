"""A memcheck stand-in (compute-sanitizer is closed on this pool): every buffer the main C-ABI entry points read or
write is carved out of a larger allocation with canary bands on both sides; after the calls the bands must be intact
and the inputs unmodified.  Ragged sizes on purpose (rows that are no multiple of any tile)."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from seldon_core_b200 import _cabi, synth

BAND = 4096
PATTERN = 0xA5


class Guarded(object):
    """A device buffer of `nbytes` bytes with BAND bytes of canary on each side (256-byte aligned payload)."""

    def __init__(self, nbytes, fill=None):
        import torch

        self.n = int(nbytes)
        pad = (-self.n) % 256
        self.raw = torch.full((BAND + self.n + pad + BAND,), PATTERN, dtype=torch.uint8, device="cuda")
        self.payload = self.raw[BAND:BAND + self.n]
        if fill is not None:
            self.payload.copy_(fill.contiguous().view(torch.uint8).view(-1))
        assert self.payload.data_ptr() % 256 == 0

    @property
    def ptr(self):
        return self.payload.data_ptr()

    def view(self, dtype, shape):
        return self.payload.view(dtype).view(shape)

    def intact(self):
        import torch

        head, tail = self.raw[:BAND], self.raw[BAND + self.n:]
        return bool(torch.all(head == PATTERN)) and bool(torch.all(tail == PATTERN))


def stream():
    import torch

    return torch.cuda.current_stream().cuda_stream


@pytest.mark.parametrize("d,rows", [(64, (3001, 12345, 7)), (256, (2600, 4097)), (18, (500, 1031))])
def test_mahalanobis_batch_stays_inside_its_buffers(d, rows):
    import torch

    L = _cabi.lib()
    h = ctypes.c_void_p()
    _cabi.check(L.sb_maha_create(ctypes.byref(h), d, 16 if d > 18 else 4, 25.0, 3.0, 50, -1))
    try:
        for i, r in enumerate(rows):
            X, _ = synth.maha_rows_numpy(r, d, 16 if d > 18 else 4, seed=40 + i)
            xt = torch.from_numpy(X).cuda()
            gx = Guarded(xt.numel() * 4, xt)
            need = ctypes.c_uint64()
            _cabi.check(L.sb_maha_workspace_bytes(h, r, ctypes.byref(need)))
            ws, sc, fl = Guarded(need.value), Guarded(4 * r), Guarded(4 * r)
            _cabi.check(L.sb_maha_batch(h, gx.ptr, r, d, sc.ptr, fl.ptr, ws.ptr, need.value, stream()))
            torch.cuda.synchronize()
            for name, g in (("X", gx), ("workspace", ws), ("score", sc), ("flag", fl)):
                assert g.intact(), "%s: canary band overwritten (d=%d rows=%d)" % (name, d, r)
            assert torch.equal(gx.view(torch.float32, (r, d)), xt), "the input batch was modified"
            s = sc.view(torch.float32, (r,))
            f = fl.view(torch.int32, (r,))
            assert bool(torch.isfinite(s).all()) and bool(((f == 0) | (f == 1)).all())
    finally:
        L.sb_maha_destroy(h)


def test_codec_and_combiner_stay_inside_their_buffers():
    import torch

    L = _cabi.lib()
    st = stream()
    for rows, cols in ((1, 1), (333, 7), (1025, 64), (50, 257)):
        x = torch.randn((rows, cols), device="cuda")
        gx = Guarded(x.numel() * 4, x)
        # Tensor: pack then unpack
        wire = Guarded(8 * rows * cols)
        _cabi.check(L.sb_pack_tensor_f64(gx.ptr, 0, rows * cols, wire.ptr, st))
        back = Guarded(4 * rows * cols)
        _cabi.check(L.sb_unpack_tensor_f64(wire.ptr, 8 * rows * cols, 0, rows, cols, None, None, back.ptr, cols, st))
        # ListValue
        nb = int(L.sb_listvalue_nbytes(rows, cols))
        lv, back2, status = Guarded(nb), Guarded(4 * rows * cols), Guarded(4)
        _cabi.check(L.sb_pack_listvalue(gx.ptr, 0, rows, cols, lv.ptr, st))
        _cabi.check(L.sb_unpack_listvalue(lv.ptr, nb, rows, cols, back2.ptr, cols, status.ptr, st))
        # JSON text both ways
        cap = int(L.sb_json_text_capacity(rows, cols))
        text, nbytes = Guarded(cap), Guarded(8)
        ews = Guarded(int(L.sb_json_encode_workspace_bytes(rows, cols)))
        _cabi.check(L.sb_pack_json_ndarray(gx.ptr, 0, rows, cols, ews.ptr, ews.n, text.ptr, nbytes.ptr, st))
        torch.cuda.synchronize()
        n = int(nbytes.view(torch.int64, (1,)).item())
        assert 0 < n <= cap
        dws = Guarded(int(L.sb_json_workspace_bytes(n)))
        counts, back3 = Guarded(32), Guarded(4 * rows * cols)
        _cabi.check(L.sb_json_ndarray_count(text.ptr, n, dws.ptr, dws.n, counts.ptr, st))
        _cabi.check(L.sb_unpack_json_ndarray(text.ptr, n, rows, cols, dws.ptr, back3.ptr, cols, status.ptr, st))
        torch.cuda.synchronize()
        for name, g in (("x", gx), ("wire", wire), ("back", back), ("listvalue", lv), ("back2", back2), ("status", status),
                        ("text", text), ("nbytes", nbytes), ("enc ws", ews), ("dec ws", dws), ("counts", counts),
                        ("back3", back3)):
            assert g.intact(), "%s: canary band overwritten (%d x %d)" % (name, rows, cols)
        for b in (back, back2, back3):
            assert torch.equal(b.view(torch.float32, (rows, cols)), x)
    # combiner, 3 children, ragged element count
    kids = [Guarded(4 * 777 * 13, torch.randn((777, 13), device="cuda")) for _ in range(3)]
    out = Guarded(4 * 777 * 13)
    ptrs = (ctypes.c_void_p * 3)(*[k.ptr for k in kids])
    _cabi.check(L.sb_combiner_mean(ptrs, 3, 777 * 13, out.ptr, st))
    torch.cuda.synchronize()
    assert out.intact() and all(k.intact() for k in kids)
    want = sum(k.view(torch.float32, (777, 13)) for k in kids) / 3.0
    assert torch.allclose(out.view(torch.float32, (777, 13)), want, rtol=1e-6, atol=0)


def test_routers_stay_inside_their_buffers():
    import torch

    L = _cabi.lib()
    for kind, n in ((0, 5), (1, 33)):
        h = ctypes.c_void_p()
        _cabi.check(L.sb_router_create(ctypes.byref(h), kind, n, 0.1, 7, 0))
        T = 1234
        p, u = synth.router_rewards(T, n)
        gu, gp = Guarded(4 * T, torch.from_numpy(u).cuda()), Guarded(4 * n, torch.from_numpy(p).cuda())
        dec = Guarded(4 * T)
        _cabi.check(L.sb_router_run(h, T, None, gu.ptr, gp.ptr, 2, dec.ptr, stream()))
        torch.cuda.synchronize()
        assert dec.intact() and gu.intact() and gp.intact()
        dv = dec.view(torch.int32, (T,))
        assert int(dv.min()) >= 0 and int(dv.max()) < n
        L.sb_router_destroy(h)
