"""GPU parity of the payload codec: device kernels vs the oracle's wire bytes (bit-exact bytes / exact fp32 values)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle import codec as ocodec
from seldon_core_b200 import codec
from seldon_core_b200.errors import SeldonB200Error


@pytest.mark.parametrize("shape", [(1, 1), (2, 3), (300, 7), (129, 200), (10000, 4), (4096, 64)])
def test_decode_tensor(shape):
    import torch

    a = np.random.RandomState(0).randn(*shape)
    out = codec.decode_tensor(ocodec.encode_tensor(a))
    assert out.dtype == torch.float32 and tuple(out.shape) == shape
    assert np.array_equal(out.cpu().numpy(), a.astype(np.float32))   # fp64 -> fp32 round-to-nearest, exact


def test_decode_tensor_fused_standardise():
    import torch

    rs = np.random.RandomState(1)
    a, mu, sigma = rs.randn(500, 16), rs.randn(16), rs.rand(16) + 0.5
    mu_d = torch.tensor(mu, dtype=torch.float32, device="cuda")
    inv_d = torch.tensor(1.0 / (sigma + 1e-10), dtype=torch.float32, device="cuda")
    out = codec.decode_tensor(ocodec.encode_tensor(a), mu=mu_d, inv_sigma=inv_d).cpu().numpy()
    ref = (a - mu_d.cpu().numpy().astype(np.float64)) * inv_d.cpu().numpy().astype(np.float64)
    assert np.max(np.abs(out - ref)) <= 1e-6 * np.max(np.abs(ref))


@pytest.mark.parametrize("shape", [(1, 1), (3, 4), (2, 64), (5, 256), (1000, 12), (777, 11)])
def test_decode_ndarray(shape):
    a = np.random.RandomState(2).randn(*shape)
    a.flat[0] = 0.0
    wire = ocodec.encode_listvalue(a)
    assert codec.ndarray_cols(wire) == shape[1]
    out = codec.decode_ndarray(wire, shape[1])
    assert np.array_equal(out.cpu().numpy(), a.astype(np.float32))


def test_decode_ndarray_rejects_malformed():
    a = np.random.RandomState(3).randn(4, 5)
    wire = bytearray(ocodec.encode_listvalue(a))
    wire[30] ^= 0xFF if wire[30] in (0x0A, 0x09, 0x11) else 0x00
    bad = bytearray(wire)
    row_bytes, hdr = ocodec.listvalue_row_layout(5)
    bad[row_bytes + hdr] = 0x1A  # string_value tag in row 1, col 0
    with pytest.raises(SeldonB200Error):
        codec.decode_ndarray(bytes(bad), 5)
    with pytest.raises(SeldonB200Error):
        codec.decode_ndarray(bytes(wire[:-3]), 5)   # truncated


def test_encode_tensor_and_ndarray_bytes():
    import torch

    rs = np.random.RandomState(4)
    for shape in ((1, 1), (3, 4), (1000, 10), (257, 64)):
        a = rs.randn(*shape).astype(np.float32)
        t = torch.from_numpy(a).cuda()
        assert codec.encode_tensor(t) == ocodec.encode_tensor(a.astype(np.float64))
        assert codec.encode_ndarray(t) == ocodec.encode_listvalue(a.astype(np.float64))
    flags = torch.tensor(rs.randint(0, 2, 1001), dtype=torch.int32, device="cuda")
    assert codec.encode_ndarray(flags) == ocodec.encode_listvalue(flags.cpu().numpy().astype(np.float64))
    assert codec.encode_tensor(flags) == ocodec.encode_tensor(flags.cpu().numpy().astype(np.float64))


def test_round_trip_property_large():
    import torch

    t = torch.randn(200000, 64, device="cuda")
    assert torch.equal(codec.decode_tensor(codec.encode_tensor(t)), t)
    t2 = t[:20000]
    assert torch.equal(codec.decode_ndarray(codec.encode_ndarray(t2), 64), t2)


@pytest.mark.parametrize("shape", [(16, 8), (17, 8), (4097, 8), (33, 128), (100, 1024)])
def test_ndarray_ring_kernels_round_trip_and_bytes(shape):
    """cols % 4 == 0 takes the bulk-copy ring kernels: chunk boundaries, the < 16-byte tail of the last chunk."""
    import torch

    a = np.random.RandomState(5).randn(*shape).astype(np.float32)
    wire = ocodec.encode_listvalue(a.astype(np.float64))
    assert np.array_equal(codec.decode_ndarray(wire, shape[1]).cpu().numpy(), a)
    assert codec.encode_ndarray(torch.from_numpy(a).cuda()) == wire


def test_ndarray_ring_kernel_rejects_bad_tag_in_a_late_chunk():
    a = np.random.RandomState(6).randn(1000, 8)
    wire = bytearray(ocodec.encode_listvalue(a))
    row_bytes, hdr = ocodec.listvalue_row_layout(8)
    wire[937 * row_bytes + hdr + 11 * 5 + 2] = 0x19  # fixed64 tag of row 937, column 5 -> another field number
    with pytest.raises(SeldonB200Error):
        codec.decode_ndarray(bytes(wire), 8)
    wire = bytearray(ocodec.encode_listvalue(a))
    wire[500 * row_bytes] = 0x12  # row header tag
    with pytest.raises(SeldonB200Error):
        codec.decode_ndarray(bytes(wire), 8)


def test_decode_tensor_rejects_shape_payload_mismatch():
    """A Tensor whose shape claims more values than the message carries must raise, as the reference does on the
    reshape (utils.py:169), both in the host parser and at the C ABI (no out-of-bounds read on the device)."""
    import torch

    from seldon_core_b200 import _cabi

    good = ocodec.encode_tensor(np.arange(6.0).reshape(2, 3))
    lying = b"\x0a\x04\xe8\x07\xe8\x07" + b"\x12\x10" + b"\x00" * 16          # shape [1000, 1000], 16 bytes of values
    for bad in (lying, good[:-8], good + b"\x00" * 8, good[:6]):
        with pytest.raises(SeldonB200Error):
            codec.decode_tensor(bad)
    wire = torch.zeros(64, dtype=torch.uint8, device="cuda")
    out = torch.empty((1000, 1000), dtype=torch.float32, device="cuda")
    rc = _cabi.lib().sb_unpack_tensor_f64(wire.data_ptr(), wire.numel(), 8, 1000, 1000, None, None, out.data_ptr(), 1000,
                                          torch.cuda.current_stream().cuda_stream)
    assert rc == -1 and b"payload has 64" in _cabi.lib().sb_last_error()
    rc = _cabi.lib().sb_unpack_tensor_f64(wire.data_ptr(), wire.numel(), 2 ** 63, 1, 1, None, None, out.data_ptr(), 1,
                                          torch.cuda.current_stream().cuda_stream)
    assert rc == -1
