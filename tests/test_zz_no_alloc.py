"""The workspace-free statistics entry point (sb_gram_stats_noalloc) against the regular one.

Some callers cannot give libsb200 a device workspace: a JNI/cgo binding has no device allocator of its own, and a
serving thread must not call cudaMalloc on the request path (it synchronises the whole device).  For them the
library keeps the sub-segment partials in a static device array; this test checks that the entry point built on it
returns the same (count, sum, Gram) record as sb_maha_pass1_stats with a caller workspace - twice in a row with
different shapes, so the second call runs on top of whatever the first one left in the shared array - and that
arguments beyond the array's capacity are refused.  Collected last on purpose (file name), so a failure here cannot
hide the result of any other test under `pytest -x`.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from seldon_core_b200 import _cabi

TOL = 1e-4  # norm-wise, the tolerance of tests/test_gpu_maha.py for covariance-type quantities


def regular_stats(X):
    """Existing entry point: fresh handle (shift = 0, no clipping), caller-provided workspace."""
    import torch

    L = _cabi.lib()
    rows, d = X.shape
    h = ctypes.c_void_p()
    _cabi.check(L.sb_maha_create(ctypes.byref(h), d, min(3, d), 25.0, 3.0, 50, -1))
    try:
        need = ctypes.c_uint64()
        _cabi.check(L.sb_maha_workspace_bytes(h, rows, ctypes.byref(need)))
        ws = torch.empty(int(need.value), dtype=torch.uint8, device=X.device)
        out = torch.empty(1 + d + d * d, dtype=torch.float64, device=X.device)
        st = torch.cuda.current_stream().cuda_stream
        _cabi.check(L.sb_maha_pass1_stats(h, X.data_ptr(), rows, X.stride(0), ws.data_ptr(), ws.numel(), out.data_ptr(), st))
        torch.cuda.synchronize()
        L.sb_maha_abort_batch(h)
        return out.cpu().numpy()
    finally:
        L.sb_maha_destroy(h)


def noalloc_stats(X, shift):
    import torch

    rows, d = X.shape
    out = torch.full((1 + d + d * d,), float("nan"), dtype=torch.float64, device=X.device)
    before = torch.cuda.memory_allocated()
    _cabi.check(_cabi.lib().sb_gram_stats_noalloc(X.data_ptr(), rows, d, X.stride(0), shift.data_ptr(), out.data_ptr(),
                                                  torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    return out.cpu().numpy()


def same_record(a, b, d):
    assert a[0] == b[0]
    s_a, s_b, g_a, g_b = a[1:1 + d], b[1:1 + d], a[1 + d:], b[1 + d:]
    assert np.max(np.abs(s_a - s_b)) <= TOL * max(np.max(np.abs(s_b)), 1e-30)
    assert np.max(np.abs(g_a - g_b)) <= TOL * np.max(np.abs(g_b))
    assert np.all(np.isfinite(a))


def test_noalloc_matches_regular_entry_point_on_two_consecutive_calls():
    import torch

    g = torch.Generator(device="cuda").manual_seed(7)
    # call 1: the tensor-core path (d = 64), a ragged row count
    X1 = torch.randn((200001, 64), generator=g, device="cuda") * 2.0 + 0.5
    z64 = torch.zeros(64, device="cuda")
    same_record(noalloc_stats(X1, z64), regular_stats(X1), 64)
    # call 2, straight after: fewer rows and fewer features, so most of the shared array still holds call 1's partials
    X2 = torch.randn((5003, 18), generator=g, device="cuda") - 1.0
    z18 = torch.zeros(18, device="cuda")
    rec2 = noalloc_stats(X2, z18)
    same_record(rec2, regular_stats(X2), 18)
    ref = X2.double()
    assert np.allclose(rec2[1:19], ref.sum(0).cpu().numpy(), rtol=1e-5, atol=1e-3)
    assert np.max(np.abs(rec2[19:].reshape(18, 18) - (ref.T @ ref).cpu().numpy())) <= 1e-5 * float((ref.T @ ref).abs().max())
    # and the first shape once more on top of call 2's leftovers
    same_record(noalloc_stats(X1, z64), regular_stats(X1), 64)


def test_noalloc_rejects_sizes_beyond_the_static_array():
    import torch

    L = _cabi.lib()
    X = torch.zeros((8, 65), device="cuda")
    out = torch.zeros(1 + 65 + 65 * 65, dtype=torch.float64, device="cuda")
    z = torch.zeros(65, device="cuda")
    assert L.sb_gram_stats_noalloc(X.data_ptr(), 8, 65, 65, z.data_ptr(), out.data_ptr(), 0) == -1
    assert b"static scratch limit" in L.sb_last_error()
    assert L.sb_gram_stats_noalloc(X.data_ptr(), (1 << 20) + 1, 64, 64, z.data_ptr(), out.data_ptr(), 0) == -1
    assert b"1048576" in L.sb_last_error()
