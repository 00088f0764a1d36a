"""Small fixed workloads for ncu captures (one mode per invocation, a handful of steady-state launches each).

    python tools/prof_workload.py c2     # 16Mi x 64 detector batch + combiner: pass1_gram_umma<64>, project_rows<1>, chain4
    python tools/prof_workload.py c5     # 4Mi x 256 fused-standardise batches: pass1_gram_umma<256>, project_rows<4>,
                                         # jacobi_cluster (first batch), subspace_topk (later batches)
    python tools/prof_workload.py codec  # tensor / ListValue / JSON codec kernels and the C3-shaped combiner
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from seldon_core_b200 import MeanCombinerB200, OutlierMahalanobisB200, StandardiseB200, _cabi, synth  # noqa: E402

dev = torch.device("cuda")
mode = sys.argv[1] if len(sys.argv) > 1 else "c2"
if mode == "c2":
    prime, _ = synth.maha_rows_torch(1 << 20, 64, 16, 101, dev)
    X, labels = synth.maha_rows_torch(1 << 24, 64, 16, 102, dev)
    kids = [torch.softmax(torch.randn((1 << 24, 8), device=dev), dim=1) for _ in range(4)]
    det = OutlierMahalanobisB200(n_components=16)
    det.score_device(prime)
    for _ in range(4):
        s, f = det.score_device(X)
        MeanCombinerB200.aggregate_device(kids)
    torch.cuda.synchronize()
    assert torch.equal(f.to(torch.int8), labels)
elif mode == "c5":
    d, rows = 256, 1 << 22
    g = torch.Generator(device=dev).manual_seed(7)
    mu = torch.randn(d, generator=g, device=dev).double().cpu().numpy()
    sigma = (0.5 + torch.rand(d, generator=g, device=dev)).double().cpu().numpy()
    det = OutlierMahalanobisB200(n_components=16)
    prime, _ = synth.maha_rows_torch(1 << 18, d, 16, 201, dev)
    det.score_device(prime)
    det.fuse_standardise(StandardiseB200(mu, sigma))
    for i in range(4):
        Xs, labels = synth.maha_rows_torch(rows, d, 16, 300 + i, dev)
        Xs.mul_(torch.tensor(sigma, dtype=torch.float32, device=dev) + 1e-10).add_(torch.tensor(mu, dtype=torch.float32, device=dev))
        s, f = det.score_device(Xs)
    torch.cuda.synchronize()
    assert torch.equal(f.to(torch.int8), labels)
else:
    L = _cabi.lib()
    st = torch.cuda.current_stream().cuda_stream
    rows, cols = 1 << 22, 64
    x = torch.randn((rows, cols), device=dev)
    wire = torch.empty(rows * cols * 8, dtype=torch.uint8, device=dev)
    y = torch.empty_like(x)
    for _ in range(3):
        _cabi.check(L.sb_pack_tensor_f64(x.data_ptr(), 0, rows * cols, wire.data_ptr(), st))
        _cabi.check(L.sb_unpack_tensor_f64(wire.data_ptr(), wire.numel(), 0, rows, cols, None, None, y.data_ptr(), cols, st))
    rows = 1 << 21
    x = torch.randn((rows, cols), device=dev)
    nb = int(L.sb_listvalue_nbytes(rows, cols))
    wire = torch.empty(nb, dtype=torch.uint8, device=dev)
    y = torch.empty_like(x)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    for _ in range(3):
        _cabi.check(L.sb_pack_listvalue(x.data_ptr(), 0, rows, cols, wire.data_ptr(), st))
        _cabi.check(L.sb_unpack_listvalue(wire.data_ptr(), nb, rows, cols, y.data_ptr(), cols, status.data_ptr(), st))
    a = np.random.RandomState(0).randn(4096, cols).astype(np.float32).astype(np.float64)
    inner = json.dumps(a.tolist())[1:-1]
    text = ("[" + ", ".join([inner] * 64) + "]").encode()
    rows = 4096 * 64
    buf = torch.frombuffer(bytearray(text), dtype=torch.uint8).to(dev)
    ws = torch.empty((int(L.sb_json_workspace_bytes(len(text))) + 7) // 8, dtype=torch.int64, device=dev)
    counts = torch.zeros(4, dtype=torch.int64, device=dev)
    y = torch.empty((rows, cols), dtype=torch.float32, device=dev)
    for _ in range(3):
        _cabi.check(L.sb_json_ndarray_count(buf.data_ptr(), len(text), ws.data_ptr(), ws.numel() * 8, counts.data_ptr(), st))
        _cabi.check(L.sb_unpack_json_ndarray(buf.data_ptr(), len(text), rows, cols, ws.data_ptr(), y.data_ptr(), cols, status.data_ptr(), st))
    from seldon_core_b200 import RandomABTestB200, codec  # noqa: E402

    xe = torch.randn((1 << 18, cols), device=dev)
    for _ in range(3):
        codec.encode_json_ndarray_device(xe)
    RandomABTestB200(ratioA=0.5).run_batched(1 << 24)
    kids = [torch.softmax(torch.randn((1 << 18, 1000), device=dev), dim=1) for _ in range(8)]
    for _ in range(3):
        MeanCombinerB200.aggregate_device(kids)
    torch.cuda.synchronize()
print("ok", mode)
