"""Where the time of a small d = 256 batch goes (the per-rank share of a 16 Mi-row C5 batch on 8 GPUs is 2 Mi rows):
step time against the sum of the profiled kernels, with and without the combiner, on one GPU (no NCCL)."""
import ctypes
import sys
import time

import torch

sys.path.insert(0, ".")
from seldon_core_b200 import MeanCombinerB200, OutlierMahalanobisB200, StandardiseB200, _cabi, synth  # noqa: E402

dev = torch.device("cuda")
d, rows = 256, 1 << 21
g = torch.Generator(device=dev).manual_seed(7)
mu = torch.randn(d, generator=g, device=dev).double().cpu().numpy()
sigma = (0.5 + torch.rand(d, generator=g, device=dev)).double().cpu().numpy()
det = OutlierMahalanobisB200(n_components=16)
det.set_option(_cabi.SB_OPT_PROFILE, 1)
prime, _ = synth.maha_rows_torch(1 << 18, d, 16, 201, dev)
det.score_device(prime)
det.fuse_standardise(StandardiseB200(mu, sigma))
X, _ = synth.maha_rows_torch(rows, d, 16, 300, dev)
X.mul_(torch.tensor(sigma, dtype=torch.float32, device=dev) + 1e-10).add_(torch.tensor(mu, dtype=torch.float32, device=dev))
kids = [torch.softmax(torch.randn((rows, 8), generator=g, device=dev), dim=1) for _ in range(4)]
for _ in range(5):
    det.score_device(X)
torch.cuda.synchronize()
for label, comb in (("detector only", False), ("detector + combiner", True)):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(20):
        det.score_device(X)
        if comb:
            MeanCombinerB200.aggregate_device(kids)
    e1.record()
    t_host = time.perf_counter() - t0
    torch.cuda.synchronize()
    km = (ctypes.c_float * 5)()
    _cabi.check(_cabi.lib().sb_maha_kernel_times(det._h, km, torch.cuda.current_stream().cuda_stream))
    print("%s: %.3f ms/step on the device, host issue time %.3f ms/step, profiled kernels %s sum %.3f ms"
          % (label, e0.elapsed_time(e1) / 20, t_host * 1e3 / 20, ["%.3f" % v for v in km], sum(km)), flush=True)
