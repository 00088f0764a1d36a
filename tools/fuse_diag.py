"""Diagnostic: fused standardise against the two-step graph, and run-to-run spread of the two-step graph itself."""
import sys
sys.path.insert(0, '/root/repo')
import numpy as np, torch
from seldon_core_b200 import OutlierMahalanobisB200, StandardiseB200, synth
dev = torch.device("cuda")
for d in (64, 256):
    g = torch.Generator(device=dev).manual_seed(5)
    mu = torch.randn(d, generator=g, device=dev).double().cpu().numpy()
    sigma = (0.5 + torch.rand(d, generator=g, device=dev)).double().cpu().numpy()
    std = StandardiseB200(mu, sigma)
    a, b, c = (OutlierMahalanobisB200(n_components=16) for _ in range(3))
    c.fuse_standardise(std)
    for i, rows in enumerate((40000, 70001, 128, 300000)):
        Xs, _ = synth.maha_rows_torch(rows, d, 16, 400 + i, dev)
        Xraw = (Xs * (torch.tensor(sigma, dtype=torch.float32, device=dev) + 1e-10) + torch.tensor(mu, dtype=torch.float32, device=dev)).contiguous()
        Xt = std.transform_device(Xraw)
        sa, fa = a.score_device(Xt); sb, fb = b.score_device(Xt); sc, fc = c.score_device(Xraw)
        rel = lambda x, y: float(((x - y).abs() / y.abs().clamp_min(1e-3)).max())
        print(d, rows, "two-step vs two-step: %.2e flags %d | fused vs two-step: %.2e flags %d" % (rel(sa, sb), int((fa != fb).sum()), rel(sc, sa), int((fc != fa).sum())))
