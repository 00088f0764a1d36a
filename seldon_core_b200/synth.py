"""Synthetic inputs for the configurations of BASELINE.json.

numpy generators for tests / fixtures (CPU) and torch generators for the benchmark (on device,
Philox).  Every value is exactly representable in fp32 so the fp64 oracle and the fp32 device path
consume the same numbers.
"""
import numpy as np


def maha_spectrum(d, k):
    """k leading eigenvalues log-spaced 8 -> 1, the other d-k equal to 0.05 (eigengap 20 at k)."""
    lam = np.full(d, 0.05)
    lam[:k] = np.logspace(np.log10(8.0), 0.0, k)
    return lam


def maha_basis(d, seed):
    q, r = np.linalg.qr(np.random.RandomState(seed).randn(d, d))
    return q * np.sign(np.diag(r))


def maha_rows_numpy(nb, d, k, seed, outlier_frac=0.01, basis_seed=1234, unconstrained=False):
    """Rows x = mu + Q diag(sqrt(lam)) z with a score gap around threshold 25.

    Inliers: the radius^2 of the leading-k part of z is rescaled into [0, 17]; planted outliers get
    radius^2 in [60, 120].  Returns (X float32 [nb,d], labels int8 [nb]).
    """
    rs = np.random.RandomState(seed)
    lam = maha_spectrum(d, k)
    Q = maha_basis(d, basis_seed)
    mu = np.random.RandomState(basis_seed + 1).uniform(-1, 1, d)
    z = rs.randn(nb, d)
    labels = np.zeros(nb, dtype=np.int8)
    if not unconstrained:
        r2 = (z[:, :k] ** 2).sum(axis=1)
        target = 17.0 * rs.rand(nb) ** (2.0 / k)  # radius^2 of a uniform point in the k-ball of radius^2 17
        labels = (rs.rand(nb) < outlier_frac).astype(np.int8)
        target = np.where(labels == 1, rs.uniform(60.0, 120.0, nb), target)
        z[:, :k] *= np.sqrt(target / np.maximum(r2, 1e-30))[:, None]
    X = mu + (z * np.sqrt(lam)) @ Q.T
    return X.astype(np.float32), labels


def maha_rows_torch(nb, d, k, seed, device, outlier_frac=0.01, basis_seed=1234, out=None, chunk=1 << 20):
    """Same recipe generated on ``device`` with torch's Philox generator (different stream from numpy)."""
    import torch

    g = torch.Generator(device=device)
    g.manual_seed(seed)
    lam = torch.tensor(maha_spectrum(d, k), dtype=torch.float32, device=device)
    Q = torch.tensor(maha_basis(d, basis_seed), dtype=torch.float32, device=device)
    mu = torch.tensor(np.random.RandomState(basis_seed + 1).uniform(-1, 1, d), dtype=torch.float32, device=device)
    X = out if out is not None else torch.empty((nb, d), dtype=torch.float32, device=device)
    labels = torch.empty(nb, dtype=torch.int8, device=device)
    for lo in range(0, nb, chunk):
        m = min(chunk, nb - lo)
        z = torch.randn((m, d), generator=g, device=device)
        r2 = (z[:, :k] ** 2).sum(dim=1)
        u = torch.rand((m, 3), generator=g, device=device)
        target = 17.0 * u[:, 0] ** (2.0 / k)
        lab = u[:, 1] < outlier_frac
        target = torch.where(lab, 60.0 + 60.0 * u[:, 2], target)
        z[:, :k] *= torch.sqrt(target / r2.clamp_min(1e-30))[:, None]
        X[lo:lo + m] = mu + (z * torch.sqrt(lam)) @ Q.T
        labels[lo:lo + m] = lab.to(torch.int8)
    return X, labels


def softmax_rows_numpy(rows, cols, seed):
    z = np.random.RandomState(seed).randn(rows, cols)
    e = np.exp(z - z.max(axis=1, keepdims=True))
    return (e / e.sum(axis=1, keepdims=True)).astype(np.float32)


def router_rewards(T, n_branches, seed=11, perm_seed=7):
    """C4: per-arm success probabilities linspace(0.05,0.95) permuted, and the uniforms u_t."""
    p = np.random.RandomState(perm_seed).permutation(np.linspace(0.05, 0.95, n_branches)).astype(np.float32)
    u = np.random.RandomState(seed).rand(T).astype(np.float32)
    return p, u
