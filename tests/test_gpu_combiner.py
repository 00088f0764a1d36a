"""GPU parity: combiner mean and standardise through the C ABI vs the oracle (tolerance 1e-6 relative: pure adds)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle.combiner import mean_combiner, standardise
from seldon_core_b200 import MeanCombinerB200, StandardiseB200, synth
from seldon_core_b200.errors import SeldonB200Error


def test_average_combiner_kats_on_device():
    c = MeanCombinerB200()
    out = c.aggregate([np.array([[1.0, 1.0]]), np.array([[1.0, 0.5]]), np.array([[2.2, 0.9]])], [])
    assert abs(out[0, 0] - (1.0 + 1.0 + 2.2) / 3) < 1e-7 * 4 and abs(out[0, 1] - (1.0 + 0.5 + 0.9) / 3) < 1e-6
    assert c.aggregate([np.array([[1.0]]), np.array([[1.0]])], [])[0, 0] == 1.0
    single = np.array([[1.0, 5.0, 0.3]])
    assert np.allclose(c.aggregate([single], []), single, rtol=1e-7)
    with pytest.raises(SeldonB200Error):
        c.aggregate([np.zeros((1, 2)), np.zeros((1, 2)), np.zeros((1, 3))], [])


@pytest.mark.parametrize("n_children", [1, 2, 3, 4, 5, 8, 16])
@pytest.mark.parametrize("shape", [(1, 1), (7, 3), (257, 10), (1024, 1000), (33, 1001)])
def test_combiner_matches_oracle(n_children, shape):
    kids = [synth.softmax_rows_numpy(shape[0], shape[1], 10 + j) for j in range(n_children)]
    out = MeanCombinerB200().aggregate(kids, [])
    ref = mean_combiner(kids)
    assert out.shape == ref.shape and out.dtype == np.float64
    assert np.max(np.abs(out - ref) / np.maximum(np.abs(ref), 1e-30)) <= 1e-6


def test_combiner_device_tensors_and_unaligned_views():
    import torch

    base = [torch.rand(1000 * 37 + 3, device="cuda") for _ in range(8)]
    views = [b[1:1 + 1000 * 37].view(1000, 37) for b in base]          # 4-byte aligned only
    out = MeanCombinerB200().aggregate(views, [])
    ref = torch.stack([v.double() for v in views]).sum(0) / 8
    assert out.is_cuda and torch.max(torch.abs(out.double() - ref) / ref.abs().clamp_min(1e-30)) <= 1e-6


def test_mnist_combiner_fixture(golden_dir):
    import os

    g = np.load(os.path.join(golden_dir, "combiner_mnist.npz"))
    out = MeanCombinerB200().aggregate([g["a"], g["b"]], [])
    assert np.max(np.abs(out - g["out"]) / np.abs(g["out"])) <= 1e-6


@pytest.mark.parametrize("shape", [(513, 37), (1000, 256), (3, 4), (70001, 64)])
def test_standardise_matches_oracle(shape):
    rs = np.random.RandomState(0)
    X = rs.randn(*shape).astype(np.float32)  # 37 columns: scalar kernel; multiples of 4: the float4 kernel
    mu, sigma = rs.randn(shape[1]), rs.rand(shape[1]) + 0.1
    out = StandardiseB200(mu, sigma).transform_input(X, None)
    ref = standardise(X, mu.astype(np.float32), sigma)
    assert np.max(np.abs(out - ref) / np.maximum(np.abs(ref), 1e-3)) < 1e-5
