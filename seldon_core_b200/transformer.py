"""Standardise input transformer: (X - mu) / (sigma + 1e-10) per feature, on the GPU.

Spec: ``components/outlier-detection/vae/CoreVAE.py:128`` of the reference; the class shape follows
``examples/transformers/mean_transformer/MeanTransformer.py:8-12`` (``transform_input(X, feature_names)``).
"""
import numpy as np

from . import _cabi
from .errors import SeldonB200Error


class StandardiseB200(object):
    def __init__(self, mu, sigma, device=None):
        self.mu = np.asarray(mu, dtype=np.float64)
        self.sigma = np.asarray(sigma, dtype=np.float64)
        self._device = device
        self._dev_params = None

    def _params(self, dev):
        import torch

        if self._dev_params is None or self._dev_params[0].device != dev:
            mu = torch.tensor(self.mu, dtype=torch.float32, device=dev)
            inv = torch.tensor(1.0 / (self.sigma + 1e-10), dtype=torch.float32, device=dev)
            self._dev_params = (mu, inv)
        return self._dev_params

    def transform_device(self, X):
        import torch

        mu, inv = self._params(X.device)
        if X.shape[1] != mu.numel():
            raise SeldonB200Error("expected %d features, got %d" % (mu.numel(), X.shape[1]))
        out = torch.empty_like(X)
        with torch.cuda.device(X.device):
            _cabi.check(_cabi.lib().sb_standardise(X.data_ptr(), X.shape[0], X.shape[1], X.stride(0), mu.data_ptr(),
                                                   inv.data_ptr(), out.data_ptr(), out.stride(0),
                                                   torch.cuda.current_stream().cuda_stream))
        return out

    def transform_input(self, X, feature_names=None):
        import torch

        if isinstance(X, torch.Tensor) and X.is_cuda:
            return self.transform_device(X.to(torch.float32).contiguous())
        if not torch.cuda.is_available():
            raise SeldonB200Error("no CUDA device: seldon_core_b200 has no CPU fallback", status_code=500)
        dev = torch.device(self._device or ("cuda:%d" % torch.cuda.current_device()))
        Xd = torch.from_numpy(np.ascontiguousarray(np.asarray(X), dtype=np.float32)).to(dev)
        return self.transform_device(Xd).double().cpu().numpy()
