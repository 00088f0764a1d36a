"""Eigen-step timing at d = 128 / 256: iterations of the subspace kernel (debug counter 7) and the device time of the
eigen step (SB_OPT_PROFILE) per batch, with the subspace iteration on and off."""
import ctypes
import sys

import torch

sys.path.insert(0, ".")
from seldon_core_b200 import OutlierMahalanobisB200, _cabi, synth  # noqa: E402

dev = torch.device("cuda")
for d in (128, 256):
    for nosub in (0, 1):
        det = OutlierMahalanobisB200(n_components=16)
        det.set_option(_cabi.SB_OPT_PROFILE, 1)
        det.set_option(_cabi.SB_OPT_NO_SUBSPACE, nosub)
        out = []
        for i in range(6):
            X, _ = synth.maha_rows_torch(1 << 18, d, 16, 900 + i, dev)
            det.score_device(X)
            km = (ctypes.c_float * 5)()
            _cabi.check(_cabi.lib().sb_maha_kernel_times(det._h, km, torch.cuda.current_stream().cuda_stream))
            c = (ctypes.c_int64 * 8)()
            _cabi.check(_cabi.lib().sb_maha_debug_counters(det._h, c, torch.cuda.current_stream().cuda_stream))
            out.append("it=%d sweeps=%d eigh=%.3fms" % (c[7], c[3], km[1]))
        print("d=%d no_subspace=%d:" % (d, nosub), " | ".join(out), flush=True)
