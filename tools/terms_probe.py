"""d = 256: 1xTF32 against 3xTF32 in the pass-1 Gram kernel - score / covariance error against the fp64 oracle and the
kernel time (SB_OPT_GRAM_TERMS; SURVEY 7 hard part 2 leaves the choice at d = 256 to the measurements)."""
import ctypes
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from oracle.mahalanobis import MahalanobisLean  # noqa: E402
from seldon_core_b200 import OutlierMahalanobisB200, _cabi, synth  # noqa: E402

d = 256
prime, _ = synth.maha_rows_numpy(6000, d, 16, seed=51)
X, labels = synth.maha_rows_numpy(60000, d, 16, seed=52)
ref = MahalanobisLean(n_components=16)
ref.predict(prime.astype(np.float64), None)
pr = ref.predict(X.astype(np.float64), None)
Xbig, _ = synth.maha_rows_torch(1 << 22, d, 16, 77, torch.device("cuda"))
for terms in (3, 1):
    det = OutlierMahalanobisB200(n_components=16)
    det.set_option(_cabi.SB_OPT_GRAM_TERMS, terms)
    det.set_option(_cabi.SB_OPT_PROFILE, 1)
    det.predict(prime, None)
    pg = det.predict(X, None)
    err = np.max(np.abs(det.score - ref.score) / np.maximum(np.abs(ref.score), 1e-6))
    cerr = np.max(np.abs(det.C - ref.C)) / np.max(np.abs(ref.C))
    for _ in range(3):
        det.score_device(Xbig)
    km = (ctypes.c_float * 5)()
    _cabi.check(_cabi.lib().sb_maha_kernel_times(det._h, km, torch.cuda.current_stream().cuda_stream))
    print("terms=%d: max rel score err %.2e, cov err (norm-wise) %.2e, flags equal %s; pass 1 on 4 Mi x 256 rows %.3f ms"
          % (terms, err, cerr, bool(np.array_equal(pg, pr)), km[0]), flush=True)
