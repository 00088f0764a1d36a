"""Small workload for compute-sanitizer: the smoke batch (d = 64 tcgen05 Gram, projection, chain, register Jacobi), a
d = 256 pair of batches (cluster Jacobi on the first, subspace iteration on the second), the combiner, the codec
kernels, the JSON codec in both directions and both routers.  Sizes are small: the tools slow kernels down ~100x."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from seldon_core_b200 import (EpsilonGreedyB200, MeanCombinerB200, OutlierMahalanobisB200, RandomABTestB200,  # noqa: E402
                              ThompsonSamplingB200, codec, synth)

for d, rows in ((64, (3000, 4000)), (256, (2500, 3000))):
    det = OutlierMahalanobisB200(n_components=16)
    for i, r in enumerate(rows):
        X, _ = synth.maha_rows_numpy(r, d, 16, seed=i + 1)
        det.predict(X, None)
    print("maha d=%d ok, n=%d" % (d, det.n), flush=True)
kids = [synth.softmax_rows_numpy(257, 10, s) for s in range(4)]
MeanCombinerB200().aggregate(kids, [])
a = np.random.RandomState(0).randn(300, 12)
t = torch.from_numpy(a.astype(np.float32)).cuda()
codec.decode_tensor(codec.encode_tensor(t))
codec.decode_ndarray(codec.encode_ndarray(t), 12)
txt = codec.encode_json_ndarray(t)
codec.decode_json_ndarray(b'{"data":{"ndarray":' + txt + b'}}')
for r in (EpsilonGreedyB200(n_branches=8, seed=1), ThompsonSamplingB200(n_branches=8, seed=1)):
    u = np.random.RandomState(1).rand(300).astype(np.float32)
    r.run_batched(300, u, reward_mode=2, arm_p=np.linspace(0.1, 0.9, 8).astype(np.float32))
    r.route(None, None)
RandomABTestB200(ratioA=0.5).run_batched(1000)
torch.cuda.synchronize()
print("sanitizer workload ok")
