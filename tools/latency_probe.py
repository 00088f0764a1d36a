"""Where the time of a one-row request goes: device time of the batch kernels (SB_OPT_PROFILE), total device time of
sb_maha_batch (events), and wall time of the phases of predict_raw."""
import ctypes
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from seldon_core_b200 import OutlierMahalanobisB200, _cabi, synth, wire  # noqa: E402

det = OutlierMahalanobisB200(n_components=16)
det.set_option(_cabi.SB_OPT_PROFILE, 1)
prime, _ = synth.maha_rows_numpy(20000, 64, 16, seed=101)
det.predict(prime, None)
for rows in (1, 1000):
    X, _ = synth.maha_rows_numpy(rows, 64, 16, seed=7)
    Xd = torch.from_numpy(X).cuda()
    for _ in range(20):
        det.score_device(Xd)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(100):
        det.score_device(Xd)
    e1.record()
    t_issue = (time.perf_counter() - t0) / 100
    torch.cuda.synchronize()
    km = (ctypes.c_float * 5)()
    _cabi.check(_cabi.lib().sb_maha_kernel_times(det._h, km, torch.cuda.current_stream().cuda_stream))
    print("rows=%d: score_device %.1f us on the device, %.1f us host issue; profiled kernels (us): pass1 %.1f eigh %.1f prefix %.1f "
          "project %.1f chain %.1f" % (rows, e0.elapsed_time(e1) * 10, t_issue * 1e6, *[v * 1e3 for v in km]), flush=True)
    req = wire.make_tensor_request_pinned(X.astype(np.float64)).tobytes()
    for _ in range(20):
        det.predict_raw(req)
    t0 = time.perf_counter()
    for _ in range(100):
        det.predict_raw(req)
    print("         predict_raw %.1f us per request (wall)" % ((time.perf_counter() - t0) * 1e4), flush=True)
    # phases
    st = det._wire
    acc = [0.0] * 5
    for _ in range(100):
        t = [time.perf_counter()]
        host = st.host_bytes(req); info = wire.parse_seldon_message(host.numpy()); t.append(time.perf_counter())
        Xm = st.decode_matrix(host, info); torch.cuda.synchronize(); t.append(time.perf_counter())
        _, flag = det.score_device(Xm); torch.cuda.synchronize(); t.append(time.perf_counter())
        m = det.metrics(); t.append(time.perf_counter())
        resp = wire.build_message(st, info["kind"], [], flag, metrics=m); t.append(time.perf_counter())
        for i in range(5):
            acc[i] += t[i + 1] - t[i]
    print("         phases (us, with a device synchronisation after each): stage+parse %.1f, decode %.1f, batch %.1f, metrics %.1f, "
          "encode %.1f" % tuple(a * 1e4 for a in acc), flush=True)
