"""Roofline measurements of the other rows of SURVEY section 8 on one B200: combiner (config 3), payload codec,
routers (config 4).  Prints one JSON line per component; CUDA-event timing, >= 3 warm-ups, inputs larger than L2.
The Mahalanobis headline lives in bench.py."""
import ctypes
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from seldon_core_b200 import EpsilonGreedyB200, MeanCombinerB200, ThompsonSamplingB200, _cabi, codec, synth  # noqa: E402

HBM = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.isfile(
    os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0


def timed(fn, warm=3, reps=10):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
    ev[0].record()
    for i in range(reps):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    ts = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(reps))
    return ts[0], ts[len(ts) // 2]


def emit(name, unit, work, nbytes, best_ms, med_ms, extra=None):
    gbs = nbytes / (med_ms * 1e-3) / 1e9
    line = {"component": name, "value": work / (med_ms * 1e-3), "unit": unit, "ms_median": med_ms, "ms_best": best_ms,
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": HBM, "unit": "GB/s", "frac": gbs / HBM,
                         "algorithmic_bytes": nbytes}}
    if extra:
        line.update(extra)
    print(json.dumps(line), flush=True)


def combiner():
    rows, cols, n = 1 << 20, 1000, 8   # BASELINE config 3: 8 children x 1M x 1000 fp32
    g = torch.Generator(device="cuda").manual_seed(1)
    kids = [torch.softmax(torch.randn((rows, cols), generator=g, device="cuda"), dim=1) for _ in range(n)]
    out = MeanCombinerB200.aggregate_device(kids)
    ref = torch.stack([k[:4096].double() for k in kids]).sum(0) / n
    err = float((out[:4096].double() - ref).abs().div(ref.abs().clamp_min(1e-30)).max())
    assert err <= 1e-6, err
    best, med = timed(lambda: MeanCombinerB200.aggregate_device(kids))
    emit("combiner_mean C3 (8 x 1Mi x 1000 fp32)", "rows/s", rows, 4 * (n + 1) * rows * cols, best, med,
         {"max_rel_err_vs_fp64": err})


def codec_bench():
    L = _cabi.lib()
    rows, cols = 1 << 22, 64
    st = torch.cuda.current_stream().cuda_stream
    x = torch.randn((rows, cols), device="cuda")
    wire = torch.empty(rows * cols * 8, dtype=torch.uint8, device="cuda")
    out = torch.empty_like(x)
    f_pack = lambda: _cabi.check(L.sb_pack_tensor_f64(x.data_ptr(), 0, rows * cols, wire.data_ptr(), st))
    f_unpack = lambda: _cabi.check(L.sb_unpack_tensor_f64(wire.data_ptr(), wire.numel(), 0, rows, cols, None, None, out.data_ptr(), cols, st))
    best, med = timed(f_pack)
    emit("pack tensor fp32->f64 (4Mi x 64)", "values/s", rows * cols, 12 * rows * cols, best, med)
    best, med = timed(f_unpack)
    assert torch.equal(out, x)
    emit("unpack tensor f64->fp32 (4Mi x 64)", "values/s", rows * cols, 12 * rows * cols, best, med)
    rows = 1 << 21
    x = torch.randn((rows, cols), device="cuda")
    nb = L.sb_listvalue_nbytes(rows, cols)
    wire = torch.empty(nb, dtype=torch.uint8, device="cuda")
    out = torch.empty_like(x)
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    f_pack = lambda: _cabi.check(L.sb_pack_listvalue(x.data_ptr(), 0, rows, cols, wire.data_ptr(), st))
    f_unpack = lambda: _cabi.check(L.sb_unpack_listvalue(wire.data_ptr(), nb, rows, cols, out.data_ptr(), cols, status.data_ptr(), st))
    best, med = timed(f_pack)
    emit("pack ndarray ListValue (2Mi x 64)", "values/s", rows * cols, nb + 4 * rows * cols, best, med)
    best, med = timed(f_unpack)
    assert torch.equal(out, x) and int(status.item()) == 0
    emit("unpack ndarray ListValue (2Mi x 64)", "values/s", rows * cols, nb + 4 * rows * cols, best, med)


def json_bench():
    """REST JSON ndarray text -> fp32: 4096 rows of json.dumps output tiled to 1 Mi rows x 64 (about 1.3 GB of text)."""
    import time

    import numpy as np

    L = _cabi.lib()
    st = torch.cuda.current_stream().cuda_stream
    block_rows, cols, tiles = 4096, 64, 256
    a = np.random.RandomState(0).randn(block_rows, cols).astype(np.float32).astype(np.float64)
    inner = json.dumps(a.tolist())[1:-1]                       # "[..], [..], ..." without the outer brackets
    text = ("[" + ", ".join([inner] * tiles) + "]").encode()
    rows = block_rows * tiles
    t0 = time.perf_counter()
    ref = np.array(json.loads(b'{"data":{"ndarray":[' + inner.encode() + b']}}')["data"]["ndarray"])
    cpu_s = time.perf_counter() - t0
    buf = torch.frombuffer(bytearray(text), dtype=torch.uint8).cuda()
    nbytes = len(text)
    ws = torch.empty((int(L.sb_json_workspace_bytes(nbytes)) + 7) // 8, dtype=torch.int64, device="cuda")
    counts = torch.zeros(4, dtype=torch.int64, device="cuda")
    out = torch.empty((rows, cols), dtype=torch.float32, device="cuda")
    status = torch.zeros(1, dtype=torch.int32, device="cuda")

    def run():
        _cabi.check(L.sb_json_ndarray_count(buf.data_ptr(), nbytes, ws.data_ptr(), ws.numel() * 8, counts.data_ptr(), st))
        _cabi.check(L.sb_unpack_json_ndarray(buf.data_ptr(), nbytes, rows, cols, ws.data_ptr(), out.data_ptr(), cols,
                                             status.data_ptr(), st))

    best, med = timed(run)
    assert int(status.item()) == 0 and [int(v) for v in counts.cpu()] == [rows * cols, rows + 1, rows + 1, 0]
    assert np.array_equal(out[:block_rows].cpu().numpy(), ref.astype(np.float32))
    assert torch.equal(out[-block_rows:], out[:block_rows])
    emit("unpack REST JSON ndarray text (1Mi x 64, %.2f GB, %.1f B/value)" % (nbytes / 1e9, nbytes / (rows * cols)),
         "values/s", rows * cols, nbytes + 4 * rows * cols, best, med,
         {"cpu_reference_values_per_s": block_rows * cols / cpu_s,
          "cpu_reference": "json.loads + np.array on one 4096-row block (utils.py:499-505), one host thread",
          "note": "two passes over the text (count, parse): DRAM traffic is about 2x the algorithmic bytes"})


def routers():
    for kind, T in (("eps", 10_000_000), ("thompson", 1_000_000)):
        p, u = synth.router_rewards(T, 32)
        ud, pd = torch.from_numpy(u).cuda(), torch.from_numpy(p).cuda()
        r = EpsilonGreedyB200(n_branches=32, epsilon=0.1, seed=1) if kind == "eps" else ThompsonSamplingB200(n_branches=32, seed=1)
        r.run_batched(1000, ud[:1000], reward_mode=_cabi.SB_REWARD_BERNOULLI, arm_p=pd)
        r = EpsilonGreedyB200(n_branches=32, epsilon=0.1, seed=1) if kind == "eps" else ThompsonSamplingB200(n_branches=32, seed=1)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        dec = r.run_batched(T, ud, reward_mode=_cabi.SB_REWARD_BERNOULLI, arm_p=pd)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        print(json.dumps({"component": "router %s C4 (32 arms, seed 1, interleaved route+feedback)" % kind, "calls": T,
                          "value": T / (ms * 1e-3), "unit": "calls/s", "ns_per_call": ms * 1e6 / T, "ms": ms,
                          "decisions_checksum": int(dec.long().sum().item()),
                          "note": "sequential bandit: latency kernel, no roofline"}), flush=True)


if __name__ == "__main__":
    print(json.dumps({"gpu": torch.cuda.get_device_name(0), "hbm_peak_gbs": HBM}))
    which = sys.argv[1:] or ["combiner", "codec", "json", "routers"]
    if "combiner" in which:
        combiner()
    if "codec" in which:
        codec_bench()
    if "json" in which:
        json_bench()
    if "routers" in which:
        routers()
