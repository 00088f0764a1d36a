#!/usr/bin/env python
"""Benchmark of the hot path: rows/s scored through the graph `Mahalanobis outlier detector -> combiner`
(BASELINE.json `metric`; the workload is BASELINE config 2: 16 Mi x 64 fp32 rows, n_components = 16, one B200).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

One "step" = one request batch through the graph: pass-1 statistics -> covariance + eigenvectors -> pass-2 scores and
flags -> state commit (CoreMahalanobis._get_preds), then the combiner's mean over the 4 child outputs (width 8) of the
same rows.  With N > 1 ranks the request batch is N x 16 Mi rows, row-sharded, one all-gather of the (count, sum,
Gram) statistics per batch (weak scaling); the combiner is rank-local.

Printed JSON (rank 0, one line):
  value     rows/s with the batch already resident in HBM (CUDA events, max over ranks)
  e2e       the same graph from request BYTES to response BYTES through the class API (`predict_raw` /
            `aggregate_raw`): serialised SeldonMessage in pinned host memory -> device decode -> detector ->
            combiner -> device encode -> response bytes in pinned host memory
  roofline  the dominant kernel of the step against the measured HBM peak, plus one line per kernel
  cpu_baseline  the reference's CPU path timed on this box's host cores on a bounded sample (N = 1 only)
  c5 / c3 / c4 / codec   the other BASELINE configurations and the codec kernels, measured in the same run
  peaks     TF32 dense tensor peak measured here the way MEASURED_PEAKS.json measured bf16
`--impl reference` times the reference's own CPU implementation (unmodified CoreMahalanobis.py from oracle/_ref
when present, else the oracle's port) and prints the same line shape with "impl": "reference".
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ROWS_PER_GPU = 1 << 24
D, K_COMP = 64, 16
CHILDREN, WIDTH = 4, 8          # combiner of the graph: 4 child outputs of width 8 per row
PRIME_ROWS = 1 << 20
UNIT = "rows/s"
CPU_SAMPLE_ROWS = 100000
CPU_PRIME_ROWS = 20000


def baseline_metric():
    try:
        with open(os.path.join(ROOT, "BASELINE.json")) as f:
            return json.load(f)["metric"]
    except Exception:  # noqa: BLE001
        return "rows/sec scored (Mahalanobis+combiner graph) at 1/2/4/8 B200; % HBM roofline"


METRIC = baseline_metric()


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(path):
        with open(path) as f:
            return json.load(f), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks and throttle reasons sampled every 200 ms while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.lines = []
        self.proc = None
        self.gpu_index = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu_index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 9:
                continue
            try:
                sm.append(float(p[1]))
                smax.append(float(p[2]))
            except ValueError:
                continue
            for nm, v in zip(names, p[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------------------------------- CPU arm
def host_threads():
    try:
        from threadpoolctl import threadpool_info

        return int(max([p.get("num_threads", 1) for p in threadpool_info()] or [1]))
    except Exception:  # noqa: BLE001
        return int(os.cpu_count() or 1)


def cpu_graph(steps=1, warmup=0, budget_s=150.0):
    """rows/s of the reference's CPU implementation of the graph on a bounded sample of config 2:
    `CoreMahalanobis.predict` (the UNMODIFIED reference file from oracle/_ref or /root/reference when present,
    kind "reference"; else the oracle's line-by-line numpy port, kind "port") on CPU_SAMPLE_ROWS rows per step after a
    CPU_PRIME_ROWS priming call, followed by the combiner mean over 4 child outputs of width 8 of the same rows (the
    N-child rule of AverageCombinerUnit.java:50-71 as restated in oracle/combiner.py; MnistCombiner.py is the 2-child
    case).  The reference materialises (nb, 64, 64) fp64, so the 16 Mi-row batch (550 GB) cannot run."""
    import numpy as np

    from oracle import ref_shim
    from oracle.combiner import mean_combiner
    from seldon_core_b200 import synth

    if ref_shim.available():
        det = ref_shim.core_mahalanobis_class()(n_components=K_COMP)
        kind = "reference"
        what = "unmodified CoreMahalanobis.predict (%s)" % ("oracle/_ref copy" if "oracle" in ref_shim.REFERENCE_ROOT else ref_shim.REFERENCE_ROOT)
    else:
        from oracle.mahalanobis import MahalanobisPort

        det = MahalanobisPort(n_components=K_COMP)
        kind = "port"
        what = "oracle.MahalanobisPort (numpy, same operations as CoreMahalanobis._get_preds)"
    prime, _ = synth.maha_rows_numpy(CPU_PRIME_ROWS, D, K_COMP, seed=101)
    det.predict(prime.astype(np.float64), None)
    X, _ = synth.maha_rows_numpy(CPU_SAMPLE_ROWS, D, K_COMP, seed=102)
    X = X.astype(np.float64)
    kids = [synth.softmax_rows_numpy(CPU_SAMPLE_ROWS, WIDTH, 900 + j).astype(np.float64) for j in range(CHILDREN)]
    times, t_start = [], time.perf_counter()
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        det.predict(X, None)
        mean_combiner(kids)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
        if time.perf_counter() - t_start > budget_s and times:
            break  # bounded: the run must end within a few minutes
    total = sum(times)
    info = {"value": CPU_SAMPLE_ROWS * len(times) / total, "unit": UNIT, "cores": host_threads(), "kind": kind,
            "sample": "%s + 4-child combiner mean on %d rows x %d of the config-2 generator per step after a %d-row "
                      "priming call; the Python loop of the reference is single-threaded, BLAS uses `cores` threads"
                      % (what, CPU_SAMPLE_ROWS, D, CPU_PRIME_ROWS)}
    return info, total / len(times), len(times)


def cpu_restatement_full(rows=ROWS_PER_GPU):
    """The lean fp64 restatement (oracle/maha_lean.c; never materialises (nb, p, p)) on the FULL
    config-2 batch - the one CPU number measured on the same configuration as the GPU arm.  Single thread."""
    import numpy as np

    from oracle.combiner import mean_combiner
    from oracle.mahalanobis import MahalanobisLean
    from seldon_core_b200 import synth

    try:
        import psutil

        if psutil.virtual_memory().available < 3 * rows * D * 8:
            rows = int(psutil.virtual_memory().available // (3 * D * 8)) // 1024 * 1024
    except Exception:  # noqa: BLE001
        pass
    det = MahalanobisLean(n_components=K_COMP)
    prime, _ = synth.maha_rows_numpy(CPU_PRIME_ROWS, D, K_COMP, seed=101)
    det.predict(prime.astype(np.float64), None)
    X = np.empty((rows, D), dtype=np.float64)
    chunk = 1 << 20
    for lo in range(0, rows, chunk):
        m = min(chunk, rows - lo)
        X[lo:lo + m] = synth.maha_rows_numpy(m, D, K_COMP, seed=5000 + lo // chunk)[0]
    kids = [np.full((rows, WIDTH), 1.0 / WIDTH) for _ in range(CHILDREN)]
    t0 = time.perf_counter()
    det.predict(X, None)
    mean_combiner(kids)
    dt = time.perf_counter() - t0
    return {"value": rows / dt, "unit": UNIT, "rows": rows, "seconds": dt, "cores": 1, "kind": "restatement",
            "same_config": rows == ROWS_PER_GPU,
            "what": "oracle/maha_lean.c (fp64 closed form of CoreMahalanobis._get_preds, one thread) + combiner mean "
                    "on the full %d x %d batch" % (rows, D)}


def run_reference(args, rank, world):
    if rank != 0:
        return
    # torchrun pins OMP_NUM_THREADS=1 for its workers; the CPU arm is meant to use every host thread it can, and
    # numpy / BLAS have not been loaded yet in this process
    if "numpy" not in sys.modules and os.environ.get("OMP_NUM_THREADS") == "1" and world > 1:
        for var in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
            os.environ[var] = str(os.cpu_count() or 1)
    steps = max(1, args.steps)
    info, sec_per_step, done = cpu_graph(steps=steps, warmup=min(args.warmup, 1))
    line = {"impl": "reference", "metric": METRIC, "value": info["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": done, "warmup": min(args.warmup, 1), "ms_per_step": sec_per_step * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "BASELINE config 2 graph (Mahalanobis 16Mi x 64, n_components=16 -> 4-child combiner), "
                                   "CPU arm on a bounded sample of %d rows per step" % CPU_SAMPLE_ROWS,
                       "sample_rows_per_step": CPU_SAMPLE_ROWS},
            "cpu_baseline": info,
            "e2e": {"value": info["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    if world == 1 and not args.no_restatement:
        try:
            line["restatement"] = cpu_restatement_full()
        except Exception as e:  # noqa: BLE001
            line["restatement"] = {"error": repr(e)}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------- GPU helpers
def timed_events(fn, warm=3, reps=10):
    """(best, median) ms of fn() with CUDA events on torch's current stream (the stream our kernels launch on)."""
    import torch

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


def hbm_line(nbytes, ms, hbm):
    gbs = nbytes / (ms * 1e-3) / 1e9
    return {"ms": ms, "achieved": gbs, "unit": "GB/s", "frac": gbs / hbm, "algorithmic_bytes": nbytes}


def tf32_peak():
    """Dense TF32 peak the way MEASURED_PEAKS.json measured bf16: torch.matmul 8192^3 with allow_tf32, best of 10
    (burst) and back to back for ~1.5 s (sustained)."""
    import torch

    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        n = 8192
        a = torch.randn((n, n), device="cuda")
        b = torch.randn((n, n), device="cuda")
        c = torch.empty((n, n), device="cuda")
        best, _ = timed_events(lambda: torch.matmul(a, b, out=c), warm=3, reps=10)
        reps = max(10, int(1500.0 / best))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            torch.matmul(a, b, out=c)
        e1.record()
        torch.cuda.synchronize()
        sus = e0.elapsed_time(e1) / reps
        fl = 2.0 * n ** 3
        return {"tf32_tflops": fl / (best * 1e-3) / 1e12, "tf32_tflops_sustained": fl / (sus * 1e-3) / 1e12,
                "how": "torch.matmul fp32 8192^3 with allow_tf32 (cuBLAS): best of 10, and %d back to back" % reps}
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old


def block_c3(dev, hbm, world):
    """BASELINE config 3: combiner mean over 8 children of 1 Mi x 1000 fp32 (each rank its own copy: weak)."""
    import torch

    from seldon_core_b200 import MeanCombinerB200

    rows, cols, n = 1 << 20, 1000, 8
    g = torch.Generator(device=dev).manual_seed(1)
    kids = [torch.softmax(torch.randn((rows, cols), generator=g, device=dev), dim=1) for _ in range(n)]
    out = MeanCombinerB200.aggregate_device(kids)
    ref = torch.stack([k[:2048].double() for k in kids]).sum(0) / n
    err = float((out[:2048].double() - ref).abs().div(ref.abs().clamp_min(1e-30)).max())
    best, med = timed_events(lambda: MeanCombinerB200.aggregate_device(kids), warm=3, reps=10)
    r = hbm_line(4 * (n + 1) * rows * cols, med, hbm)
    r.update({"workload": "8 children x 1Mi x 1000 fp32 per GPU", "value": rows * world / (med * 1e-3), "unit_value": "rows/s",
              "ms_best": best, "max_rel_err_vs_fp64": err, "kernel": "combiner_mean_vec4"})
    return r


def block_c5(dev, hbm, rank, world, tf32):
    """BASELINE config 5 graph: standardise (fused into the detector's loads) -> Mahalanobis d = 256, k = 16 ->
    4-child combiner, ONE 16 Mi-row request batch split contiguously over the ranks (strong scaling)."""
    import torch

    from seldon_core_b200 import MeanCombinerB200, OutlierMahalanobisB200, StandardiseB200, _cabi, synth
    from seldon_core_b200 import dist as sdist

    d5, total = 256, 1 << 24
    r0, r1 = sdist.shard_bounds(total, world, rank)
    rows = r1 - r0
    g = torch.Generator(device=dev).manual_seed(7)
    mu = torch.randn(d5, generator=g, device=dev).double().cpu().numpy()
    sigma = (0.5 + torch.rand(d5, generator=g, device=dev)).double().cpu().numpy()
    std = StandardiseB200(mu, sigma, device=dev)
    det = OutlierMahalanobisB200(n_components=K_COMP, device=dev)
    det.set_option(_cabi.SB_OPT_PROFILE, 1)
    kids = [torch.softmax(torch.randn((rows, WIDTH), generator=g, device=dev), dim=1) for _ in range(CHILDREN)]
    prime, _ = synth.maha_rows_torch(1 << 18, d5, K_COMP, 201, dev)
    det.score_device(prime)
    del prime
    det.fuse_standardise(std)
    Xraw, labels = synth.maha_rows_torch(rows, d5, K_COMP, 300 + rank, dev)
    Xraw.mul_(torch.tensor(sigma, dtype=torch.float32, device=dev) + 1e-10).add_(torch.tensor(mu, dtype=torch.float32, device=dev))
    sharded = sdist.ShardedMahalanobis(det) if world > 1 else None

    def batch():
        s, f = sharded.score_shard(Xraw, total) if sharded is not None else det.score_device(Xraw)
        return s, f, MeanCombinerB200.aggregate_device(kids)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            import torch.distributed as dist

            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(5):
        s, f, _ = batch()
    barrier()
    mism = int((f.to(torch.int8) != labels).sum().item())
    nb = 20
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(nb):
        batch()
    e1.record()
    barrier()
    km = (ctypes.c_float * 5)()
    _cabi.check(_cabi.lib().sb_maha_kernel_times(det._h, km, torch.cuda.current_stream().cuda_stream))
    t = torch.tensor([e0.elapsed_time(e1), float(mism)], dtype=torch.float64, device=dev)
    if world > 1:
        import torch.distributed as dist

        tm = t.clone()
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        ts = t.clone()
        dist.all_reduce(ts, op=dist.ReduceOp.SUM)
        t = torch.stack([tm[0], ts[1]])
    ms = float(t[0]) / nb
    bytes_row = (8 * d5 + 8) + 4 * WIDTH * (CHILDREN + 1)
    gbs = bytes_row * total / (ms * 1e-3) / 1e9 / world
    out = {"workload": "standardise (fused) -> Mahalanobis d=256 k=16 -> 4-child combiner, one 16Mi-row request batch "
                       "split over %d rank(s)" % world,
           "scaling": "strong", "value": total / (ms * 1e-3), "unit": UNIT, "ms_per_batch": ms,
           "flags_vs_planted_labels_mismatches": int(t[1]),
           "roofline": {"bound": "hbm", "algorithmic_bytes_per_row": bytes_row, "achieved_per_gpu": gbs, "peak": hbm,
                        "unit": "GB/s", "frac": gbs / hbm},
           "kernel_ms_rank0": {"pass1_gram": km[0], "eigh": km[1], "prefix": km[2], "project": km[3], "chain": km[4]}}
    if tf32 and km[0] > 0:
        fl = 2.0 * d5 * d5 * 3 * rows  # 3xTF32: three tensor products per Gram element
        out["pass1_tensor"] = {"bound": "tensor", "flops_per_row": 2 * d5 * d5 * 3, "achieved": fl / (km[0] * 1e-3) / 1e12,
                               "peak": tf32["tf32_tflops"], "unit": "TFLOP/s",
                               "frac": fl / (km[0] * 1e-3) / 1e12 / tf32["tf32_tflops"],
                               "hbm_frac": 4 * d5 * rows / (km[0] * 1e-3) / 1e9 / hbm}
    return out


def block_c4(dev):
    """BASELINE config 4: both routers, 32 arms, seed 1, strictly interleaved route + send_feedback; batched (all
    calls in one launch; a bounded prefix of the 10 M-call configuration, the full size is a test) and call by call
    through route() / send_feedback() of the class API."""
    import numpy as np
    import torch

    from seldon_core_b200 import EpsilonGreedyB200, ThompsonSamplingB200, _cabi, synth

    out = {}
    for name, cls, T in (("epsilon_greedy", EpsilonGreedyB200, 4_000_000), ("thompson", ThompsonSamplingB200, 200_000)):
        p, u = synth.router_rewards(T, 32)
        ud, pd = torch.from_numpy(u).to(dev), torch.from_numpy(p).to(dev)
        mk = (lambda: cls(n_branches=32, epsilon=0.1, seed=1, device=dev)) if cls is EpsilonGreedyB200 else (
            lambda: cls(n_branches=32, seed=1, device=dev))
        r = mk()
        r.run_batched(1000, ud[:1000], reward_mode=_cabi.SB_REWARD_BERNOULLI, arm_p=pd)
        r = mk()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        dec = r.run_batched(T, ud, reward_mode=_cabi.SB_REWARD_BERNOULLI, arm_p=pd)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        # call by call: the API the wrapper uses (one route, one feedback per request)
        r2 = mk()
        feats = np.zeros((1, 4))
        n_calls = 3000
        decs = dec[:n_calls].cpu().numpy()
        for t in range(200):
            a = r2.route(feats, None)
            r2.send_feedback(feats, None, a, 1.0 if u[t] < p[a] else 0.0, None)
        t0 = time.perf_counter()
        same = True
        for t in range(200, n_calls):
            a = r2.route(feats, None)
            same &= a == int(decs[t])
            r2.send_feedback(feats, None, a, 1.0 if u[t] < p[a] else 0.0, None)
        dt = time.perf_counter() - t0
        out[name] = {"batched": {"calls": T, "value": T / (ms * 1e-3), "unit": "calls/s", "ns_per_call": ms * 1e6 / T},
                     "call_by_call": {"calls": n_calls - 200, "value": (n_calls - 200) / dt, "unit": "calls/s",
                                      "us_per_call": dt * 1e6 / (n_calls - 200),
                                      "equal_to_batched_decisions": bool(same)},
                     "counters": r.counters()}
    out["note"] = "sequential bandits: latency kernels, no roofline; reference classes: 23 us (eps-greedy) / 51 us (Thompson) per call"
    return out


def block_codec(dev, hbm):
    import numpy as np
    import torch

    from seldon_core_b200 import _cabi

    L = _cabi.lib()
    st = torch.cuda.current_stream().cuda_stream
    out = {}
    rows, cols = 1 << 22, 64
    x = torch.randn((rows, cols), device=dev)
    wire = torch.empty(rows * cols * 8, dtype=torch.uint8, device=dev)
    y = torch.empty_like(x)
    f_pack = lambda: _cabi.check(L.sb_pack_tensor_f64(x.data_ptr(), 0, rows * cols, wire.data_ptr(), st))
    f_unpack = lambda: _cabi.check(L.sb_unpack_tensor_f64(wire.data_ptr(), wire.numel(), 0, rows, cols, None, None,
                                                          y.data_ptr(), cols, st))
    _, med = timed_events(f_pack)
    out["pack_tensor"] = hbm_line(12 * rows * cols, med, hbm)
    _, med = timed_events(f_unpack)
    assert torch.equal(x, y)
    out["unpack_tensor"] = hbm_line(12 * rows * cols, med, hbm)
    rows = 1 << 21
    x = torch.randn((rows, cols), device=dev)
    nb = int(L.sb_listvalue_nbytes(rows, cols))
    wire = torch.empty(nb, dtype=torch.uint8, device=dev)
    y = torch.empty_like(x)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    f_pack = lambda: _cabi.check(L.sb_pack_listvalue(x.data_ptr(), 0, rows, cols, wire.data_ptr(), st))
    f_unpack = lambda: _cabi.check(L.sb_unpack_listvalue(wire.data_ptr(), nb, rows, cols, y.data_ptr(), cols,
                                                         status.data_ptr(), st))
    _, med = timed_events(f_pack)
    out["pack_listvalue"] = hbm_line(nb + 4 * rows * cols, med, hbm)
    _, med = timed_events(f_unpack)
    assert torch.equal(x, y) and int(status.item()) == 0
    out["unpack_listvalue"] = hbm_line(nb + 4 * rows * cols, med, hbm)
    # REST JSON text of an ndarray -> fp32
    block_rows, tiles = 4096, 64
    a = np.random.RandomState(0).randn(block_rows, cols).astype(np.float32).astype(np.float64)
    inner = json.dumps(a.tolist())[1:-1]
    text = ("[" + ", ".join([inner] * tiles) + "]").encode()
    rows = block_rows * tiles
    buf = torch.frombuffer(bytearray(text), dtype=torch.uint8).to(dev)
    nbytes = len(text)
    ws = torch.empty((int(L.sb_json_workspace_bytes(nbytes)) + 7) // 8, dtype=torch.int64, device=dev)
    counts = torch.zeros(4, dtype=torch.int64, device=dev)
    y = torch.empty((rows, cols), dtype=torch.float32, device=dev)

    def run():
        _cabi.check(L.sb_json_ndarray_count(buf.data_ptr(), nbytes, ws.data_ptr(), ws.numel() * 8, counts.data_ptr(), st))
        _cabi.check(L.sb_unpack_json_ndarray(buf.data_ptr(), nbytes, rows, cols, ws.data_ptr(), y.data_ptr(), cols,
                                             status.data_ptr(), st))

    _, med = timed_events(run)
    assert int(status.item()) == 0
    out["unpack_json_ndarray"] = hbm_line(nbytes + 4 * rows * cols, med, hbm)
    out["unpack_json_ndarray"]["values_per_s"] = rows * cols / (med * 1e-3)
    # fp32 matrix -> the text json.dumps writes (shortest round-trip decimals)
    x = torch.randn((rows, cols), device=dev)
    cap = int(L.sb_json_text_capacity(rows, cols))
    text = torch.empty(cap, dtype=torch.uint8, device=dev)
    ews = torch.empty((int(L.sb_json_encode_workspace_bytes(rows, cols)) + 7) // 8, dtype=torch.int64, device=dev)
    nb_out = torch.zeros(1, dtype=torch.int64, device=dev)
    f_enc = lambda: _cabi.check(L.sb_pack_json_ndarray(x.data_ptr(), 0, rows, cols, ews.data_ptr(), ews.numel() * 8,
                                                       text.data_ptr(), nb_out.data_ptr(), st))
    _, med = timed_events(f_enc)
    tlen = int(nb_out.item())
    out["pack_json_ndarray"] = hbm_line(tlen + 4 * rows * cols, med, hbm)
    out["pack_json_ndarray"]["values_per_s"] = rows * cols / (med * 1e-3)
    out["pack_json_ndarray"]["text_bytes_per_value"] = tlen / (rows * cols)
    sizes = {"pack_tensor": 1 << 28, "unpack_tensor": 1 << 28, "pack_listvalue": 1 << 27, "unpack_listvalue": 1 << 27}
    for k, v in out.items():
        if "values_per_s" not in v:
            v["values_per_s"] = sizes[k] / (v["ms"] * 1e-3)
    return out


def block_latency(dev):
    """Request latency of the wire path: `predict_raw` on small serialised requests (1, 100 and 10 000 rows x 64, f64
    Tensor), wall clock per call, median of 200 - the regime the reference's wrapper serves most of the time."""
    import numpy as np

    from seldon_core_b200 import OutlierMahalanobisB200, synth, wire

    det = OutlierMahalanobisB200(n_components=K_COMP, device=dev)
    prime, _ = synth.maha_rows_numpy(20000, D, K_COMP, seed=101)
    det.predict(prime, None)
    out = {}
    for rows in (1, 100, 10000):
        X, _ = synth.maha_rows_numpy(rows, D, K_COMP, seed=7 + rows)
        req = wire.make_tensor_request_pinned(X.astype(np.float64)).tobytes()
        for _ in range(20):
            det.predict_raw(req)
        ts = []
        for _ in range(200):
            t0 = time.perf_counter()
            det.predict_raw(req)
            ts.append(time.perf_counter() - t0)
        ts.sort()
        out["rows_%d" % rows] = {"p50_us": ts[100] * 1e6, "p99_us": ts[197] * 1e6, "request_bytes": len(req)}
    out["note"] = "bytes in (pageable) -> bytes out, one request at a time, state updated per request like the reference"
    return out


def guarded(fn, *a):
    import torch

    try:
        r = fn(*a)
    except Exception as e:  # noqa: BLE001  - a failing side block must not lose the headline line
        r = {"error": repr(e)[:300]}
    torch.cuda.empty_cache()
    return r


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rows", type=int, default=ROWS_PER_GPU, help="rows per GPU (default: config 2)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-blocks", action="store_true", help="skip the c5 / c3 / c4 / codec blocks")
    ap.add_argument("--no-restatement", action="store_true", help="reference arm: skip the full-size lean restatement")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch

    from seldon_core_b200 import MeanCombinerB200, OutlierMahalanobisB200, _cabi, synth
    from seldon_core_b200 import dist as sdist

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback)"
    if args.warmup < 3:
        args.warmup = 3  # timing rules: at least 3 warm-up steps (the printed line carries the value used)
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    distributed = world > 1
    if distributed:
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    L = _cabi.lib()
    rows = args.rows
    total_rows = rows * world
    peaks, which = measured_peaks()
    hbm = float(peaks["hbm_gbs"])

    # synthetic request batch (config-2 recipe), this rank's contiguous row block; primed state identical on all ranks
    prime, _ = synth.maha_rows_torch(PRIME_ROWS, D, K_COMP, 101, dev)
    X, labels = synth.maha_rows_torch(rows, D, K_COMP, 102 + rank, dev)
    g = torch.Generator(device=dev).manual_seed(900 + rank)
    kids = [torch.softmax(torch.randn((rows, WIDTH), generator=g, device=dev), dim=1) for _ in range(CHILDREN)]
    det = OutlierMahalanobisB200(n_components=K_COMP, device=dev)
    det.set_option(_cabi.SB_OPT_PROFILE, 1)
    det.score_device(prime)
    del prime
    sharded = sdist.ShardedMahalanobis(det) if distributed else None
    ce0, ce1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def step():
        s, f = sharded.score_shard(X, total_rows) if distributed else det.score_device(X)
        ce0.record()
        out = MeanCombinerB200.aggregate_device(kids)
        ce1.record()
        return s, f, out

    def barrier():
        torch.cuda.synchronize()
        if distributed:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(args.warmup):
        s, f, out = step()
    barrier()
    assert torch.equal(f.to(torch.int8), labels), "flags differ from the planted labels"
    ref = torch.stack([k[:4096].double() for k in kids]).sum(0) / CHILDREN
    assert float((out[:4096].double() - ref).abs().max()) <= 1e-6, "combiner output differs from the fp64 mean"

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = L.sb_launch_count()
    t_host0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    elapsed_ms = e0.elapsed_time(e1)
    wall_s = time.perf_counter() - t_host0
    launches = L.sb_launch_count() - launches0
    # per-kernel device time of the last timed step (events recorded by the library on the same stream)
    km = (ctypes.c_float * 5)()
    _cabi.check(L.sb_maha_kernel_times(det._h, km, torch.cuda.current_stream().cuda_stream))
    kern_ms = [float(v) for v in km]
    comb_ms = ce0.elapsed_time(ce1)
    clocks = sampler.stop() if rank == 0 else None

    t = torch.tensor([elapsed_ms, float(launches)], dtype=torch.float64, device=dev)
    if distributed:
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone()
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        elapsed_ms, launches = float(tmax[0]), int(tsum[1])
    ms_per_step = elapsed_ms / args.steps
    value = total_rows / (ms_per_step * 1e-3)

    # ---- end to end, bytes in -> bytes out through the class API
    e2e = None
    if not args.no_e2e:
        e2e = guarded(e2e_wire, det, sharded, kids, rows, total_rows, world, dev, labels, X, args, barrier)
        if distributed and "value" in e2e:
            te = torch.tensor([e2e["seconds"]], dtype=torch.float64, device=dev)
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
            e2e["value"] = total_rows * e2e["steps"] / float(te[0])
            e2e["seconds"] = float(te[0])
    del X, kids
    torch.cuda.empty_cache()

    blocks = {}
    if not args.no_blocks:
        tf32 = guarded(tf32_peak) if rank == 0 or True else None
        blocks["peaks"] = tf32
        blocks["c5"] = guarded(block_c5, dev, hbm, rank, world, tf32 if "tf32_tflops" in tf32 else None)
        blocks["c3"] = guarded(block_c3, dev, hbm, world)
        if rank == 0:
            blocks["c4"] = guarded(block_c4, dev)
            blocks["codec"] = guarded(block_codec, dev, hbm)
            blocks["latency"] = guarded(block_latency, dev)
        if distributed:
            barrier()

    if rank == 0:
        pass2_bytes = (4 * D + 8) * rows          # read X once, write score + flag
        pass1_bytes = 4 * D * rows
        comb_bytes = 4 * WIDTH * (CHILDREN + 1) * rows
        graph_bytes = (8 * D + 8) * rows + comb_bytes
        tpath = os.path.join(ROOT, "profiles", "roofline_traffic.json")
        pass2_ms = kern_ms[3] + kern_ms[4]
        # the dominant kernel of the step: pass-1 Gram statistics or pass 2 (projection + chain, one algorithmic unit)
        if kern_ms[0] >= pass2_ms:
            dom = {"kernel": "pass1_gram_umma<64> (clip, row sums, 3xTF32 Gram on tcgen05.mma kind::tf32)", "ms": kern_ms[0],
                   "bytes": pass1_bytes, "traffic_key": "pass1_dram_bytes_per_launch"}
        else:
            dom = {"kernel": "pass 2 = project_rows<1,clip> + chain4<clip> (project on mma.sync 3xTF32, inverse chain, score, flag)",
                   "ms": pass2_ms, "bytes": pass2_bytes, "traffic_key": "pass2_dram_bytes_per_launch"}
        traffic = None
        if os.path.isfile(tpath):
            with open(tpath) as f:
                traffic = json.load(f).get(dom["traffic_key"])
        dom_gbs = dom["bytes"] / (dom["ms"] * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "BASELINE config 2 graph: Mahalanobis outlier scoring, %d x %d fp32 rows per GPU, "
                                   "n_components=%d, state primed with %d rows -> combiner mean over %d child outputs "
                                   "of width %d" % (rows, D, K_COMP, PRIME_ROWS, CHILDREN, WIDTH),
                       "rows_per_gpu": rows, "d": D, "n_components": K_COMP, "request_batch_rows": total_rows,
                       "combiner_children": CHILDREN, "combiner_width": WIDTH,
                       "parallelism": "rows sharded over %d rank(s); one all-gather of (count,sum,Gram) per batch"
                                      % world if distributed else "single GPU",
                       "l2": "inputs (%.1f GB per GPU) are larger than L2 (126 MB); no flush needed" % (rows * D * 4 / 1e9)},
            "roofline": {"bound": "hbm", "kernel": dom["kernel"],
                         "achieved": dom_gbs, "peak": hbm, "unit": "GB/s", "frac": dom_gbs / hbm,
                         "traffic": traffic, "peak_source": "of " + which,
                         "algorithmic_bytes_per_launch": dom["bytes"],
                         "kernel_ms": {"pass1_gram": kern_ms[0], "eigh": kern_ms[1], "prefix": kern_ms[2],
                                       "project": kern_ms[3], "chain": kern_ms[4], "combiner": comb_ms},
                         "pass1": hbm_line(pass1_bytes, kern_ms[0], hbm),
                         "pass2": hbm_line(pass2_bytes, pass2_ms, hbm),
                         "combiner": hbm_line(comb_bytes, comb_ms, hbm),
                         "whole_graph": hbm_line(graph_bytes, ms_per_step, hbm)},
            "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
            "wall_s_timed_region": wall_s,
        }
        line.update(blocks)
        if world == 1 and not args.no_cpu:
            info, _, _ = cpu_graph(steps=1, warmup=0)
            line["cpu_baseline"] = info
        print(json.dumps(line), flush=True)
    if distributed:
        dist.destroy_process_group()


def e2e_wire(det, sharded, kids, rows, total_rows, world, dev, labels, X, args, barrier):
    """bytes in -> bytes out: the request is a serialised SeldonMessage (data.tensor, packed doubles) in pinned host
    memory, the responses are the serialised SeldonMessages of `predict_raw` (flags) and `aggregate_raw` (means) in
    pinned host memory.  Timed with the wall clock between device synchronisations (host work is part of it)."""
    import torch

    from seldon_core_b200 import MeanCombinerB200, wire

    req = wire.make_tensor_request_pinned(X)                # fp64 payload built once, outside the timed region
    kid_reqs = [wire.make_tensor_request_pinned(k) for k in kids]
    comb = MeanCombinerB200(device=dev)
    sh = sharded

    def one():
        r1 = det.predict_raw(req, shard=(sh, total_rows) if sh is not None else None)
        r2 = comb.aggregate_raw(kid_reqs)
        return r1, r2

    for _ in range(2):
        r1, r2 = one()
    barrier()
    flags = wire.parse_tensor_response(r1)
    assert int(flags.sum()) == int(labels.sum()) and flags.shape[0] == rows
    n_e2e = max(2, min(args.steps, 5))
    t0 = time.perf_counter()
    for _ in range(n_e2e):
        one()
    barrier()
    dt = time.perf_counter() - t0
    h2d = (req.nbytes + sum(k.nbytes for k in kid_reqs)) * world
    d2h = (r1.nbytes + r2.nbytes) * world
    return {"value": total_rows * n_e2e / dt, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
            "steps": n_e2e, "seconds": dt,
            "note": "serialised SeldonMessage (Tensor of f64) in pinned host memory -> predict_raw / aggregate_raw -> "
                    "serialised responses in pinned host memory; wall clock between device synchronisations, max over ranks"}


if __name__ == "__main__":
    main()
