"""Diagnostics for a GPU box: per-fixture error profile and coarse timings (not a test, not the bench)."""
import glob
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from seldon_core_b200 import OutlierMahalanobisB200, _cabi, synth  # noqa: E402


def rel(a, b):
    return np.abs(a - b) / np.maximum(np.abs(b), 1e-6)


def fixtures():
    for path in sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "maha_*.npz"))):
        g = np.load(path)
        kw = {str(k): (int(v) if float(v).is_integer() else float(v)) for k, v in zip(g["kw_keys"], g["kw_vals"])}
        for generic in (0, 1):
            d = g["X0"].shape[1]
            if generic and d != 64:
                continue
            det = OutlierMahalanobisB200(**kw)
            det.set_option(_cabi.SB_OPT_FORCE_GENERIC, generic)
            for i in range(int(g["n_batches"])):
                if i > 3:
                    break
                try:
                    pred = det.predict(g["X%d" % i], None)
                except Exception as e:  # noqa: BLE001
                    print(os.path.basename(path), "generic", generic, "batch", i, "FAILED:", e)
                    break
                s, r = det.score, g["score%d" % i]
                e = rel(s, r)
                fo = int(np.flatnonzero(r)[0]) if np.any(r) else len(r)
                fo_g = int(np.flatnonzero(s)[0]) if np.any(s) else len(s)
                det._last_sums()
                segs = [(fo, fo + 50), (fo + 50, fo + 200), (fo + 200, len(r))]
                prof = " ".join("%d-%d:%.1e" % (a, min(b, len(r)), e[a:b].max()) for a, b in segs if a < len(r))
                print("%-22s gen=%d b%d nb=%5d first_open ref=%d gpu=%d gate_row=%d | %s | flags!= %d  nan=%d" % (
                    os.path.basename(path)[5:-4], generic, i, len(r), fo, fo_g, det.gate_row, prof,
                    int(np.sum(pred != g["pred%d" % i])), int(np.isnan(s).sum())))
            st = det._state()
            print("    state: n=%d  C err %.1e  mean err %.1e" % (
                st["n"], np.max(np.abs(st["C"] - g["C"])) / np.max(np.abs(g["C"])),
                np.max(np.abs(st["mean"] - g["mean"])) / max(np.max(np.abs(g["mean"])), 1e-30)))


def timings():
    dev = torch.device("cuda")
    for nb in (1 << 20, 1 << 24):
        prime, _ = synth.maha_rows_torch(1 << 20, 64, 16, 101, dev)
        X, labels = synth.maha_rows_torch(nb, 64, 16, 102, dev)
        for terms in (3, 1):
            det = OutlierMahalanobisB200(n_components=16)
            det.set_option(_cabi.SB_OPT_GRAM_TERMS, terms)
            det.score_device(prime)
            torch.cuda.synchronize()
            ts = []
            for _ in range(4):
                t0 = time.time()
                s, f = det.score_device(X)
                torch.cuda.synchronize()
                ts.append(time.time() - t0)
            print("nb=%d terms=%d batch times ms: %s  -> %.2f G rows/s; flags==labels: %s" % (
                nb, terms, ["%.2f" % (t * 1e3) for t in ts], nb / min(ts) / 1e9,
                bool(torch.equal(f.to(torch.int8), labels))))


def stages():
    """CUDA-event time of each C-ABI stage of one 16 Mi x 64 batch."""
    import ctypes

    dev = torch.device("cuda")
    L = _cabi.lib()
    for nb in (1 << 20, 1 << 24):
        prime, _ = synth.maha_rows_torch(1 << 20, 64, 16, 101, dev)
        X, _ = synth.maha_rows_torch(nb, 64, 16, 102, dev)
        det = OutlierMahalanobisB200(n_components=16)
        det.score_device(prime)
        h = det._h
        ws = det._workspace(nb)
        score = torch.empty(nb, dtype=torch.float32, device=dev)
        flag = torch.empty(nb, dtype=torch.int32, device=dev)
        st = torch.cuda.current_stream().cuda_stream
        for it in range(3):
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
            ev[0].record()
            _cabi.check(L.sb_maha_pass1_stats(h, X.data_ptr(), nb, 64, ws.data_ptr(), ws.numel(), None, st))
            ev[1].record()
            _cabi.check(L.sb_maha_finalize(h, None, 1, 0, nb, 0, ws.data_ptr(), ws.numel(), st))
            ev[2].record()
            _cabi.check(L.sb_maha_pass2_score(h, X.data_ptr(), nb, 64, score.data_ptr(), flag.data_ptr(), ws.data_ptr(),
                                              ws.numel(), st))
            ev[3].record()
            _cabi.check(L.sb_maha_commit(h, st))
            ev[4].record()
            torch.cuda.synchronize()
            c = (ctypes.c_int64 * 8)()
            _cabi.check(L.sb_maha_debug_counters(h, c, st))
            print("nb=%d it=%d pass1+reduce %.3f ms | finalize(cov+eigh+prefix) %.3f ms | pass2 %.3f ms | commit %.3f ms"
                  " | sweeps=%d nsub=%d sublen=%d bstar=%d" % (
                      nb, it, ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2]), ev[2].elapsed_time(ev[3]),
                      ev[3].elapsed_time(ev[4]), c[3], c[5], c[6], c[0]))


if __name__ == "__main__":
    print(torch.cuda.get_device_name(0))
    if "--stages" in sys.argv:
        stages()
        sys.exit(0)
    fixtures()
    timings()
