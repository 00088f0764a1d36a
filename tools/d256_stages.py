"""Per-stage kernel times of one Mahalanobis batch of 2 Mi rows (usage: d256_stages.py [d ...]; ITERS=n batches)."""
import sys, ctypes, torch
sys.path.insert(0, '/root/repo')
from seldon_core_b200 import OutlierMahalanobisB200, _cabi, synth
dev = torch.device('cuda')
for d in [int(a) for a in sys.argv[1:]] or (128, 256):
    det = OutlierMahalanobisB200(n_components=16)
    det.set_option(_cabi.SB_OPT_PROFILE, 1)
    prime, _ = synth.maha_rows_torch(1 << 18, d, 16, 201, dev)
    X, _ = synth.maha_rows_torch(int(__import__("os").environ.get("ROWS", 1 << 21)), d, 16, 202, dev)
    det.score_device(prime)
    for it in range(int(__import__('os').environ.get('ITERS', 3))):
        det.score_device(X)
        km = (ctypes.c_float * 5)()
        _cabi.check(_cabi.lib().sb_maha_kernel_times(det._h, km, torch.cuda.current_stream().cuda_stream))
        c = (ctypes.c_int64 * 8)()
        _cabi.check(_cabi.lib().sb_maha_debug_counters(det._h, c, torch.cuda.current_stream().cuda_stream))
        print("d=%d rows=%d: pass1 %.2f eigh %.2f prefix %.2f project %.2f chain %.2f ms sweeps=%d nsub=%d" % (d, X.shape[0], km[0], km[1], km[2], km[3], km[4], c[3], c[5]))
