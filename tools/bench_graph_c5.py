"""BASELINE config 5: standardise transformer -> Mahalanobis outlier detector (d = 256, k = 16) -> 4-child combiner,
row-sharded over the ranks of one box (torchrun) with one all-gather of the covariance statistics per request batch.

    python tools/bench_graph_c5.py [--rows-per-batch N] [--batches B]
    torchrun --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 tools/bench_graph_c5.py ...

Default is the full configuration: 16 request batches of 16 Mi rows x 256 fp32 (256 Mi rows in total), each batch split
contiguously over the ranks; children = 4 outputs of width 8 per row.  Prints one JSON line (rank 0).
Algorithmic bytes per row: standardise 8*d (unfused) + Mahalanobis 8*d + 8 + combiner 4*8*(4+1).
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from seldon_core_b200 import MeanCombinerB200, OutlierMahalanobisB200, StandardiseB200, synth  # noqa: E402
from seldon_core_b200 import dist as sdist  # noqa: E402

D, K, CHILDREN, WIDTH = 256, 16, 4, 8


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows-per-batch", type=int, default=1 << 24)
    ap.add_argument("--batches", type=int, default=16)
    ap.add_argument("--prime-rows", type=int, default=1 << 18)
    ap.add_argument("--unfused", action="store_true",
                    help="run the standardise transformer as its own pass (default: fused into the detector's loads)")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    total = args.rows_per_batch
    r0, r1 = sdist.shard_bounds(total, world, rank)
    rows = r1 - r0
    g = torch.Generator(device=dev).manual_seed(7)
    mu = torch.randn(D, generator=g, device=dev).double().cpu().numpy()
    sigma = (0.5 + torch.rand(D, generator=g, device=dev)).double().cpu().numpy()
    std = StandardiseB200(mu, sigma, device=dev)
    det = OutlierMahalanobisB200(n_components=K, device=dev)
    sharded = sdist.ShardedMahalanobis(det) if world > 1 else None
    comb_in = [torch.softmax(torch.randn((rows, WIDTH), generator=g, device=dev), dim=1) for _ in range(CHILDREN)]
    # raw inputs = de-standardised config-2-style rows, so the standardised data has the planted structure
    prime, _ = synth.maha_rows_torch(args.prime_rows, D, K, 201, dev)
    det.score_device(prime)
    if not args.unfused:
        det.fuse_standardise(std)  # after priming: the priming rows are already in standardised units
    Xs, labels = synth.maha_rows_torch(rows, D, K, 300 + rank, dev)
    mu_t = torch.tensor(mu, dtype=torch.float32, device=dev)
    sg_t = torch.tensor(sigma, dtype=torch.float32, device=dev)
    Xraw = Xs * (sg_t + 1e-10) + mu_t
    del Xs

    def batch():
        X = std.transform_device(Xraw) if args.unfused else Xraw
        if sharded is not None:
            s, f = sharded.score_shard(X, total)
        else:
            s, f = det.score_device(X)
        out = MeanCombinerB200.aggregate_device(comb_in)
        return s, f, out

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(2):
        s, f, out = batch()
    barrier()
    mism = int((f.to(torch.int8) != labels).sum().item())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.batches):
        batch()
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    mm = torch.tensor([mism], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(mm, op=dist.ReduceOp.SUM)
    if rank == 0:
        sec = float(ms[0]) * 1e-3
        rows_total = total * args.batches
        # standardise as its own pass reads and writes the batch once more (8*d B/row); fused it costs nothing
        bytes_row = (8 * D if args.unfused else 0) + (8 * D + 8) + 4 * WIDTH * (CHILDREN + 1)
        hbm = 6549.4
        p = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.isfile(p):
            hbm = json.load(open(p))["hbm_gbs"]
        gbs = bytes_row * rows_total / sec / 1e9 / world
        print(json.dumps({"config": "C5 graph: standardise -> Mahalanobis(d=256,k=16) -> 4-child combiner (%s)"
                                    % ("standardise as a separate pass" if args.unfused else "standardise fused into the detector"),
                          "n_gpus": world, "rows_per_request_batch": total, "batches": args.batches,
                          "value": rows_total / sec, "unit": "rows/s", "ms_per_batch": sec * 1e3 / args.batches,
                          "flags_vs_planted_labels_mismatches": int(mm[0]),
                          "roofline": {"bound": "hbm", "algorithmic_bytes_per_row": bytes_row,
                                       "achieved_per_gpu": gbs, "peak": hbm, "unit": "GB/s", "frac": gbs / hbm}}),
              flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
