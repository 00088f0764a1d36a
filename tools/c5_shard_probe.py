"""torchrun probe (N ranks): what a sharded d = 256 batch costs beyond its kernels.  Per step on every rank:
  A  the sharded step as the class API runs it (pass 1 -> all-gather of the records -> finalize, pass 2, commit)
  B  the same with the all-gather replaced by a local copy (every rank uses its own record N times: wrong numbers,
     same kernels) - the difference to A is the collective plus the waiting for the slowest rank
  C  an all-gather of the record alone, back to back
Prints max-over-ranks device times."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, ".")
from seldon_core_b200 import OutlierMahalanobisB200, StandardiseB200, synth  # noqa: E402
from seldon_core_b200 import dist as sdist  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
d, total = 256, 1 << 24
r0, r1 = sdist.shard_bounds(total, world, rank)
rows = r1 - r0
det = OutlierMahalanobisB200(n_components=16, device=dev)
prime, _ = synth.maha_rows_torch(1 << 18, d, 16, 201, dev)
det.score_device(prime)
X, _ = synth.maha_rows_torch(rows, d, 16, 300 + rank, dev)
sh = sdist.ShardedMahalanobis(det)
SL = 1 + d + d * d


def step_a():
    sh.score_shard(X, total)


def step_b():
    local_st = torch.empty(SL, dtype=torch.float64, device=dev)
    det.pass1_device(X, local_st)
    allst = local_st.repeat(world).view(world, SL)
    det.finish_device(X, allst, world, rank, total, r0)


rec = torch.zeros(SL, dtype=torch.float64, device=dev)


def step_c():
    sdist.gather_stats(rec)


def timed(fn, n=20):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / n], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0])


a, b, c = timed(step_a), timed(step_b), timed(step_c, 100)
if rank == 0:
    print("ranks %d, %d rows per rank: sharded step %.3f ms; same kernels without the collective %.3f ms; all-gather of one "
          "%d-byte record per rank alone %.3f ms" % (world, rows, a, b, SL * 8, c), flush=True)
dist.destroy_process_group()
