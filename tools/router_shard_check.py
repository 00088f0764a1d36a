"""torchrun check (NCCL, N ranks): dist.ShardedRouter with the device routers on rank 0 against the C oracle - the
requests of all ranks played in rank order, decisions broadcast, a rank without requests takes part."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, ".")
from oracle import routers as R  # noqa: E402
from seldon_core_b200 import EpsilonGreedyB200, ThompsonSamplingB200, synth  # noqa: E402
from seldon_core_b200 import dist as sdist  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
T, n = 5000, 32
p, u = synth.router_rewards(T, n)
counts = [T // world + (1 if r < T % world else 0) for r in range(world)]
if world > 1:
    counts[1], counts[0] = 0, counts[0] + counts[1]          # rank 1 has no requests
lo = sum(counts[:rank])
ok = True
for kind, cls in ((0, EpsilonGreedyB200), (1, ThompsonSamplingB200)):
    router = None
    if rank == 0:
        router = cls(n_branches=n, epsilon=0.1, seed=1, device=dev) if kind == 0 else cls(n_branches=n, seed=1, device=dev)
    sr = sdist.ShardedRouter(router, arm_p=torch.from_numpy(p).to(dev) if rank == 0 else None)
    mine = sr.play(torch.from_numpy(u[lo:lo + counts[rank]]).to(dev)).cpu().numpy()
    want, _ = R.run_batched(kind, n, 0.1, 1, 0, T, reward=u, reward_mode=R.REWARD_BERNOULLI, arm_p=p)
    good = np.array_equal(mine, want[lo:lo + counts[rank]])
    flag = torch.tensor([int(good)], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    ok = ok and bool(flag.item())
    if rank == 0:
        print("kind %d: decisions of all %d ranks equal the oracle: %s" % (kind, world, bool(flag.item())), flush=True)
dist.destroy_process_group()
sys.exit(0 if ok else 1)
