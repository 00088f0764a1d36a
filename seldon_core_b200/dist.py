"""Row-sharded request batches over the GPUs of one box (one process per GPU, torch.distributed).

The reference has no multi-device path (it scales by replicas, proto/seldon_deployment.proto:53-64).  Here a request
batch is split into contiguous row blocks, one per rank; the only exchange is one all-gather of the per-rank
statistics records (1 + d + d*d doubles: count, row sums, Gram about the start-of-call mean) between pass 1 and the
covariance update - 33 KB per rank at d = 64, 514 KB at d = 256.  Every rank then forms the identical covariance,
eigenvectors and state, and derives the prefix of the ranks before it from the same records, so no
second exchange is needed.  Works with the nccl backend (GPU) and, for the stats algebra, with gloo (CPU tests).
"""
import numpy as np


def shard_bounds(total_rows, world_size, rank):
    """Contiguous block [row0, row1) of `rank`; the first (total % world) ranks get one extra row."""
    base, extra = divmod(int(total_rows), int(world_size))
    row0 = rank * base + min(rank, extra)
    return row0, row0 + base + (1 if rank < extra else 0)


def gather_stats(local_stats, group=None):
    """all-gather of one record per rank -> tensor (world, len) in rank order."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    flat = torch.empty(world * local_stats.numel(), dtype=local_stats.dtype, device=local_stats.device)
    dist.all_gather_into_tensor(flat, local_stats.contiguous().view(-1), group=group)
    return flat.view(world, local_stats.numel())


def combine_stats_reference(all_stats, rank):
    """What sb_maha_finalize does with the gathered records, restated in numpy (host logic / gloo tests):
    totals in rank order and the prefix of the ranks before `rank`."""
    a = np.asarray(all_stats, dtype=np.float64)
    total = np.zeros_like(a[0])
    prefix = None
    for r in range(a.shape[0]):
        if r == rank:
            prefix = total.copy()
        total = total + a[r]
    return total, prefix


class ShardedMahalanobis(object):
    """Scores this rank's row block of a batch of `total_rows` rows; all ranks end in the same detector state."""

    def __init__(self, detector, group=None):
        import torch.distributed as dist

        self.det = detector
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)

    def score_shard(self, X_local, total_rows):
        import torch

        row0, row1 = shard_bounds(total_rows, self.world, self.rank)
        assert X_local.shape[0] == row1 - row0, "shard has %d rows, expected %d" % (X_local.shape[0], row1 - row0)
        d = X_local.shape[1]
        local = torch.empty(self.det.stats_len(d), dtype=torch.float64, device=X_local.device)
        self.det.pass1_device(X_local, local)
        all_stats = gather_stats(local, self.group)
        return self.det.finish_device(X_local, all_stats, self.world, self.rank, int(total_rows), row0)


class ShardedRouter(object):
    """The bandit routers are sequential and are not sharded: every rank holds the requests of its row
    block, rank 0 holds the router.  Per batch of requests: the per-request reward inputs (one float per rank-local
    request) are gathered to rank 0, which plays all requests of all ranks in rank order in one `run_batched` launch,
    and the decisions are broadcast back; every rank returns the decisions of its own requests.  The call sequence it
    serves is route (seldon_methods.py:196-243) followed by send_feedback (L59-103) per request.

    `router` is an EpsilonGreedyB200 / ThompsonSamplingB200 on rank 0 (ignored elsewhere) or any callable
    `(T, rewards float32[T]) -> int32[T]` (the gloo tests pass the C oracle)."""

    def __init__(self, router, group=None, reward_mode=2, arm_p=None):
        import torch.distributed as dist

        self.router, self.group = router, group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.reward_mode, self.arm_p = reward_mode, arm_p

    def play(self, local_rewards):
        """local_rewards: float32 tensor [m], one entry per request of this rank (with reward mode BERNOULLI the
        uniform u_t that decides success against the chosen arm's probability).  Returns int32 [m]."""
        import torch
        import torch.distributed as dist

        dev = local_rewards.device
        m = torch.tensor([local_rewards.numel()], dtype=torch.int64, device=dev)
        counts = torch.empty(self.world, dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(counts, m, group=self.group)
        counts = counts.cpu().tolist()
        mmax, total = max(counts), sum(counts)
        padded = torch.zeros(max(mmax, 1), dtype=torch.float32, device=dev)
        padded[:local_rewards.numel()] = local_rewards.to(torch.float32)
        gathered = [torch.empty_like(padded) for _ in range(self.world)] if self.rank == 0 else None
        dist.gather(padded, gathered, dst=0, group=self.group)
        decisions = torch.empty(max(total, 1), dtype=torch.int32, device=dev)
        if self.rank == 0 and total > 0:
            rew = torch.cat([g[:c] for g, c in zip(gathered, counts)])
            if callable(self.router) and not hasattr(self.router, "run_batched"):
                dec = self.router(total, rew)
            else:
                dec = self.router.run_batched(total, rew, reward_mode=self.reward_mode, arm_p=self.arm_p)
            decisions[:total] = torch.as_tensor(dec, dtype=torch.int32, device=dev)
        dist.broadcast(decisions, src=0, group=self.group)
        lo = sum(counts[:self.rank])
        return decisions[lo:lo + counts[self.rank]]
