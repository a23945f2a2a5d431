"""Multi-GPU plumbing: environments shard trivially (one process per GPU, no collective on the step path);
the only exchange is the all-reduce of the episode-statistics vector."""
from __future__ import annotations


def shard_range(total_envs: int, rank: int, world: int):
    """Contiguous block of global env ids owned by `rank`: [lo, hi)."""
    base, rem = divmod(int(total_envs), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def reduce_stats(stats: dict, device):
    """Sum the episode statistics over all ranks (NCCL on GPUs, gloo in the CPU tests)."""
    import torch
    import torch.distributed as dist
    keys = tuple(k for k in ("return_sum", "length_sum", "fatal", "episodes", "overflow_steps", "resets", "steps") if k in stats)
    t = torch.tensor([float(stats[k]) for k in keys], dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return {k: float(v) for k, v in zip(keys, t.tolist())}
