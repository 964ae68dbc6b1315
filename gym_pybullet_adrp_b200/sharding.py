"""Multi-GPU layout of the path: envs are independent, so a job of ``total_envs`` is cut into contiguous
env ranges, one per rank, one process per GPU, **no collective on the step path**.  The Philox subsequence of an
env is its GLOBAL index (``adrp_config.env_offset`` + local index), so results do not depend on the world size.
The only collective is the optional episode-statistics reduce after a rollout."""
from __future__ import annotations

import os


def shard_range(total_envs: int, world_size: int, rank: int) -> tuple[int, int]:
    """(first global env, number of envs) of ``rank``: contiguous, sizes differ by at most one."""
    if not 0 <= rank < world_size:
        raise ValueError("rank out of range")
    base, extra = divmod(int(total_envs), int(world_size))
    count = base + (1 if rank < extra else 0)
    offset = rank * base + min(rank, extra)
    return offset, count


def env_world() -> tuple[int, int, int]:
    """(rank, local_rank, world_size) from the torchrun environment (1-process defaults)."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")))


def reduce_episode_stats(stats, group=None):
    """Sum a small tensor of episode statistics (returns, episode count, crashes, gates passed ...) over the ranks.
    Off the step path; a no-op when torch.distributed is not initialised."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.SUM, group=group)
    return stats
