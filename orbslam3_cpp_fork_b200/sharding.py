"""Frame sharding across GPUs: every frame / stereo pair is an independent unit,
so ranks take disjoint frame ranges and there is no data-path collective.  torch.distributed is
used only for the barrier around the timed region, the max-over-ranks time and the frame count."""
from __future__ import annotations


def shard_range(n_frames: int, rank: int, world: int) -> range:
    """Contiguous, balanced shard of [0, n_frames) for `rank` (sizes differ by at most 1)."""
    base, rem = divmod(n_frames, world)
    start = rank * base + min(rank, rem)
    return range(start, start + base + (1 if rank < rem else 0))


def round_robin(n_frames: int, rank: int, world: int) -> range:
    """frame k -> GPU k mod world."""
    return range(rank, n_frames, world)


def reduce_time_and_count(seconds: float, frames: int, device=None):
    """(max over ranks of seconds, sum over ranks of frames); identity when not distributed."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return seconds, frames
    t = torch.tensor([seconds], dtype=torch.float64, device=device)
    c = torch.tensor([frames], dtype=torch.int64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(c, op=dist.ReduceOp.SUM)
    return float(t[0]), int(c[0])
