"""Request sharding across per-GPU replicas (one process per GPU under torchrun).

The path shards by request: every replica holds the whole model, requests are
independent, so there is NO collective on the data path. torch.distributed is used only as
plumbing around a measurement or an offline batch: a barrier, the max over ranks of a timing,
and gathering each rank's results to rank 0.
"""
from __future__ import annotations

from typing import Sequence, TypeVar

T = TypeVar("T")


def shard_bounds(n_items: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous, balanced split: the first n % world ranks get one extra item."""
    base, extra = divmod(n_items, world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def shard_requests(items: Sequence[T], rank: int, world: int) -> Sequence[T]:
    lo, hi = shard_bounds(len(items), rank, world)
    return items[lo:hi]


def max_over_ranks(values: Sequence[float], device=None) -> list[float]:
    """Element-wise MAX of per-rank timings (device-timed numbers are reported as the slowest rank's)."""
    import torch
    import torch.distributed as dist

    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.tolist()


def gather_to_rank0(obj):
    """Rank 0 receives [obj_rank0, obj_rank1, ...]; other ranks receive None."""
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return [obj]
    out = [None] * dist.get_world_size() if dist.get_rank() == 0 else None
    dist.gather_object(obj, out, dst=0)
    return out
