"""Row sharding for the multi-GPU path: rows are independent, so G ranks
score G contiguous row blocks with a replicated model and never exchange data on the inference
path.  The helpers here are pure host logic shared by the estimator, bench.py and the refit driver;
the only collective in the whole design is the refit gradient all-reduce (`allreduce_sum`).
"""
from __future__ import annotations

from typing import List, Tuple


def shard_bounds(n_rows: int, world: int) -> List[Tuple[int, int]]:
    """Contiguous, disjoint, exhaustive [lo, hi) row blocks; sizes differ by at most one row."""
    if world < 1:
        raise ValueError("world must be >= 1")
    if n_rows < 0:
        raise ValueError("n_rows must be >= 0")
    return [(n_rows * g // world, n_rows * (g + 1) // world) for g in range(world)]


def my_rows(n_rows: int, rank: int, world: int) -> Tuple[int, int]:
    if not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside world of {world}")
    return shard_bounds(n_rows, world)[rank]


def max_over_ranks(value: float, device=None) -> float:
    """Step time of the whole job = slowest rank (bench.py's timing rule)."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def allreduce_sum(tensor):
    """In-place sum across ranks (refit: one fused [loss | grad] buffer per iteration)."""
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(tensor, op=dist.ReduceOp.SUM)
    return tensor
