"""Multi-GPU partition of the hot path: one process per GPU, studies sharded over ranks, no data-path collective.

The reference processes its inputs one after the other in one process (reference ts2d/main.py:80-83).  Studies are
independent, so rank r of W takes studies r, r+W, r+2W, ... (all five group models are resident on every GPU); the only
communication is the bookkeeping below (a max of device times, a sum of counts, and an ordered gather of small
per-study records), which works over NCCL on the GPU box and over gloo in the CPU tests.
"""
from __future__ import annotations

import os
from typing import Any, List, Sequence, Tuple


def rank_world() -> Tuple[int, int, int]:
    """(rank, world size, local rank) from the torchrun environment; (0, 1, 0) when launched plainly"""
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def shard_indices(n_items: int, rank: int, world: int) -> List[int]:
    """indices of the studies rank `rank` processes: i with i % world == rank (round robin keeps ranks balanced to +-1)"""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"invalid rank {rank} of world size {world}")
    return list(range(rank, n_items, world))


def shard(items: Sequence[Any], rank: int, world: int) -> List[Any]:
    return [items[i] for i in shard_indices(len(items), rank, world)]


def _dist():
    import torch.distributed as dist
    return dist if dist.is_available() and dist.is_initialized() else None


def _device():
    import torch
    d = _dist()
    if d is not None and d.get_backend() == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def reduce_max(x: float) -> float:
    """max over ranks (device time of a step = the slowest rank)"""
    import torch
    d = _dist()
    if d is None:
        return float(x)
    t = torch.tensor([float(x)], dtype=torch.float64, device=_device())
    d.all_reduce(t, op=d.ReduceOp.MAX)
    return float(t.item())


def reduce_sum(x: float) -> float:
    import torch
    d = _dist()
    if d is None:
        return float(x)
    t = torch.tensor([float(x)], dtype=torch.float64, device=_device())
    d.all_reduce(t, op=d.ReduceOp.SUM)
    return float(t.item())


def gather_in_order(local_records: Sequence[Any], n_items: int) -> List[Any]:
    """inverse of `shard`: every rank contributes the records of its studies, every rank gets the list in study order"""
    d = _dist()
    if d is None:
        return list(local_records)
    world, rank = d.get_world_size(), d.get_rank()
    if len(local_records) != len(shard_indices(n_items, rank, world)):
        raise ValueError("rank {} holds {} records for {} studies".format(rank, len(local_records), len(shard_indices(n_items, rank, world))))
    parts: List[Any] = [None] * world
    d.all_gather_object(parts, list(local_records))
    out: List[Any] = [None] * n_items
    for r in range(world):
        for k, i in enumerate(shard_indices(n_items, r, world)):
            out[i] = parts[r][k]
    return out
