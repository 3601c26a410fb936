"""File/chunk sharding across the GPUs of one box: the path partitions by independent
files (or, inside one file in slice mode, by slices) and has NO exchange step, so every rank works on
its own shard and only result METADATA (sizes, times) is gathered.  `torch.distributed` is plumbing:
NCCL on the GPU box, gloo in the CPU tests.

The assignment is a pure function of (weights, world_size): every rank computes the same table
without communicating."""
from typing import List, Sequence


def assign(weights: Sequence[int], world_size: int) -> List[List[int]]:
    """Longest-processing-time-first partition: item indices per rank, deterministic, disjoint, covering.
    Heaviest items go first to the currently lightest rank (ties -> lowest rank), which bounds the
    makespan at 4/3 of optimal -- a file is the unit in reference-parity mode (one serial chunk), a
    slice in chunk mode."""
    if world_size < 1:
        raise ValueError("world_size must be >= 1")
    order = sorted(range(len(weights)), key=lambda i: (-int(weights[i]), i))
    load = [0] * world_size
    out: List[List[int]] = [[] for _ in range(world_size)]
    for i in order:
        r = min(range(world_size), key=lambda k: (load[k], k))
        out[r].append(i)
        load[r] += int(weights[i])
    for lst in out:
        lst.sort()
    return out


def my_items(weights: Sequence[int], rank: int, world_size: int) -> List[int]:
    return assign(weights, world_size)[rank]


def gather_sizes(local: dict, group=None) -> dict:
    """All ranks contribute {item index: value}; every rank gets the merged table (metadata only -- the
    coded bytes themselves never cross ranks)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return dict(local)
    parts = [None] * dist.get_world_size(group)
    dist.all_gather_object(parts, dict(local), group=group)
    merged = {}
    for p in parts:
        for k, v in p.items():
            if k in merged:
                raise RuntimeError("item %r was processed by two ranks" % (k,))
            merged[k] = v
    return merged


def max_over_ranks(x: float, device=None, group=None) -> float:
    """Timing rule of the bench: a multi-GPU time is the MAX over ranks."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return float(x)
    t = torch.tensor([float(x)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())


def sum_over_ranks(x: float, device=None, group=None) -> float:
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return float(x)
    t = torch.tensor([float(x)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return float(t.item())
