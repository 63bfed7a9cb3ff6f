"""Batch sharding for multi-GPU inference: the path shards by sample, one process per GPU, no
data-path collective.  Only rank bookkeeping and an optional gather of results live here."""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_range(n: int, rank: int, world: int):
    """Contiguous [lo, hi) slice of n samples owned by `rank` (sizes differ by at most one)."""
    if not (0 <= rank < world):
        raise ValueError("bad rank/world")
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_batch(x: torch.Tensor, rank: int, world: int) -> torch.Tensor:
    lo, hi = shard_range(x.shape[0], rank, world)
    return x[lo:hi]


def gather_batch(y_local: torch.Tensor, n_total: int, group=None) -> torch.Tensor:
    """All ranks get the full [n_total, ...] result (variable shard sizes handled by padding to the largest)."""
    world = dist.get_world_size(group)
    sizes = [shard_range(n_total, r, world) for r in range(world)]
    mx = max(hi - lo for lo, hi in sizes)
    pad = torch.zeros((mx,) + tuple(y_local.shape[1:]), dtype=y_local.dtype, device=y_local.device)
    pad[: y_local.shape[0]] = y_local
    outs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(outs, pad, group=group)
    return torch.cat([o[: hi - lo] for o, (lo, hi) in zip(outs, sizes)], 0)
