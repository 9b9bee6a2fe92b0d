"""Batch sharding of independent depth frames across ranks (one process per GPU, no collective on the hot path).

Every frame's attention is independent (``B`` is a pure batch dimension in every op of the reference block,
``models/ITA/QAT/layers.py:101-127``), so rank ``r`` of ``W`` simply owns a contiguous chunk of the batch; weights and
requant parameters (< 200 KB) are replicated.  The only communication is the optional gather of results / the
max-over-ranks reduction of timings, both off the timed path.
"""
from __future__ import annotations

from typing import List, Tuple

import torch
import torch.distributed as dist


def shard_range(total: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous [begin, end) of ``total`` frames owned by ``rank``; sizes differ by at most one, ragged tails and
    ``total < world`` (empty shards) included."""
    if world <= 0 or not (0 <= rank < world) or total < 0:
        raise ValueError(f"bad shard request total={total} rank={rank} world={world}")
    base, rem = divmod(total, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def all_shards(total: int, world: int) -> List[Tuple[int, int]]:
    return [shard_range(total, r, world) for r in range(world)]


def gather_frames(local: torch.Tensor, total: int, group=None) -> torch.Tensor:
    """Concatenate every rank's ``[n_r, S, E]`` result in rank order (works for ragged shards; any backend)."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return local
    world = dist.get_world_size(group)
    sizes = [e - b for b, e in all_shards(total, world)]
    mx = max(sizes)
    pad = torch.zeros((mx,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    bufs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(bufs, pad, group=group)
    return torch.cat([b[:n] for b, n in zip(bufs, sizes)], dim=0)


def max_over_ranks(value: float, device=None, group=None) -> float:
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())
