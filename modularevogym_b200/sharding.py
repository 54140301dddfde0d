"""Multi-GPU plumbing of the batched engine: environments are independent (one reference ``Sim`` each,
no shared state), so they shard across ranks with NO collective on the step path.  The only exchange is
one all-gather of per-environment results (reward, done flag, observation rows) at rollout end
(SURVEY 8e).  ``torch.distributed`` only: NCCL on the GPU box, gloo in the CPU tests."""
from __future__ import annotations

from typing import Tuple

import torch
import torch.distributed as dist


def shard_range(n_total: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block [lo, hi) of global environment ids owned by ``rank``; sizes differ by at most one."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    base, extra = divmod(n_total, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def owner_of(env_id: int, n_total: int, world: int) -> int:
    """Rank that owns global environment ``env_id`` under :func:`shard_range`."""
    base, extra = divmod(n_total, world)
    split = extra * (base + 1)
    return env_id // (base + 1) if env_id < split else extra + (env_id - split) // max(base, 1)


def gather_env_rows(local: torch.Tensor, n_total: int, group=None) -> torch.Tensor:
    """All-gather of per-environment rows at rollout end: ``local`` is ``(n_local, ...)`` on this rank, the
    result is ``(n_total, ...)`` in global environment order on every rank.  Shards may differ by one row, so
    rows are padded to the largest shard for the collective and trimmed afterwards."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return local
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    n_max = -(-n_total // world)
    lo, hi = shard_range(n_total, rank, world)
    if local.shape[0] != hi - lo:
        raise ValueError(f"rank {rank} owns {hi - lo} environments but passed {local.shape[0]} rows")
    padded = local
    if local.shape[0] < n_max:
        pad = torch.zeros((n_max - local.shape[0],) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        padded = torch.cat([local, pad], dim=0)
    out = torch.empty((world * n_max,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, padded.contiguous(), group=group)
    parts = []
    for r in range(world):
        a, b = shard_range(n_total, r, world)
        parts.append(out[r * n_max: r * n_max + (b - a)])
    return torch.cat(parts, dim=0)
