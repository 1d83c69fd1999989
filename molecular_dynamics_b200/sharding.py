"""Replica sharding across ranks: replicas never interact (TorchMD_Fix/forces.py:104 loops
over them), so rank g owns a contiguous block and the only collective is the all-gather of per-replica
observables at reporting intervals.  Works with any torch.distributed backend (nccl on GPUs, gloo in tests)."""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_range(total: int, rank: int, world: int):
    """Contiguous block [lo, hi) of `total` replicas owned by `rank`; sizes differ by at most one."""
    base, extra = divmod(total, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def gather_observables(local: torch.Tensor, total: int, group=None) -> torch.Tensor:
    """All-gather a rank's (R_local, K) observable block into the (total, K) array in global replica order."""
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return local
    world = dist.get_world_size(group)
    sizes = [shard_range(total, r, world) for r in range(world)]
    nmax = max(hi - lo for lo, hi in sizes)
    pad = torch.zeros((nmax,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    out = torch.empty((world * nmax,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, pad, group=group)
    return torch.cat([out[r * nmax: r * nmax + (hi - lo)] for r, (lo, hi) in enumerate(sizes)], dim=0)
