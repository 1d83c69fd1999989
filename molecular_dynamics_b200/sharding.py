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


def jarzynski_free_energy(work: torch.Tensor, temperature: float, group=None):
    """Free-energy profile along the pulling coordinate from the SMD work of the whole replica ensemble
    (the reference's SMD data set is the input of exactly this estimate):
    ``dF = -kT ln < exp(-W / kT) >`` over ALL replicas of all ranks, plus the second-cumulant estimate
    ``<W> - var(W) / (2 kT)``.  ``work``: this rank's (..., R_local) block, e.g.
    ``Integrator.observables[:, :, OBS_SMD_WORK]`` -> one value per report.  The exchange is three all-reduces of
    (...)-sized tensors (max, then the shifted exponential sum with the count and the first two moments)."""
    from .integrator import BOLTZMAN

    kT = BOLTZMAN * float(temperature)
    w = work.to(torch.float64)
    x = -w / kT
    multi = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
    m = x.amax(dim=-1) if x.shape[-1] else torch.full(x.shape[:-1], -float("inf"), dtype=x.dtype, device=x.device)
    if multi:
        dist.all_reduce(m, op=dist.ReduceOp.MAX, group=group)
    n = torch.full_like(m, float(x.shape[-1]))
    acc = torch.stack([torch.exp(x - m.unsqueeze(-1)).sum(-1), n, w.sum(-1), (w * w).sum(-1)])
    if multi:
        dist.all_reduce(acc, op=dist.ReduceOp.SUM, group=group)
    se, n, s1, s2 = acc
    mean = s1 / n
    var = (s2 / n - mean * mean).clamp_min(0.0) * n / (n - 1).clamp_min(1.0)
    return {"dF": -kT * (m + torch.log(se / n)), "dF_cumulant2": mean - var / (2.0 * kT), "mean_work": mean, "n": n}
