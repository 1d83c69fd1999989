"""Energy minimisation around the CUDA force path.

``minimize_bfgs(system, forces, fmax=0.5, steps=1000)`` keeps upstream's call (tutorial.ipynb cell 14,
``torchmd.minimizers.minimize_bfgs``: scipy L-BFGS-B over the coordinates of ONE replica, gradients from
``forces.compute``).  ``minimize_fire(system, forces, ...)`` is the batched form for thousands of replicas: FIRE
(Bitzek et al., PRL 97, 170201) with per-replica time step and mixing, one ``Forces.compute_batch`` launch per
iteration, state and arithmetic on the GPU — the tool to relax perturbed conformers before a batched MD run.
Host plumbing (torch ops), not a kernel: the force evaluation is the hot part and is the CUDA path.
"""
from __future__ import annotations

import numpy as np
import torch


def minimize_bfgs(system, forces, fmax=0.5, steps=1000):
    from scipy.optimize import minimize

    if steps == 0:
        return
    if system.pos.shape[0] != 1:
        raise RuntimeError("System minimization currently doesn't support replicas")  # upstream's message
    dev, dt = system.pos.device, system.pos.dtype
    frc = torch.zeros_like(system.forces)

    def evalfunc(coords):
        pos = torch.tensor(coords.reshape(1, -1, 3), dtype=dt, device=dev)
        epot = forces.compute_batch(pos, system.box, frc)[0].sum().item()
        return epot, -frc[0].detach().cpu().numpy().astype(np.float64).reshape(-1)

    x0 = system.pos.detach().cpu().numpy()[0].astype(np.float64).reshape(-1)
    res = minimize(evalfunc, x0, method="L-BFGS-B", jac=True, options={"gtol": fmax, "maxiter": steps})
    system.pos = torch.tensor(res.x.reshape(1, -1, 3), dtype=dt, device=dev)
    forces.compute_batch(system.pos, system.box, system.forces)
    return res


def minimize_fire(system, forces, steps=1000, fmax=0.5, dt_start=0.02, dt_max=0.2, max_move=0.1, n_min=5, f_inc=1.1,
                  f_dec=0.5, alpha_start=0.1, f_alpha=0.99):
    """Batched FIRE on every replica of ``system`` (in place).  Stops when max |F| <= fmax (kcal/mol/A) on all replicas
    or after ``steps`` force evaluations.  Returns dict(epot (R,), fmax (R,), iterations)."""
    pos = system.pos
    R = pos.shape[0]
    dev = pos.device
    frc = system.forces
    vel = torch.zeros_like(pos)
    dt = torch.full((R, 1, 1), dt_start, device=dev, dtype=pos.dtype)
    alpha = torch.full((R, 1, 1), alpha_start, device=dev, dtype=pos.dtype)
    npos = torch.zeros((R, 1, 1), device=dev, dtype=pos.dtype)
    ene = forces.compute_batch(pos, system.box, frc)
    it = 0
    for it in range(1, steps + 1):
        fnorm = frc.flatten(1).norm(dim=1).clamp_min(1e-30).view(R, 1, 1)
        if float(frc.abs().amax()) <= fmax:
            break
        power = (frc * vel).flatten(1).sum(1).view(R, 1, 1)
        vnorm = vel.flatten(1).norm(dim=1).view(R, 1, 1)
        vel = (1.0 - alpha) * vel + alpha * frc * (vnorm / fnorm)
        up = power > 0
        npos = torch.where(up, npos + 1, torch.zeros_like(npos))
        grow = up & (npos > n_min)
        dt = torch.where(grow, (dt * f_inc).clamp_max(dt_max), dt)
        alpha = torch.where(grow, alpha * f_alpha, alpha)
        dt = torch.where(up, dt, dt * f_dec)
        alpha = torch.where(up, alpha, torch.full_like(alpha, alpha_start))
        vel = torch.where(up, vel, torch.zeros_like(vel))
        vel = vel + dt * frc
        step = dt * vel
        smax = step.flatten(1).abs().amax(1).view(R, 1, 1)
        step = step * (max_move / smax.clamp_min(max_move))  # no atom moves more than max_move per iteration
        pos += step
        ene = forces.compute_batch(pos, system.box, frc)
    return {"epot": ene.sum(1), "fmax": frc.flatten(1).abs().amax(1), "iterations": it}
