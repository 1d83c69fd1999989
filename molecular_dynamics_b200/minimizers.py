"""Energy minimisation around the CUDA force path.

``minimize_bfgs(system, forces, fmax=0.5, steps=1000)`` keeps upstream's call (tutorial.ipynb cell 14,
``torchmd.minimizers.minimize_bfgs``: scipy L-BFGS-B over the coordinates of ONE replica, gradients from
``forces.compute``).  ``minimize_fire(system, forces, ...)`` is the batched form for thousands of replicas: FIRE
(Bitzek et al., PRL 97, 170201) with per-replica time step and mixing, one ``Forces.compute_batch`` launch per
iteration and one fused update kernel (``b200md_fire_update``, csrc/b200md_minimize.cuh) -- the tool to relax
perturbed conformers before a batched MD run.
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
                  f_dec=0.5, alpha_start=0.1, f_alpha=0.99, check_every=10):
    """Batched FIRE on every replica of ``system`` (in place).  Stops when max |F| <= fmax (kcal/mol/A) on all replicas
    or after ``steps`` force evaluations.  Returns dict(epot (R,), fmax (R,), iterations).

    Two launches per iteration and nothing else: ``b200md_fire_update`` (one kernel: the per-replica reductions F.v,
    |v|, |F|, max|F|, the time-step / mixing logic, velocity and position update with the displacement cap) and a
    force-only ``Forces.compute_batch``.  The convergence flag (per-replica max|F|, left on the device by the update
    kernel) is read every ``check_every`` iterations instead of every one -- the only host synchronisation of the loop;
    energies are evaluated once, at the end."""
    import ctypes as C

    from . import _cabi

    pos, frc = system.pos, system.forces
    if not pos.is_cuda or pos.dtype != torch.float32 or not pos.is_contiguous() or not frc.is_contiguous():
        raise RuntimeError("minimize_fire needs contiguous float32 CUDA tensors in System.pos / System.forces")
    R, N, _ = pos.shape
    dev = pos.device
    lib = _cabi.load_library()
    vel = torch.zeros_like(pos)
    state = torch.zeros((R, 4), dtype=torch.float32, device=dev)
    state[:, 0], state[:, 1] = dt_start, alpha_start
    opts = _cabi.FireOpts(dt_max, max_move, f_inc, f_dec, alpha_start, f_alpha, int(n_min))
    forces.compute_batch(pos, system.box, frc, energies=False)
    it = 0
    index = dev.index if dev.index is not None else torch.cuda.current_device()
    while it < steps:
        rc = lib.b200md_fire_update(index, R, N, C.c_void_p(pos.data_ptr()), C.c_void_p(vel.data_ptr()), C.c_void_p(frc.data_ptr()),
                                    C.c_void_p(state.data_ptr()), C.byref(opts), C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
        if rc != _cabi.OK:
            raise RuntimeError(f"b200md error {rc}: {_cabi.last_error()}")
        it += 1
        if it % check_every == 0 and float(state[:, 3].max()) <= fmax:
            # state[:, 3] is max|F| of the forces the update just consumed: converged one iteration ago -- the move it made
            # from there is downhill by construction, keep it
            forces.compute_batch(pos, system.box, frc, energies=False)
            break
        forces.compute_batch(pos, system.box, frc, energies=False)
    ene = forces.compute_batch(pos, system.box, frc)
    return {"epot": ene.sum(1), "fmax": frc.flatten(1).abs().amax(1), "iterations": it}
