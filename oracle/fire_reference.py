"""TEST INFRASTRUCTURE -- never imported by molecular_dynamics_b200/.

FP64 restatement of one batched FIRE iteration (Bitzek et al., PRL 97, 170201), the arithmetic b200md_fire_update
(csrc/b200md_minimize.cuh) performs in fp32: the checker for tests/test_gpu_parity.py::test_fire_update_kernel_vs_restatement.
Upstream TorchMD has no batched minimiser (torchmd.minimizers.minimize_bfgs relaxes one replica with scipy), so this
path's parity is "unpinned": it is checked against this restatement and by its properties (energies fall, max|F| falls).
"""
import numpy as np


def fire_update(pos, vel, frc, dt, alpha, npos, dt_max=0.2, max_move=0.1, n_min=5, f_inc=1.1, f_dec=0.5, alpha_start=0.1,
                f_alpha=0.99):
    """pos, vel, frc (R, N, 3) float64; dt, alpha, npos (R,) float64 -> updated copies of all six."""
    R = pos.shape[0]
    f, v = frc.reshape(R, -1), vel.reshape(R, -1).copy()
    power = (f * v).sum(1)
    vnorm, fnorm = np.sqrt((v * v).sum(1)), np.maximum(np.sqrt((f * f).sum(1)), 1e-30)
    up = power > 0
    v = np.where(up[:, None], (1.0 - alpha)[:, None] * v + (alpha * vnorm / fnorm)[:, None] * f, 0.0)
    npos = np.where(up, npos + 1.0, 0.0)
    grow = up & (npos > n_min)
    dt = np.where(grow, np.minimum(dt * f_inc, dt_max), dt)
    alpha = np.where(grow, alpha * f_alpha, alpha)
    dt = np.where(up, dt, dt * f_dec)
    alpha = np.where(up, alpha, alpha_start)
    v = v + dt[:, None] * f
    step = dt[:, None] * v
    smax = np.abs(step).max(1)
    step = step * (max_move / np.maximum(smax, max_move))[:, None]
    return pos + step.reshape(pos.shape), v.reshape(pos.shape), dt, alpha, npos
