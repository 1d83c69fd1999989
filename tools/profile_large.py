"""Large-solute kernels on a bundle of `copies` deca-alanine chains in one topology: times Forces.compute with CUDA
events (and is the command ncu wraps for profiles/*_large_pairs_kernel.md).
usage: python tools/profile_large.py [--copies 10] [--replicas 256] [--mode plain|rf] [--reps 5]"""
import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__  # noqa: E402

__graft_entry__.build()
from molecular_dynamics_b200 import Forces, Parameters, System  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--copies", type=int, default=10)
    ap.add_argument("--replicas", type=int, default=256)
    ap.add_argument("--mode", default="plain")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--integrate", type=int, default=0, help="> 0: also time Integrator.step(niter) (Langevin 300 K, 1 fs)")
    ap.add_argument("--rigid", action="store_true", help="rigid X-H bonds (SHAKE / RATTLE), 2 fs steps")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    with np.load(os.path.join(ROOT, "tests", "golden", "ala10.npz")) as z:
        g = {k: z[k] for k in z.files}
    par = Parameters.from_arrays({k[4:]: v for k, v in g.items() if k.startswith("par_")}, device=dev).tiled(a.copies)
    N = 104 * a.copies
    base = torch.tensor(g["base_frames"][np.arange(a.replicas) % len(g["base_frames"])], dtype=torch.float32, device=dev)
    shift = torch.arange(a.copies, device=dev, dtype=torch.float32).repeat_interleave(104)[None, :, None] * \
        torch.tensor([40.0, 0.0, 0.0], device=dev)
    s = System(N, a.replicas, torch.float, dev)
    s.pos[:] = base.repeat(1, a.copies, 1) + shift
    kw = {} if a.mode == "plain" else {"cutoff": 9.0, "rfa": True, "switch_dist": 7.5}
    F = Forces(par, **kw)
    F.check_nan = False
    F.compute_batch(s.pos, s.box, s.forces)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.reps):
        F.compute_batch(s.pos, s.box, s.forces)
    e1.record()
    torch.cuda.synchronize()
    t = e0.elapsed_time(e1) * 1e-3 / a.reps
    npair = int(F.ava_idx.shape[0])
    flop = (36 + (8 if kw else 0)) * npair
    print(f"N {N} R {a.replicas} mode {a.mode}: {1e3 * t:.3f} ms per Forces.compute, {a.replicas / t:.4g} evaluations/s, "
          f"{a.replicas * npair / t:.4g} unique pairs/s, {a.replicas * flop / t / 1e12:.2f} TFLOP/s (pair flops, unique pairs)")


    if a.integrate > 0:
        from molecular_dynamics_b200 import Integrator

        s.vel[:] = 0.01 * torch.randn(s.vel.shape, device=dev, generator=torch.Generator(device=dev).manual_seed(1))
        dt_fs = 2.0 if a.rigid else 1.0
        integ = Integrator(s, F, timestep=dt_fs, device=dev, gamma=1.0, T=300.0, seed=3, rigid_bonds=a.rigid)
        integ.energy_every, integ.check_nan = 10, False
        integ.step(niter=40)  # (a synthetic start is not on the constraint surface: the first steps project it)
        integ.step(niter=a.integrate)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(a.reps):
            integ.step(niter=a.integrate)
        e1.record()
        torch.cuda.synchronize()
        t = e0.elapsed_time(e1) * 1e-3 / (a.reps * a.integrate)
        print(f"N {N} R {a.replicas} mode {a.mode}{' rigid bonds 2 fs' if a.rigid else ''}: Langevin step {1e6 * t:.1f} us, "
              f"{a.replicas / t:.4g} replica-steps/s, {a.replicas / t * dt_fs * 1e-6 * 86400:.4g} aggregate ns/day")


if __name__ == "__main__":
    main()
