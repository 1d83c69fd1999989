#!/usr/bin/env python
"""Measured budget of the fused NVE step (8 192 ALA10 replicas x 500 steps, device-timed) and upper bounds of levers:

  * term subsets -> where the time of a step goes (pairs / bonded kinds), measured, not argued;
  * "uniform types": every atom gets LJ type 0 (timing only, physically wrong) -> all 32 i-lanes of a pass read ONE row of
    the (type, atom j) table, i.e. the best case a type-sorted internal atom order (lever 3a) could reach.

usage: python tools/lever_probe.py [rf]
"""
import os
import statistics
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from molecular_dynamics_b200 import Forces, Integrator, Parameters, System  # noqa: E402

dev = torch.device("cuda", 0)
g = bench.load_ala10()
arrs = {k[4:]: v for k, v in g.items() if k.startswith("par_")}
R, S = 8192, 500
pos_h, vel_h = bench.synth_inputs(g, R)
kw = bench.FORCES_MODES["rf" if "rf" in sys.argv[1:] else "plain"]


def rate(par, terms=None):
    F = Forces(par, terms=terms, **kw)
    F.check_nan = False
    s = System(104, R, torch.float, dev)
    s.pos[:] = torch.tensor(pos_h, device=dev)
    s.vel[:] = torch.tensor(vel_h * (0.0 if terms is not None else 1.0), device=dev)  # subsets: start at rest (no blow-up)
    F.compute_batch(s.pos, s.box, s.forces, energies=False)
    integ = Integrator(s, F, timestep=0.01 if terms is not None else 1.0, device=dev)
    integ.energy_every, integ.check_nan = 0, False
    integ.step(niter=S)
    torch.cuda.synchronize()
    ts = []
    for _ in range(3):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        integ.step(niter=S)
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    t = statistics.median(ts)
    return R * S / (t * 1e-3), t * 1e3 / S  # replica-steps/s, us per step of the whole batch


par = Parameters.from_arrays(arrs, device=dev)
full, us_full = rate(par)
print(f"{'all terms':34s} {full:.4e} replica-steps/s   {us_full:8.2f} us per batch step")
for name, terms in (("pairs (lj + electrostatics)", ("lj", "electrostatics")), ("bonds", ("bonds",)), ("angles", ("angles",)),
                    ("dihedrals + 1-4", ("dihedrals", "1-4", "lj", "electrostatics")), ("dihedrals", ("dihedrals",)),
                    ("bonded (no pair tiles)", ("bonds", "angles", "dihedrals"))):
    r, us = rate(par, terms)
    print(f"{name:34s} {r:.4e} replica-steps/s   {us:8.2f} us per batch step   {us / us_full:5.2f} of a full step")
u = dict(arrs)
u["mapped_atom_types"] = np.zeros_like(arrs["mapped_atom_types"])
ru, _ = rate(Parameters.from_arrays(u, device=dev))
print(f"{'all terms, uniform LJ types':34s} {ru:.4e} replica-steps/s   ({ru / full:.3f} x: upper bound of a type-sorted atom order)")
