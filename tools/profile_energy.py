#!/usr/bin/env python
"""Cost of the FP64 energy pass by term: Forces.compute over 65 536 ALA10 conformers with and without energies, for term
subsets (device-timed, median of 3 groups of 5 launches).  usage: python tools/profile_energy.py [rf]"""
import statistics
import sys
import os

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from molecular_dynamics_b200 import Forces, Parameters, System  # noqa: E402

dev = torch.device("cuda", 0)
g = bench.load_ala10()
par = Parameters.from_arrays({k[4:]: v for k, v in g.items() if k.startswith("par_")}, device=dev)
R = 65536
pos, _ = bench.synth_inputs(g, 8192)
s = System(104, R, torch.float, dev)
s.pos[:] = torch.tensor(pos, device=dev).repeat(8, 1, 1)
kw = bench.FORCES_MODES["rf" if "rf" in sys.argv[1:] else "plain"]


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    out = []
    for _ in range(3):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        out.append(a.elapsed_time(b) / reps)
    return statistics.median(out)


SETS = {"all": None, "pairs": ("lj", "electrostatics"), "lj": ("lj",), "bonded": ("bonds", "angles", "dihedrals", "1-4", "lj", "electrostatics"),
        "bonds": ("bonds",), "angles": ("angles",), "dihedrals": ("dihedrals",)}
for name, terms in SETS.items():
    F = Forces(par, terms=terms, **kw)
    F.check_nan = False
    te = timed(lambda: F.compute_batch(s.pos, s.box, s.forces))
    tf = timed(lambda: F.compute_batch(s.pos, s.box, s.forces, energies=False))
    print(f"{name:10s} with energies {te:7.3f} ms   force-only {tf:7.3f} ms   energy pass {te - tf:7.3f} ms")
