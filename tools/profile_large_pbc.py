#!/usr/bin/env python
"""Force-only Forces.compute of the 256 x 1 040-atom bundle (tools/profile_large.py) in cutoff mode: vacuum against a
500 A periodic box (same geometry, same pairs within the cutoff): the periodic tile-box test must cost nothing extra."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from molecular_dynamics_b200 import Forces, System  # noqa: E402
from molecular_dynamics_b200.parameters import Parameters  # noqa: E402

dev = torch.device("cuda:0")
with np.load(os.path.join(ROOT, "tests", "golden", "ala10.npz")) as z:
    g = {k: z[k] for k in z.files}
copies, R = 10, 256
par = Parameters.from_arrays({k[4:]: v for k, v in g.items() if k.startswith("par_")}, device=dev).tiled(copies)
N = 104 * copies
base = torch.tensor(g["base_frames"][np.arange(R) % len(g["base_frames"])], dtype=torch.float32, device=dev)
shift = torch.arange(copies, device=dev, dtype=torch.float32).repeat_interleave(104)[None, :, None] * torch.tensor([40.0, 0.0, 0.0], device=dev)
s = System(N, R, torch.float, dev)
s.pos[:] = base.repeat(1, copies, 1) + shift
F = Forces(par, cutoff=9.0, rfa=True, switch_dist=7.5)
F.check_nan = False
for label, L in (("vacuum", 0.0), ("periodic 500 A", 500.0)):
    s.box.zero_()
    for c in range(3):
        s.box[:, c, c] = L
    for _ in range(3):
        F.compute_batch(s.pos, s.box, s.forces, energies=False)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        F.compute_batch(s.pos, s.box, s.forces, energies=False)
    b.record()
    torch.cuda.synchronize()
    print(f"{label}: {a.elapsed_time(b) / 20 * 1e3:.1f} us per force-only Forces.compute, checksum {float(s.forces.double().abs().sum()):.6f}")
