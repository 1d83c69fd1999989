#!/usr/bin/env python
"""Single-shot Forces.compute over 65 536 ALA10 conformers (BASELINE.json configs[1]); used under ncu."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from molecular_dynamics_b200 import Forces, Parameters, System  # noqa: E402

g = bench.load_ala10()
dev = torch.device("cuda", 0)
par = Parameters.from_arrays({k[4:]: v for k, v in g.items() if k.startswith("par_")}, device=dev)
R = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
pos, _ = bench.synth_inputs(g, 4096)
s = System(104, R, torch.float, dev)
s.pos[:] = torch.tensor(pos, device=dev).repeat(R // 4096, 1, 1)
F = Forces(par)
F.check_nan = False
for _ in range(4):
    F.compute_batch(s.pos, s.box, s.forces)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(5):
    F.compute_batch(s.pos, s.box, s.forces)
b.record()
torch.cuda.synchronize()
print(f"{R / (a.elapsed_time(b) * 1e-3 / 5):.4e} replica-evaluations/s")
