#!/usr/bin/env python
"""Per-launch device times of back-to-back Integrator.step calls (clock / power behaviour under sustained load)."""
import os, sys, subprocess
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from molecular_dynamics_b200 import Forces, Integrator, Parameters, System
R = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
g = bench.load_ala10(); dev = torch.device("cuda", 0)
par = Parameters.from_arrays({k[4:]: v for k, v in g.items() if k.startswith("par_")}, device=dev)
pos, vel = bench.synth_inputs(g, min(R, 8192))
s = System(104, R, torch.float, dev)
rep = -(-R // pos.shape[0])
s.pos[:] = torch.tensor(pos, device=dev).repeat(rep, 1, 1)[:R]; s.vel[:] = torch.tensor(vel, device=dev).repeat(rep, 1, 1)[:R]
F = Forces(par); F.check_nan = False
F.compute_batch(s.pos, s.box, s.forces)
integ = Integrator(s, F, timestep=1.0, device=dev); integ.energy_every = 100; integ.check_nan = False
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
for a, b in ev:
    a.record(); integ.step(niter=1000); b.record()
torch.cuda.synchronize()
t = [a.elapsed_time(b) for a, b in ev]
print(R, " ".join(f"{x:.1f}" for x in t))
print(subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,clocks_event_reasons.active", "--format=csv,noheader"], capture_output=True, text=True).stdout.strip())
