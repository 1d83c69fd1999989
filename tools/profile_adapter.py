#!/usr/bin/env python
"""Where a forward + backward of the batched Energy adapter (the ML callers' loss, cgan.py:345-360) spends its time:
device time by CUDA events, wall time, and the torch profiler's view of the host side (synchronisations, copies, op
dispatch).  usage: python tools/profile_adapter.py [batch]"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from molecular_dynamics_b200 import Energy  # noqa: E402
from molecular_dynamics_b200.parameters import Parameters  # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    dev = torch.device("cuda:0")
    with np.load(os.path.join(ROOT, "tests", "golden", "ala10.npz")) as z:
        g = {k: z[k] for k in z.files}
    par = Parameters.from_arrays({k[4:]: v for k, v in g.items() if k.startswith("par_")}, device=dev)
    colvar = {"name": "E2End Harm", "fk": 1.0, "cent_0": 12.0, "cent_1": 34.0, "T": 1.0, "group1": [3], "group2": [98]}
    en = Energy(parameters=par, colvar=colvar, device=dev)
    en.forces.check_nan = False
    x = torch.tensor(np.resize(g["base_frames"], (B, 104, 3)).astype(np.float32), device=dev).requires_grad_(True)
    t = torch.linspace(0.0, 1.0, B, device=dev)

    def step():
        x.grad = None
        ene, phis = en.calc_energy_batch(x, t)
        loss = sum(v for k, v in ene.items() if k != "E2End Harm").sum() / B + 200.0 * ene["E2End Harm"].sum() / B
        loss.backward()

    for _ in range(5):
        step()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    a.record()
    for _ in range(20):
        step()
    b.record()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / 20
    print(f"batch {B}: {a.elapsed_time(b) / 20:.3f} ms per step by events, {1e3 * wall:.3f} ms wall, {B / wall:.4g} conformers/s")
    from torch.profiler import ProfilerActivity, profile

    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        for _ in range(5):
            step()
        torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by="self_cpu_time_total", row_limit=25, max_name_column_width=60))
    print(prof.key_averages().table(sort_by="self_cuda_time_total", row_limit=12, max_name_column_width=60))


if __name__ == "__main__":
    main()
