#!/usr/bin/env python
"""Golden outputs of the reference's coarse-grain pair terms (evaluate_repulsion,
evaluate_repulsion_CG, TorchMD_Fix/forces.py:416-448) from the UNMODIFIED reference forces.py, on the BB40 and ALA10
fixture systems.  Runs only where /root/reference exists; writes tests/golden/repulsion.npz.

    python tools/make_golden_repulsion.py [--ref /root/reference]
"""
import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
from make_golden import load_by_path  # noqa: E402
from molecular_dynamics_b200.parameters import Parameters  # noqa: E402

CASES = {  # name -> (terms, Forces kwargs)
    "rep_plain": (("repulsion",), {}),
    "rep_cut": (("repulsion", "bonds"), {"cutoff": 9.0, "switch_dist": 7.5}),
    "cg_plain": (("repulsioncg", "electrostatics", "bonds"), {}),
    "cg_rf": (("repulsioncg", "electrostatics", "angles"), {"cutoff": 9.0, "rfa": True, "switch_dist": 7.5}),
    # several pair terms of the lj family at once (forces.py:268-316 evaluates each enabled one; only "lj" is switched)
    "combo_plain": (("lj", "repulsion", "electrostatics", "bonds"), {}),
    "combo_cut": (("repulsioncg", "lj", "repulsion", "angles"), {"cutoff": 9.0, "switch_dist": 7.5}),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", default="/root/reference")
    a = ap.parse_args()
    rf = load_by_path("ref_forces", f"{a.ref}/TorchMD_Fix/forces.py")
    out = {}
    for system, key in (("bb40", "frames"), ("ala10", "pos")):
        with np.load(os.path.join(ROOT, "tests", "golden", f"{system}.npz")) as z:
            arrs = {k[4:]: z[k] for k in z.files if k.startswith("par_")}
            pos32 = z[key][:4].astype(np.float32)
        out[f"{system}_pos"] = pos32
        for dtype, dn in ((torch.float, "f32"), (torch.double, "f64")):
            par = Parameters.from_arrays(arrs, precision=dtype)
            for name, (terms, kw) in CASES.items():
                pos = torch.tensor(pos32).type(dtype)
                R, N, _ = pos.shape
                frc = torch.zeros(R, N, 3, dtype=dtype)
                F = rf.Forces(par, terms=terms, **kw)
                E = F.compute(pos, torch.zeros(R, 3, 3, dtype=dtype), frc, returnDetails=True)
                out[f"{system}_{name}_{dn}_forces"] = frc.numpy().copy()
                for t in terms:
                    out[f"{system}_{name}_{dn}_E_{t}"] = np.array([float(e[t]) for e in E], dtype=np.float64)
    path = os.path.join(ROOT, "tests", "golden", "repulsion.npz")
    np.savez_compressed(path, **out)
    print(f"wrote {path}: {os.path.getsize(path) / 1024:.1f} KiB")


if __name__ == "__main__":
    main()
