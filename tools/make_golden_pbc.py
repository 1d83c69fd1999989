#!/usr/bin/env python
"""Golden outputs of the reference's minimum-image path (wrap_dist, TorchMD_Fix/forces.py:358-363,
applied inside calculate_distances to every term) from the UNMODIFIED reference forces.py: the BB40 and ALA10 fixture
conformers placed in orthorhombic boxes smaller than the molecule, so images do interact.  Writes
tests/golden/pbc.npz.  Runs only where /root/reference exists.
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
from molecular_dynamics_b200 import _cabi  # noqa: E402
from molecular_dynamics_b200.parameters import Parameters  # noqa: E402

MODES = {"plain": {}, "rf": {"cutoff": 9.0, "rfa": True, "switch_dist": 7.5}}
BOXES = {"bb40": [[26.0, 24.0, 22.0], [30.0, 20.0, 25.0], [0.0, 0.0, 0.0], [21.0, 23.0, 27.5]],
         "ala10": [[28.0, 25.0, 23.0], [22.0, 31.0, 24.5], [0.0, 0.0, 0.0], [25.0, 25.0, 25.0]]}


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
        box = np.zeros((4, 3, 3), dtype=np.float32)
        for r, b in enumerate(BOXES[system]):
            box[r, 0, 0], box[r, 1, 1], box[r, 2, 2] = b
        out[f"{system}_pos"], out[f"{system}_box"] = pos32, box
        for dtype, dn in ((torch.float, "f32"), (torch.double, "f64")):
            par = Parameters.from_arrays(arrs, precision=dtype)
            for mode, kw in MODES.items():
                pos = torch.tensor(pos32).type(dtype)
                frc = torch.zeros_like(pos)
                F = rf.Forces(par, **kw)
                E = F.compute(pos, torch.tensor(box).type(dtype), frc, returnDetails=True)
                out[f"{system}_{mode}_{dn}_forces"] = frc.numpy().copy()
                out[f"{system}_{mode}_{dn}_energies"] = np.array([[float(e[k]) for k in _cabi.ENERGY_SLOTS] for e in E])
    path = os.path.join(ROOT, "tests", "golden", "pbc.npz")
    np.savez_compressed(path, **out)
    print(f"wrote {path}: {os.path.getsize(path) / 1024:.1f} KiB")


if __name__ == "__main__":
    main()
