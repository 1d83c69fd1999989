#!/usr/bin/env python
"""tests/golden/ala10x3.npz: the unmodified reference (TorchMD_Fix/forces.py, imported by path) on a solute above 128
atoms -- three deca-alanine chains 10 A apart as ONE 312-atom topology (Parameters.tiled(3) of the ALA10 fixture) -- in fp64
on fp32-rounded positions, plain and cutoff / reaction field / switch.  Pins the oracle and the large-solute CUDA kernels
(csrc/b200md_large.cuh) against the reference itself.  Runs only where /root/reference exists.

Usage: python tools/make_golden_large.py [--ref /root/reference]
"""
import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

from make_golden import MODES, load_by_path, run_reference  # noqa: E402
from molecular_dynamics_b200.parameters import Parameters  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", default="/root/reference")
    a = ap.parse_args()
    rf = load_by_path("ref_forces", os.path.join(a.ref, "TorchMD_Fix", "forces.py"))
    with np.load(os.path.join(ROOT, "tests", "golden", "ala10.npz")) as z:
        g = {k: z[k] for k in z.files}
    par = Parameters.from_arrays({k[4:]: v for k, v in g.items() if k.startswith("par_")}).tiled(3)
    copies, R = 3, 2
    shift = np.repeat(np.arange(copies), 104)[None, :, None] * np.array([0.0, 10.0, 10.0])
    pos = (g["base_frames"][:R * copies].reshape(R, copies * 104, 3) + shift).astype(np.float32)
    out = {"pos": pos}
    for mode, kw in MODES.items():
        f, e, ava = run_reference(rf, par, pos, torch.float64, **kw)
        out[f"ref_{mode}_f64_forces"], out[f"ref_{mode}_f64_energies"] = f, e
        out["ava_idx"] = ava.astype(np.int32)
    for k, v in par.to_arrays().items():
        out["par_" + k] = v
    path = os.path.join(ROOT, "tests", "golden", "ala10x3.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes;", {k: np.round(out[f"ref_{k}_f64_energies"][0, :5], 4).tolist() for k in MODES})


if __name__ == "__main__":
    main()
