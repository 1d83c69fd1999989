#!/usr/bin/env python
"""Achieved parity of the CUDA path vs the fp64 outputs of the reference forces.py (tests/golden) and
vs the FP64 oracle on a 4096-conformer ALA10 batch.  Run on a GPU box; writes a markdown table."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import oracle  # noqa: E402  (checker)
from conftest import MODES, SYSTEMS, load_golden  # noqa: E402
from molecular_dynamics_b200 import Forces, System, _cabi  # noqa: E402
from molecular_dynamics_b200.parameters import Parameters  # noqa: E402

DEV = "cuda:0"


def gpu(g, pos, **kw):
    par = Parameters.from_arrays(g.par.to_arrays(), device=DEV)
    F = Forces(par, **kw)
    p = torch.tensor(np.asarray(pos, dtype=np.float32), device=DEV)
    s = System(p.shape[1], p.shape[0], torch.float, DEV)
    e = F.compute_batch(p, s.box, s.forces)
    return s.forces.cpu().numpy().astype(np.float64), e.cpu().numpy()


def row(name, mode, f, e, fr, er):
    ferr = (np.abs(f - fr).reshape(len(f), -1).max(1) / np.abs(fr).reshape(len(f), -1).max(1)).max()
    cells = []
    for k in (0, 1, 2, 3, 4, 6):
        ref = np.abs(er[:, k])
        if ref.max() == 0:
            cells.append("-")
            continue
        rel = (np.abs(e[:, k] - er[:, k]) / np.maximum(ref, 1e-30))
        cells.append(f"{rel.max():.1e}")
    tot = np.abs(e.sum(1) - er.sum(1)) / np.abs(er.sum(1))
    return f"| {name} | {mode} | {len(f)} | {ferr:.1e} | " + " | ".join(cells) + f" | {tot.max():.1e} |"


def main():
    out = ["| system | mode | replicas | max\\|dF\\|/max\\|F\\| | electrostatics | lj | bonds | angles | dihedrals | impropers | total |",
           "|---|---|---|---|---|---|---|---|---|---|---|"]
    for name in SYSTEMS:
        g = load_golden(name)
        for mode, kw in MODES.items():
            f, e = gpu(g, g["pos"], **kw)
            out.append(row(name, mode, f, e, g[f"ref_{mode}_f64_forces"], g[f"ref_{mode}_f64_energies"]))
    g = load_golden("ala10")
    rng = np.random.default_rng(1234)
    base = g["base_frames"]
    R = 4096
    pos = (base[np.arange(R) % len(base)] + 0.05 * rng.standard_normal((R, 104, 3))).astype(np.float32)
    for mode, kw in MODES.items():
        f, e = gpu(g, pos, **kw)
        fr, er = oracle.forces(g.packed, pos, **kw)
        out.append(row("ala10 perturbed (vs C oracle)", mode, f, e, fr, er))
    # solutes above 128 atoms (csrc/b200md_large.cuh): 312 atoms against the reference itself, the 1 040-atom benchmark
    # bundle (ten chains 40 A apart) against the oracle
    gl = load_golden("ala10x3")
    for mode, kw in MODES.items():
        f, e = gpu(gl, gl["pos"], **kw)
        out.append(row("ala10x3, 312 atoms (large-solute kernels)", mode, f, e, gl[f"ref_{mode}_f64_forces"], gl[f"ref_{mode}_f64_energies"]))
    copies, R = 10, 8
    parL = g.par.tiled(copies)
    shift = np.repeat(np.arange(copies), 104)[None, :, None] * np.array([40.0, 0.0, 0.0])
    posL = (base[:R * copies].reshape(R, copies * 104, 3) + shift).astype(np.float32)
    packedL = _cabi.PackedTopology(parL, _cabi.make_pair_index(104 * copies, parL.get_exclusions()))

    class _G:
        par = parL

    for mode, kw in MODES.items():
        f, e = gpu(_G, posL, **kw)
        fr, er = oracle.forces(packedL, posL, **kw)
        out.append(row("10 x ala10, 1 040 atoms (vs C oracle)", mode, f, e, fr, er))
    print("\n".join(out))


if __name__ == "__main__":
    main()
