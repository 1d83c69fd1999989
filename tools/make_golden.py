#!/usr/bin/env python
"""Generate tests/golden/*.npz from the reference checkout (runs only where /root/reference exists).

The reference's own TorchMD_Fix/forces.py is imported *unmodified, by path* and run in fp32 and fp64 on
fp32-rounded inputs; its outputs, the parsed parameter tables and the input conformers are stored as
small .npz fixtures that travel to the GPU box (where /root/reference does not exist).

Also stored: the notebook known-answer vector (V_Calculations/Test-1/tutorial.ipynb:231-271, parsed
from the stored cell output), the colvar known answers (smd_out.colvars.traj last lines) and short
trajectories of the *restated upstream* velocity-Verlet loop driving the reference forces.py in fp64
("upstream, unpinned": the reference holds no integrator output).

Usage: python tools/make_golden.py [--ref /root/reference]
"""
import argparse
import importlib.util
import json
import os
import re
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from molecular_dynamics_b200 import _cabi  # noqa: E402
from molecular_dynamics_b200.parameters import Parameters, parameters_from_prmtop, parameters_from_psf_yaml  # noqa: E402
from molecular_dynamics_b200.topology_io import NAMD_VEL_TO_ANGSTROM_PER_PS, read_namd_bin, read_pdb_coords  # noqa: E402

TIMEFACTOR = 48.88821
BOLTZMAN = 0.001987191
MODES = {"plain": dict(), "rf": dict(cutoff=9, rfa=True, switch_dist=7.5)}


def load_by_path(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def run_reference(rf, par, pos32, dtype, external=None, **kw):
    """Reference Forces.compute on (R,N,3) fp32-rounded positions in `dtype`."""
    p2 = Parameters.from_arrays(par.to_arrays(), precision=dtype)
    pos = torch.tensor(np.asarray(pos32, dtype=np.float32)).type(dtype)
    R, N, _ = pos.shape
    frc = torch.zeros(R, N, 3, dtype=dtype)
    box = torch.zeros(R, 3, 3, dtype=dtype)
    F = rf.Forces(p2, external=external, **kw)
    E = F.compute(pos, box, frc, returnDetails=True)
    ene = np.array([[float(e[k]) for k in _cabi.ENERGY_SLOTS] for e in E], dtype=np.float64)
    return frc.numpy().copy(), ene, F.ava_idx.numpy().copy()


class SmdExternal:
    """Forces(external=...) object (forces.py:318-323) carrying the SMD bias of potcalc.py:75-89;
    the force is -dE/dr of that energy (the reference obtains it by autograd)."""

    def __init__(self, a1, a2, fk, center):
        self.a1, self.a2, self.fk, self.center = a1, a2, fk, center

    def calculate(self, pos, box):
        d = pos[:, self.a2] - pos[:, self.a1]
        dist = torch.sqrt((d * d).sum(1))
        dev = self.center - dist
        ene = 0.5 * self.fk * dev * dev
        frc = torch.zeros_like(pos)
        g = (self.fk * dev / dist)[:, None] * d
        frc[:, self.a2] += g
        frc[:, self.a1] -= g
        return ene, frc


def reference_outputs(rf, par, pos32, out, tag=""):
    for mode, kw in MODES.items():
        f32, e32, ava = run_reference(rf, par, pos32, torch.float, **kw)
        f64, e64, _ = run_reference(rf, par, pos32, torch.double, **kw)
        out[f"{tag}ref_{mode}_f32_forces"], out[f"{tag}ref_{mode}_f32_energies"] = f32, e32
        out[f"{tag}ref_{mode}_f64_forces"], out[f"{tag}ref_{mode}_f64_energies"] = f64, e64
    out["ava_idx"] = ava


def upstream_vv(rf, par, pos32, vel32, nsteps, dt_fs, **kw):
    """Restated upstream Integrator.step in fp64, NVE, driving reference forces.py."""
    p2 = Parameters.from_arrays(par.to_arrays(), precision=torch.double)
    pos = torch.tensor(np.asarray(pos32, dtype=np.float32)).double()
    vel = torch.tensor(np.asarray(vel32, dtype=np.float32)).double()
    R, N, _ = pos.shape
    frc = torch.zeros_like(pos)
    box = torch.zeros(R, 3, 3, dtype=torch.double)
    F = rf.Forces(p2, **kw)
    m = p2.masses  # (N,1)
    dt = dt_fs / TIMEFACTOR
    F.compute(pos, box, frc)
    f0 = frc.numpy().copy()
    epot, ekin = [], []
    for _ in range(nsteps):
        acc = frc / m
        pos += vel * dt + 0.5 * acc * dt * dt
        vel += 0.5 * dt * acc
        E = F.compute(pos, box, frc, returnDetails=True)
        vel += 0.5 * dt * (frc / m)
        epot.append([sum(float(v) for v in e.values()) for e in E])
        ekin.append((0.5 * (vel * vel).sum(2, keepdim=True) * m).sum(1).reshape(-1).numpy().copy())
    return dict(pos=pos.numpy(), vel=vel.numpy(), forces=frc.numpy(), forces0=f0, epot=np.array(epot),
                ekin=np.array(ekin))


def parse_notebook_kat(path):
    nb = json.load(open(path))
    for c in nb["cells"]:
        for o in c.get("outputs", []):
            text = "".join(o.get("text", []))
            if "'electrostatics': 627" in text:
                head, tail = text.split("]\n", 1)
                ene = eval(head + "]")[0]
                nums = re.findall(r"[-+]?\d\.\d+e[-+]\d+", tail)
                frc = np.array(nums, dtype=np.float64).reshape(-1, 3)
                return np.array([ene[k] for k in _cabi.ENERGY_SLOTS], dtype=np.float64), frc
    raise RuntimeError("known-answer cell not found")


def save(name, par, out):
    arrs = {f"par_{k}": v for k, v in par.to_arrays().items()}
    arrs.update(out)
    path = os.path.join(ROOT, "tests", "golden", name)
    np.savez_compressed(path, **arrs)
    print(f"wrote {path}: {os.path.getsize(path) / 1024:.1f} KiB")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", default="/root/reference")
    a = ap.parse_args()
    REF = a.ref
    rf = load_by_path("ref_forces", f"{REF}/TorchMD_Fix/forces.py")
    rng = np.random.default_rng(20261104)
    VC = f"{REF}/V_Calculations"
    SMD = f"{REF}/All_ML_Training_Data/210905_SMD_decaalanine/SMD/output"

    # ---- BB40: the known-answer system -----------------------------------------------------
    par = parameters_from_psf_yaml(f"{VC}/Test-1/data/backbone-no-improp.psf", f"{VC}/Test-1/data/param_bb-3.0.yaml")
    pdb = read_pdb_coords(f"{VC}/Test-1/data/backbone.pdb").astype(np.float32)
    out = {"pos": pdb[None]}
    out["kat_energies"], out["kat_forces"] = parse_notebook_kat(f"{VC}/Test-1/tutorial.ipynb")
    reference_outputs(rf, par, pdb[None], out)
    # 32 real backbone conformers (xyz columns of the processed phi/psi arrays), both modes
    frames = np.load(f"{SMD}/processed_orient_phi_psi/0.npy")[::32, :, :3][:32].astype(np.float32)
    out["frames"] = frames
    reference_outputs(rf, par, frames, out, tag="frames_")
    vel = (np.sqrt(300 * BOLTZMAN / par.masses.numpy().reshape(1, -1, 1)) * rng.standard_normal((2, 40, 3))).astype(np.float32)
    tr = upstream_vv(rf, par, frames[:2], vel, nsteps=50, dt_fs=1.0)
    out.update({f"vv_{k}": v for k, v in tr.items()})
    out["vv_vel0"], out["vv_dt_fs"], out["vv_nsteps"] = vel, np.array(1.0), np.array(50)
    save("bb40.npz", par, out)

    # ---- BB40 with 9 CHARMM-harmonic impropers (per = 0) --------------------------------------
    par = parameters_from_psf_yaml(f"{VC}/Test-5_bias_n_improper/data/backbone.psf",
                                   f"{VC}/Test-5_bias_n_improper/data/param_bb-4.0.yaml")
    out = {"pos": frames[:8]}
    reference_outputs(rf, par, frames[:8], out)
    # SMD bias as Forces(external=...) on the backbone groups [0], [39] (potcalc.py:99-107)
    ext = SmdExternal(0, 39, 1.0, 13.5)
    for dt_, nm in ((torch.float, "f32"), (torch.double, "f64")):
        f, e, _ = run_reference(rf, par, frames[:8], dt_, external=ext, **MODES["rf"])
        out[f"ref_smd_rf_{nm}_forces"], out[f"ref_smd_rf_{nm}_energies"] = f, e
    out["smd"] = np.array([0, 39, 1.0, 13.5])
    save("bb40_improper.npz", par, out)

    # ---- ALA10 (CHAMBER prmtop, 104 atoms): the benchmark system ------------------------------
    par = parameters_from_prmtop(f"{VC}/Test-6_full_system/data/full_da-1.3.prmtop")
    base = [read_pdb_coords(f"{SMD}/0/my.pdb")]
    vels, kat_dist = [], []
    for r in range(50):
        base.append(read_namd_bin(f"{SMD}/{r}/smd_out.coor"))
        base.append(read_namd_bin(f"{SMD}/{r}/smd_out.restart.coor"))
        vels.append(read_namd_bin(f"{SMD}/{r}/smd_out.vel") * NAMD_VEL_TO_ANGSTROM_PER_PS)
        vels.append(read_namd_bin(f"{SMD}/{r}/smd_out.restart.vel") * NAMD_VEL_TO_ANGSTROM_PER_PS)
        last = open(f"{SMD}/{r}/smd_out.colvars.traj").read().strip().splitlines()[-1].split()
        kat_dist.append([float(last[1]), float(last[2])])
    base = np.array(base)  # (101,104,3) f64
    out = {"base_frames": base, "base_velocities_A_per_ps": np.array(vels), "kat_colvar": np.array(kat_dist)}
    sel = base[[0, 1, 2, 17, 40, 63, 88, 100]].astype(np.float32)
    out["pos"] = sel
    reference_outputs(rf, par, sel, out)
    ext = SmdExternal(3, 98, 1.0, 20.0)
    for dt_, nm in ((torch.float, "f32"), (torch.double, "f64")):
        f, e, _ = run_reference(rf, par, sel, dt_, external=ext)
        out[f"ref_smd_plain_{nm}_forces"], out[f"ref_smd_plain_{nm}_energies"] = f, e
    out["smd"] = np.array([3, 98, 1.0, 20.0])
    vel = (np.sqrt(300 * BOLTZMAN / par.masses.numpy().reshape(1, -1, 1)) * rng.standard_normal((2, 104, 3))).astype(np.float32)
    tr = upstream_vv(rf, par, sel[:2], vel, nsteps=40, dt_fs=1.0)
    out.update({f"vv_{k}": v for k, v in tr.items()})
    out["vv_vel0"], out["vv_dt_fs"], out["vv_nsteps"] = vel, np.array(1.0), np.array(40)
    save("ala10.npz", par, out)

    # ---- alanine dipeptide (22 atoms, CHAMBER) and ALA10 ff14SB (101 atoms, AMBER impropers) ---
    par = parameters_from_prmtop(f"{VC}/Test-10_diALA/data/diALA-1.0.prmtop")
    pos = read_pdb_coords(f"{VC}/Test-10_diALA/data/alad.pdb").astype(np.float32)[None]
    pos = np.concatenate([pos, pos + 0.03 * rng.standard_normal((3,) + pos.shape[1:]).astype(np.float32)])
    out = {"pos": pos}
    reference_outputs(rf, par, pos, out)
    save("diala.npz", par, out)

    par = parameters_from_prmtop(f"{VC}/Test-6_full_system/data/full_da-3.0.prmtop")
    crd = open(f"{VC}/Test-6_full_system/data/full_da-3.0.crd").read().splitlines()
    n = int(crd[1].split()[0])
    vals = [float(l[i:i + 12]) for l in crd[2:] for i in range(0, len(l), 12) if l[i:i + 12].strip()]
    pos = np.array(vals[:3 * n], dtype=np.float32).reshape(1, n, 3)
    pos = np.concatenate([pos, pos + 0.03 * rng.standard_normal((3,) + pos.shape[1:]).astype(np.float32)])
    out = {"pos": pos}
    reference_outputs(rf, par, pos, out)
    save("ala10_ff14sb.npz", par, out)


if __name__ == "__main__":
    main()
