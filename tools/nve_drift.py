#!/usr/bin/env python
"""NVE energy-drift statistics of the fused integrator (BASELINE.json configs[2] check) next to the FP64 oracle.

    python tools/nve_drift.py [--replicas 64] [--steps 100000] [--oracle-steps 10000] [--dt 1.0]

Prints one JSON line: per-replica drift of E_tot = E_pot + E_kin (least-squares slope x run length, kcal/mol),
RMS fluctuation about the fit, for the CUDA path over the full run and for both paths over the oracle's horizon.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def drift_stats(etot):
    """etot (reports, R) -> (drift per replica over the run, rms about the linear fit)."""
    n = etot.shape[0]
    t = np.arange(n, dtype=np.float64)
    t -= t.mean()
    slope = (t[:, None] * (etot - etot.mean(0))).sum(0) / (t * t).sum()
    fit = etot.mean(0) + slope * t[:, None]
    return slope * (n - 1), np.sqrt(((etot - fit) ** 2).mean(0))


def compare_ensembles(gpu, ora):
    """(drift, rms) per replica of two ensembles -> difference of the ensemble means with its standard error
    sqrt(s_g^2 / n_g + s_o^2 / n_o): the yardstick of "the drift matches the reference"."""
    (dg, rg), (do, ro) = gpu, ora
    se = lambda a, b: float(np.sqrt(a.var(ddof=1) / len(a) + b.var(ddof=1) / len(b)))
    return {"gpu_drift_mean": float(dg.mean()), "gpu_drift_std": float(dg.std(ddof=1)), "gpu_rms_mean": float(rg.mean()),
            "oracle_drift_mean": float(do.mean()), "oracle_drift_std": float(do.std(ddof=1)), "oracle_rms_mean": float(ro.mean()),
            "n_gpu": int(len(dg)), "n_oracle": int(len(do)),
            "drift_diff": float(dg.mean() - do.mean()), "drift_diff_se": se(dg, do),
            "rms_diff": float(rg.mean() - ro.mean()), "rms_diff_se": se(rg, ro)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--replicas", type=int, default=8192)
    ap.add_argument("--steps", type=int, default=100000)
    ap.add_argument("--oracle-steps", type=int, default=100000)
    ap.add_argument("--oracle-replicas", type=int, default=16)
    ap.add_argument("--dt", type=float, default=1.0)
    ap.add_argument("--every", type=int, default=200)
    ap.add_argument("--copies", type=int, default=1,
                    help="> 1: a bundle of this many chains 40 A apart as ONE topology (Parameters.tiled): the large-solute kernels")
    ap.add_argument("--rigid", action="store_true", help="rigid X-H bonds (SHAKE / RATTLE), e.g. with --dt 2.0")
    a = ap.parse_args()
    import torch

    import bench
    import oracle
    from molecular_dynamics_b200 import Forces, Integrator, Parameters, System, _cabi

    g = bench.load_ala10()
    dev = torch.device("cuda", 0)
    par = Parameters.from_arrays({k[4:]: v for k, v in g.items() if k.startswith("par_")}, device=dev)
    pos, vel = bench.synth_inputs(g, a.replicas * a.copies)
    par_cpu = Parameters.from_arrays({k[4:]: v for k, v in g.items() if k.startswith("par_")})
    pair_idx = g["ava_idx"]
    if a.copies > 1:
        par, par_cpu = par.tiled(a.copies), par_cpu.tiled(a.copies)
        shift = (np.repeat(np.arange(a.copies), 104)[None, :, None] * np.array([40.0, 0.0, 0.0])).astype(np.float32)
        pos = pos.reshape(a.replicas, a.copies * 104, 3) + shift
        vel = vel.reshape(a.replicas, a.copies * 104, 3)
        pair_idx = _cabi.make_pair_index(104 * a.copies, par_cpu.get_exclusions())
    s = System(104 * a.copies, a.replicas, torch.float, dev)
    s.pos[:] = torch.tensor(pos, device=dev)
    s.vel[:] = torch.tensor(vel, device=dev)
    F = Forces(par)
    F.compute_batch(s.pos, s.box, s.forces)
    f0 = s.forces.cpu().numpy().astype(np.float64)
    integ = Integrator(s, F, timestep=a.dt, device=dev, rigid_bonds=a.rigid)
    if a.rigid:  # the synthetic start is not on the constraint surface: project it, then restart the clock
        integ.step(niter=1)
        F.compute_batch(s.pos, s.box, s.forces)
        f0 = s.forces.cpu().numpy().astype(np.float64)
        pos, vel = s.pos.cpu().numpy(), s.vel.cpu().numpy()
    integ.energy_every = a.every
    chunks = []
    done = 0
    while done < a.steps:  # launches of <= 20 000 steps (observable buffer stays small)
        n = min(20000, a.steps - done)
        integ.step(niter=n)
        chunks.append(integ.observables[:, :, :_cabi.OBS_EKIN + 1].cpu().numpy())
        done += n
    obs = np.concatenate(chunks, 0)
    etot = obs[:, :, :8].sum(-1) + obs[:, :, _cabi.OBS_EKIN]
    d, r = drift_stats(etot)
    out = {"dt_fs": a.dt, "natoms": 104 * a.copies, "rigid_bonds": bool(a.rigid), "steps": a.steps, "replicas": a.replicas, "energy_every": a.every,
           "gpu": {"etot0_mean": float(etot[0].mean()), "ekin_mean": float(obs[:, :, _cabi.OBS_EKIN].mean()),
                   "drift_mean": float(d.mean()), "drift_abs_max": float(np.abs(d).max()), "drift_std": float(d.std()),
                   "rms_fluct_mean": float(r.mean()), "rms_fluct_max": float(r.max())}}
    if a.oracle_steps > 0:
        packed = _cabi.PackedTopology(par_cpu, pair_idx)
        ro = a.oracle_replicas
        o = oracle.integrate(packed, pos[:ro], vel[:ro], f0[:ro], nsteps=a.oracle_steps, dt=np.float32(integ.dt), energy_every=a.every,
                             rigid_bonds=a.rigid)
        eo = o["observables"][:, :, :8].sum(-1) + o["observables"][:, :, _cabi.OBS_EKIN]
        do, ro_ = drift_stats(eo)
        nrep = a.oracle_steps // a.every
        dg, rg = drift_stats(etot[:nrep, :ro])
        out["oracle_fp64"] = {"steps": a.oracle_steps, "replicas": ro, "drift_mean": float(do.mean()), "drift_abs_max": float(np.abs(do).max()),
                              "rms_fluct_mean": float(ro_.mean())}
        out["gpu_same_horizon"] = {"drift_mean": float(dg.mean()), "drift_abs_max": float(np.abs(dg).max()), "rms_fluct_mean": float(rg.mean()),
                                   "etot_first_report_max_abs_diff": float(np.abs(etot[0, :ro] - eo[0]).max())}
        # the whole GPU ensemble over the oracle's horizon against the oracle's ensemble: difference of means, standard error
        out["ensemble_comparison"] = compare_ensembles(drift_stats(etot[:nrep]), (do, ro_))
    print(json.dumps(out))


if __name__ == "__main__":
    main()
