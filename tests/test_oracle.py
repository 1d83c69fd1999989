"""Pin the CPU oracle (oracle/md_oracle.c) before anything is compared with it.

Anchors, in order of authority:
  1. the notebook known-answer vector of the reference (V_Calculations/Test-1/tutorial.ipynb:231-271);
  2. outputs of the reference's unmodified TorchMD_Fix/forces.py in fp64 / fp32 (tests/golden/*.npz,
     produced by tools/make_golden.py);
  3. the colvar known answers (smd_out.colvars.traj) for the SMD end-to-end distance;
  4. the restated upstream velocity-Verlet loop driving the reference forces.py (unpinned upstream).
"""
import os

import numpy as np
import pytest

import oracle
from conftest import MODES, SYSTEMS, load_golden
from molecular_dynamics_b200 import _cabi

TIMEFACTOR = 48.88821


def test_known_answer_vector_energies_and_forces():
    g = load_golden("bb40")
    f, e = oracle.forces(g.packed, g["pos"], **MODES["rf"])
    kat_e, kat_f = g["kat_energies"], g["kat_forces"]
    # stored at full fp32 precision: the fp64 oracle must agree to fp32 rounding of the reference path
    np.testing.assert_allclose(e[0], kat_e, rtol=2e-6, atol=1e-9)
    assert kat_f.shape == (40, 3)
    # forces are printed to 5 significant figures
    np.testing.assert_allclose(f[0], kat_f, rtol=2e-4, atol=2e-4 * np.abs(kat_f).max() * 1e-1)
    # the reference itself, rerun here in fp32, reproduced the notebook bit-for-bit
    np.testing.assert_array_equal(g["ref_rf_f32_energies"][0].astype(np.float32), kat_e.astype(np.float32))


@pytest.mark.parametrize("system", SYSTEMS)
@pytest.mark.parametrize("mode", list(MODES))
def test_oracle_matches_reference_fp64(system, mode):
    g = load_golden(system)
    f, e = oracle.forces(g.packed, g["pos"], **MODES[mode])
    ref_f, ref_e = g[f"ref_{mode}_f64_forces"], g[f"ref_{mode}_f64_energies"]
    np.testing.assert_allclose(e, ref_e, rtol=1e-11, atol=1e-10)
    assert np.abs(f - ref_f).max() <= 1e-11 * np.abs(ref_f).max()
    # and sits within fp32 rounding of the reference's fp32 run
    ref32 = g[f"ref_{mode}_f32_forces"]
    assert np.abs(f - ref32).max() <= 2e-5 * np.abs(ref_f).max()


def test_oracle_batched_frames_bb40():
    g = load_golden("bb40")
    for mode in MODES:
        f, e = oracle.forces(g.packed, g["frames"], **MODES[mode])
        np.testing.assert_allclose(e, g[f"frames_ref_{mode}_f64_energies"], rtol=1e-11, atol=1e-10)
        assert np.abs(f - g[f"frames_ref_{mode}_f64_forces"]).max() < 1e-9


def test_pair_index_is_bit_exact():
    """Forces._make_indeces (forces.py:346-355) restated in _cabi.make_pair_index."""
    for system in SYSTEMS:
        g = load_golden(system)
        mine = _cabi.make_pair_index(g.par.natoms, g.par.get_exclusions(("bonds", "angles", "1-4")))
        assert mine.dtype == np.int64
        np.testing.assert_array_equal(mine, g.ava_idx)


def test_smd_bias_energy_and_force():
    g = load_golden("ala10")
    a1, a2, fk, c = g["smd"]
    smd = dict(atom1=int(a1), atom2=int(a2), fk=fk, center0=c)
    f, e = oracle.forces(g.packed, g["pos"], smd=smd)
    np.testing.assert_allclose(e, g["ref_smd_plain_f64_energies"], rtol=1e-11, atol=1e-10)
    assert np.abs(f - g["ref_smd_plain_f64_forces"]).max() < 1e-9
    g = load_golden("bb40_improper")
    a1, a2, fk, c = g["smd"]
    smd = dict(atom1=int(a1), atom2=int(a2), fk=fk, center0=c)
    f, e = oracle.forces(g.packed, g["pos"], smd=smd, **MODES["rf"])
    np.testing.assert_allclose(e, g["ref_smd_rf_f64_energies"], rtol=1e-11, atol=1e-10)
    assert np.abs(f - g["ref_smd_rf_f64_forces"]).max() < 1e-9


def test_colvar_known_answers():
    """|r98 - r3| of each run's final frame equals the End2EndDist NAMD logged (50 runs)."""
    g = load_golden("ala10")
    base, kat = g["base_frames"], g["kat_colvar"]
    for r in range(50):
        final = base[1 + 2 * r]
        d = np.linalg.norm(final[98] - final[3])
        assert abs(d - kat[r, 0]) < 5e-7 * kat[r, 0]
        assert kat[r, 1] == 34.0  # moving centre reached 12 + 22 * step/500000


def test_per_replica_smd_centres():
    g = load_golden("ala10")
    pos = g["pos"]
    cen = np.linspace(12, 34, len(pos)).astype(np.float32)
    smd = dict(atom1=3, atom2=98, fk=1.0, center0=0.0)
    _, e = oracle.forces(g.packed, pos, smd=smd, smd_centers=cen)
    d = np.linalg.norm(pos[:, 98].astype(np.float64) - pos[:, 3], axis=1)
    np.testing.assert_allclose(e[:, 7], 0.5 * (cen - d) ** 2, rtol=1e-12)


def test_nan_guard():
    g = load_golden("diala")
    pos = g["pos"].copy()
    pos[1, 3, 2] = np.nan
    with pytest.raises(RuntimeError, match="Found NaN coordinates."):
        oracle.forces(g.packed, pos)


@pytest.mark.parametrize("system", ["bb40", "ala10"])
def test_oracle_integrator_matches_restated_upstream_loop(system):
    g = load_golden(system)
    n = len(g["vv_vel0"])
    pos0 = (g["frames"] if system == "bb40" else g["pos"])[:n]
    f0, _ = oracle.forces(g.packed, pos0)
    assert np.abs(f0 - g["vv_forces0"]).max() < 1e-9
    nsteps, dt = int(g["vv_nsteps"]), float(g["vv_dt_fs"]) / TIMEFACTOR
    out = oracle.integrate(g.packed, pos0, g["vv_vel0"], f0, nsteps=nsteps, dt=np.float32(dt), energy_every=1)
    # dt is rounded to fp32 at the C-ABI (as the GPU path sees it): 6e-8 relative on dt
    np.testing.assert_allclose(out["pos"], g["vv_pos"], atol=2e-6)
    np.testing.assert_allclose(out["vel"], g["vv_vel"], rtol=0, atol=2e-6 * np.abs(g["vv_vel"]).max() * 10)
    epot = out["observables"][:, :, :8].sum(-1)
    np.testing.assert_allclose(epot, g["vv_epot"], rtol=1e-6, atol=1e-5)
    np.testing.assert_allclose(out["observables"][:, :, _cabi.OBS_EKIN], g["vv_ekin"], rtol=1e-5)
    # NVE: total energy conserved over the short run
    etot = epot + out["observables"][:, :, _cabi.OBS_EKIN]
    assert np.abs(etot - etot[0]).max() < 2.0  # 1 fs VV fluctuation (free X-H bonds), no drift


def test_philox_known_answer_and_normals():
    # Random123 known-answer test for philox4x32-10, zero counter and key
    np.testing.assert_array_equal(oracle.philox(0, 0, 0, 0, 0),
                                  np.array([0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8], dtype=np.uint32))
    # Random123 kat: counter/key all ones
    np.testing.assert_array_equal(oracle.philox(0xFFFFFFFFFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFF),
                                  np.array([0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD], dtype=np.uint32))
    xs = np.array([oracle.normal3(7, r, a, s) for r in range(8) for a in range(20) for s in range(20)])
    assert abs(xs.mean()) < 0.03 and abs(xs.std() - 1.0) < 0.03


def test_langevin_thermostat_reaches_temperature():
    g = load_golden("diala")
    pos0 = np.repeat(g["pos"][:1], 16, axis=0)
    f0, _ = oracle.forces(g.packed, pos0)
    dt, gamma = 1.0 / TIMEFACTOR, 10.0 / (1000.0 / TIMEFACTOR)
    out = oracle.integrate(g.packed, pos0, np.zeros_like(pos0), f0, nsteps=4000, dt=dt, gamma=gamma,
                           temperature=300.0, seed=3, energy_every=100)
    ekin = out["observables"][10:, :, _cabi.OBS_EKIN]
    T = 2.0 / (3 * 22 * 0.001987191) * ekin.mean()
    assert 255 < T < 345


@pytest.mark.parametrize("system", SYSTEMS)
def test_torch_port_matches_reference(system):
    """oracle/torchmd_port.py (the timed CPU baseline) against the reference's own outputs."""
    import torch

    from oracle import torchmd_port as port

    g = load_golden(system)
    for mode, kw in MODES.items():
        top = port.PortTopology(g.packed, torch.float64)
        pos = torch.tensor(g["pos"]).double()
        frc = torch.zeros_like(pos)
        ene = port.compute(top, pos, frc, **kw).numpy()
        np.testing.assert_allclose(ene, g[f"ref_{mode}_f64_energies"], rtol=1e-10, atol=1e-9)
        assert np.abs(frc.numpy() - g[f"ref_{mode}_f64_forces"]).max() < 1e-9
        top32 = port.PortTopology(g.packed, torch.float32)
        pos32 = torch.tensor(g["pos"])
        frc32 = torch.zeros_like(pos32)
        ene32 = port.compute(top32, pos32, frc32, **kw).numpy()
        ref32 = g[f"ref_{mode}_f32_forces"]
        assert np.abs(frc32.numpy() - ref32).max() <= 2e-5 * np.abs(ref32).max()
        np.testing.assert_allclose(ene32, g[f"ref_{mode}_f32_energies"], rtol=2e-5, atol=1e-4)


def test_torch_port_integrator_matches_c_oracle():
    import torch

    from oracle import torchmd_port as port

    g = load_golden("diala")
    pos0 = g["pos"]
    vel0 = (0.05 * np.random.default_rng(0).standard_normal(pos0.shape)).astype(np.float32)
    f0, _ = oracle.forces(g.packed, pos0)
    dt = float(np.float32(1.0 / TIMEFACTOR))
    out = oracle.integrate(g.packed, pos0, vel0, f0, nsteps=25, dt=dt)
    top = port.PortTopology(g.packed, torch.float64)
    p, v, f = torch.tensor(pos0).double(), torch.tensor(vel0).double(), torch.tensor(f0)
    ekin, ene = port.integrate(top, p, v, f, 25, dt)
    np.testing.assert_allclose(p.numpy(), out["pos"], atol=1e-9)
    np.testing.assert_allclose(ekin.numpy(), out["observables"][0, :, _cabi.OBS_EKIN], rtol=1e-9)


REPULSION_CASES = {  # tools/make_golden_repulsion.py
    "rep_plain": (("repulsion",), {}),
    "rep_cut": (("repulsion", "bonds"), {"cutoff": 9.0, "switch_dist": 7.5}),
    "cg_plain": (("repulsioncg", "electrostatics", "bonds"), {}),
    "cg_rf": (("repulsioncg", "electrostatics", "angles"), {"cutoff": 9.0, "rfa": True, "switch_dist": 7.5}),
    # several pair terms of the lj family at once (forces.py:268-316 evaluates each enabled one; only "lj" is switched)
    "combo_plain": (("lj", "repulsion", "electrostatics", "bonds"), {}),
    "combo_cut": (("repulsioncg", "lj", "repulsion", "angles"), {"cutoff": 9.0, "switch_dist": 7.5}),
}


@pytest.mark.parametrize("system", ["bb40", "ala10"])
@pytest.mark.parametrize("case", sorted(REPULSION_CASES))
def test_oracle_repulsion_terms_vs_reference(system, case):
    """evaluate_repulsion / evaluate_repulsion_CG (forces.py:416-448) against the unmodified
    reference in fp64; the ABI reports the enabled lj-family term in the "lj" column."""
    g = load_golden(system)
    with np.load(os.path.join(os.path.dirname(__file__), "golden", "repulsion.npz")) as z:
        r = {k: z[k] for k in z.files if k.startswith(system)}
    terms, kw = REPULSION_CASES[case]
    f, e = oracle.forces(g.packed, r[f"{system}_pos"], terms=terms, **kw)
    fr = r[f"{system}_{case}_f64_forces"]
    assert np.abs(f - fr).max() <= 1e-9 * np.abs(fr).max()
    family = [t for t in terms if t in ("lj", "repulsion", "repulsioncg")]
    for t in terms:
        if t not in family:
            np.testing.assert_allclose(e[:, _cabi.energy_slot(t)], r[f"{system}_{case}_f64_E_{t}"], rtol=1e-10)
    # the oracle (like one launch of the ABI) books every enabled term of the lj family in the "lj" column
    np.testing.assert_allclose(e[:, 1], sum(r[f"{system}_{case}_f64_E_{t}"] for t in family), rtol=1e-10)


@pytest.mark.parametrize("system", ["bb40", "ala10"])
@pytest.mark.parametrize("mode", sorted(MODES))
def test_oracle_minimum_image_vs_reference(system, mode):
    """wrap_dist (forces.py:358-363) in every term, boxes smaller than the molecule, one replica in
    vacuum (all-zero box) in the same batch; unmodified reference in fp64 (tools/make_golden_pbc.py)."""
    g = load_golden(system)
    with np.load(os.path.join(os.path.dirname(__file__), "golden", "pbc.npz")) as z:
        r = {k: z[k] for k in z.files if k.startswith(system)}
    f, e = oracle.forces(g.packed, r[f"{system}_pos"], box=r[f"{system}_box"], **MODES[mode])
    fr, er = r[f"{system}_{mode}_f64_forces"], r[f"{system}_{mode}_f64_energies"]
    assert np.abs(f - fr).max() <= 1e-9 * np.abs(fr).max()
    np.testing.assert_allclose(e, er, rtol=1e-10, atol=1e-9)


def test_smd_work_definition_matches_reference_colvars_record():
    """W_harmonic1 of the reference's own NAMD/colvars record (tests/golden/smd_out_head.colvars.traj, copied by
    tools/make_golden_colvars.py) is sum fk (x0 - x) dx0 with fk = 1: the definition OBS_SMD_WORK restates.  The
    record is sampled every 50 steps, so the check integrates by trapezoids and allows the sampling error."""
    from molecular_dynamics_b200.utils import read_colvars_traj

    steps, d, c, w = read_colvars_traj(os.path.join(os.path.dirname(__file__), "golden", "smd_out_head.colvars.traj"))
    dev = c - d
    trap = np.concatenate([[0.0], np.cumsum(0.5 * (dev[1:] + dev[:-1]) * np.diff(c))])
    assert np.abs(trap - w).max() < 2e-3 and abs(w[-1]) > 0.05
    assert np.sign(w[1]) == np.sign(dev[0])


def test_oracle_smd_work_is_the_running_sum_of_its_own_records():
    g = load_golden("ala10")
    pos0 = g["pos"][:3]
    vel0 = (0.05 * np.random.default_rng(11).standard_normal(pos0.shape)).astype(np.float32)
    rate, fk, c0 = np.float32(0.01), 2.5, 13.0
    smd = dict(atom1=3, atom2=98, fk=fk, center0=c0, center_rate=rate)
    f0, _ = oracle.forces(g.packed, pos0, smd=dict(smd, center_rate=0.0))
    out = oracle.integrate(g.packed, pos0, vel0, f0, nsteps=40, dt=np.float32(1.0 / TIMEFACTOR), energy_every=1, smd=smd)
    obs = out["observables"]
    centers = c0 + float(rate) * np.arange(1, 41)[:, None]
    expect = np.cumsum(fk * (centers - obs[:, :, _cabi.OBS_E2E_DIST]) * float(rate), axis=0)
    np.testing.assert_allclose(obs[:, :, _cabi.OBS_SMD_WORK], expect, rtol=1e-12, atol=1e-14)
    # the restraint energy recorded in the external slot is 1/2 fk (c - d)^2 at the same centre
    np.testing.assert_allclose(obs[:, :, 7], 0.5 * fk * (centers - obs[:, :, _cabi.OBS_E2E_DIST]) ** 2, rtol=1e-10)
    # without a bias the column stays zero
    out0 = oracle.integrate(g.packed, pos0, vel0, f0, nsteps=5, dt=np.float32(1.0 / TIMEFACTOR))
    assert not out0["observables"][:, :, _cabi.OBS_SMD_WORK].any()


def _chamber_system(**kw):
    import torch

    from molecular_dynamics_b200.parameters import parameters_from_prmtop
    from molecular_dynamics_b200.topology_io import read_pdb_coords

    gold = os.path.join(os.path.dirname(__file__), "golden")
    par = parameters_from_prmtop(os.path.join(gold, "full_da-1.3.prmtop"), precision=torch.float64, **kw)
    pos = read_pdb_coords(os.path.join(gold, "smd_ini.pdb"))[None].astype(np.float64)
    pidx = _cabi.make_pair_index(par.natoms, par.get_exclusions(("bonds", "angles", "1-4")))
    return par, pos, _cabi.PackedTopology(par, pidx)


def test_chamber_prmtop_route_against_namd_step0_energies():
    """Loose external anchor of the prmtop -> Parameters route (otherwise "unpinned"): NAMD's own
    energies of the reference's start structure with the same CHARMM22 force field
    (tests/golden/namd_step0_energies.json, made by tools/make_golden_chamber.py).  NAMD constrains the X-H bonds
    before evaluating (rigidBonds ALL), so BOND is not comparable, ELECT contains the GB term, and the rest agrees to
    a fraction of a percent: enough to catch a wrong unit, phase convention, 1-4 table or dropped term family."""
    import json

    namd = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "namd_step0_energies.json")))["energies"]
    slot = _cabi.energy_slot
    # the reference's behaviour: CHARMM impropers and Urey-Bradley silently dropped
    par, pos, packed = _chamber_system()
    assert par.impropers is None and len(par.bonds) == 103
    _, e0 = oracle.forces(packed, pos)
    assert e0[0, slot("impropers")] == 0.0
    assert abs(e0[0, slot("dihedrals")] - namd["DIHED"]) < 1e-3 * namd["DIHED"]
    # opt-in: all CHAMBER term families
    par, pos, packed = _chamber_system(charmm_impropers=True, urey_bradley=True)
    assert len(par.impropers) == 22 and len(par.bonds) == 103 + 65
    assert bool((par.improper_params[0]["params"][:, 2] == 0).all())
    _, e = oracle.forces(packed, pos)
    assert abs(e[0, slot("impropers")] - namd["IMPRP"]) < 5e-3 * namd["IMPRP"]
    np.testing.assert_allclose(e[0, slot("dihedrals")], e0[0, slot("dihedrals")], rtol=1e-14)
    ub = e[0, slot("bonds")] - e0[0, slot("bonds")]
    assert 4.0 < ub < 7.0  # 65 Urey-Bradley terms
    assert abs(e[0, slot("angles")] + ub - namd["ANGLE"]) < 0.03 * namd["ANGLE"]  # hydrogens moved by the constraints
    # VDW = LJ + 1-4 LJ (CHARMM 1-4 table); evaluate without charges so the slot holds only that
    par.charges = par.charges * 0
    _, ev = oracle.forces(_cabi.PackedTopology(par, _cabi.make_pair_index(par.natoms, par.get_exclusions())), pos)
    assert ev[0, slot("electrostatics")] == 0.0
    assert abs(ev[0, slot("lj")] - namd["VDW"]) < 5e-3 * abs(namd["VDW"])
    # the exclusion set is unchanged by the Urey-Bradley rows (they are 1-3 pairs)
    par0, _, _ = _chamber_system()
    a = _cabi.make_pair_index(par.natoms, par.get_exclusions())
    b = _cabi.make_pair_index(par0.natoms, par0.get_exclusions())
    np.testing.assert_array_equal(a, b)


def _restart_state():
    from molecular_dynamics_b200.topology_io import read_namd_bin

    gold = os.path.join(os.path.dirname(__file__), "golden")
    return (read_namd_bin(os.path.join(gold, "smd_out.restart.coor"))[None],
            read_namd_bin(os.path.join(gold, "smd_out.restart.vel"))[None])


def test_oracle_rigid_bonds_hold_and_conserve_energy_at_2fs():
    """rigid_bonds (NAMD rigidBonds all) from the reference's own restart of replica 3, which satisfies the same
    constraint set: the 53 X-H distances stay at req, the velocities carry no component along them, and the
    integrator stays second order in dt."""
    par, _, packed = _chamber_system()
    pos0, vel0 = _restart_state()
    f0, _ = oracle.forces(packed, pos0)
    dt = np.float32(2.0 / TIMEFACTOR)
    out = oracle.integrate(packed, pos0, vel0, f0, nsteps=400, dt=dt, energy_every=10, rigid_bonds=True)
    b, bp, m = par.bonds.numpy(), par.bond_params.numpy(), par.masses.numpy().ravel()
    bp = bp.astype(np.float32).astype(np.float64)  # the C-ABI description carries float parameters
    xh = (m[b[:, 0]] < 1.5) | (m[b[:, 1]] < 1.5)
    x, v = out["pos"][0], out["vel"][0]
    rij = x[b[xh, 0]] - x[b[xh, 1]]
    assert np.abs(np.linalg.norm(rij, axis=1) - bp[xh, 1]).max() < 1e-9
    assert np.abs((rij * (v[b[xh, 0]] - v[b[xh, 1]])).sum(1)).max() < 1e-9
    assert np.abs(np.linalg.norm(x[b[~xh, 0]] - x[b[~xh, 1]], axis=1) - bp[~xh, 1]).max() > 1e-3  # the rest is free
    # SHAKE + RATTLE keep velocity Verlet second order: the energy fluctuation of the same 0.8 ps scales as dt^2
    # (measured 0.19 / 0.87 kcal/mol at 1 / 2 fs on ~77 kcal/mol of kinetic energy) and does not drift away
    etot = out["observables"][:, 0, :8].sum(-1) + out["observables"][:, 0, _cabi.OBS_EKIN]
    half = oracle.integrate(packed, pos0, vel0, f0, nsteps=800, dt=np.float32(1.0 / TIMEFACTOR), energy_every=20,
                            rigid_bonds=True)
    etot_half = half["observables"][:, 0, :8].sum(-1) + half["observables"][:, 0, _cabi.OBS_EKIN]
    assert np.ptp(etot_half) < 0.3 and 2.5 < np.ptp(etot) / np.ptp(etot_half) < 6.5
    # a topology without hydrogens: the option is a no-op, bit for bit
    g = load_golden("bb40")
    p0, v0 = g["frames"][:2], np.zeros_like(g["frames"][:2])
    f0, _ = oracle.forces(g.packed, p0)
    a = oracle.integrate(g.packed, p0, v0, f0, nsteps=5, dt=dt, rigid_bonds=True)
    c = oracle.integrate(g.packed, p0, v0, f0, nsteps=5, dt=dt)
    np.testing.assert_array_equal(a["pos"], c["pos"])


def test_oracle_constrain_projects_a_perturbed_start():
    par, _, packed = _chamber_system()
    pos0, vel0 = _restart_state()
    rng = np.random.default_rng(3)
    bad = pos0 + 0.05 * rng.standard_normal(pos0.shape)
    p, v = oracle.constrain(packed, bad, vel0)
    b, bp, m = par.bonds.numpy(), par.bond_params.numpy().astype(np.float32).astype(np.float64), par.masses.numpy().ravel()
    xh = (m[b[:, 0]] < 1.5) | (m[b[:, 1]] < 1.5)
    rij = p[0, b[xh, 0]] - p[0, b[xh, 1]]
    assert np.abs(np.linalg.norm(rij, axis=1) - bp[xh, 1]).max() < 1e-9
    assert np.abs((rij * (v[0, b[xh, 0]] - v[0, b[xh, 1]])).sum(1)).max() < 1e-9
    # mass-weighted: the centre of mass of every group stays, so does the molecule's (masses as the C-ABI carries them)
    mf = m.astype(np.float32).astype(np.float64)
    np.testing.assert_allclose((mf[:, None] * p[0]).sum(0), (mf[:, None] * bad[0]).sum(0), rtol=0, atol=1e-8)
    np.testing.assert_allclose((mf[:, None] * v[0]).sum(0), (mf[:, None] * vel0[0]).sum(0), rtol=0, atol=1e-8)
    # a structure that is already on the surface does not move
    p2, v2 = oracle.constrain(packed, p, v)
    assert np.abs(p2 - p).max() < 1e-10 and np.abs(v2 - v).max() < 1e-10


def _phi_psi_fixture():
    """frames (64, 40, 3) and, per frame, the reference data set's own phi (residues 1..8) and psi (residues 0..7) in
    degrees, with the rows of par.dihedrals that are those angles (backbone order N, CA, C, O per residue)."""
    g = load_golden("bb40")
    a = np.load(os.path.join(os.path.dirname(__file__), "golden", "bb40_phi_psi.npy"))
    quads = [tuple(int(x) for x in q) for q in g.par.dihedrals.numpy()]
    rows, want = [], []
    for r in range(8):
        rows.append(quads.index((4 * r, 4 * r + 1, 4 * r + 2, 4 * r + 4)))  # psi_r: N CA C N'
        want.append(a[:, 4 * r, 4])
        rows.append(quads.index((4 * r + 2, 4 * r + 4, 4 * r + 5, 4 * r + 6)))  # phi_{r+1}: C N' CA' C'
        want.append(a[:, 4 * (r + 1), 3])
    return g, a[:, :, :3], np.array(rows), np.stack(want, 1).astype(np.float64)


def test_dihedral_angles_against_the_reference_data_sets_phi_psi():
    """Energy.calc_phi restated (oracle_dihedral_angles) reproduces the phi / psi columns the reference's data set
    stores beside its backbone coordinates (tests/golden/bb40_phi_psi.npy, tools/make_golden_phipsi.py): sign
    convention, atom order and units pinned on 64 frames x 16 angles."""
    g, frames, rows, want = _phi_psi_fixture()
    phi = np.degrees(oracle.dihedral_angles(g.packed, frames.astype(np.float64)))[:, rows]
    diff = (phi - want + 180.0) % 360.0 - 180.0
    assert np.abs(diff).max() < 0.02, np.abs(diff).max()
    assert np.abs(want).max() > 150.0 and (want > 0).any() and (want < 0).any()  # both signs, near the branch cut


def test_chamber_prmtop_route_pinned_by_namd_energies_at_the_restart_frame():
    """Tight external anchor of prmtop -> Parameters -> forces: NAMD's own ENERGY line of step
    500000 of replica 3 (tests/golden/namd_step0_energies.json, "restart_step_500000") belongs to the binary restart
    files in tests/golden/.  Those coordinates satisfy NAMD's rigid bonds, so the bonded columns compare directly:
    BOND, ANGLE (= angles + Urey-Bradley), DIHED, IMPRP to the 4 decimals NAMD prints, and KINETIC / TEMP pin the
    velocity unit and the 3 N - 53 degrees of freedom.  VDW differs by NAMD's force-switched 15-16 A cutoff; ELECT
    contains the generalised-Born term."""
    import json

    from molecular_dynamics_b200 import BOLTZMAN

    namd = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "namd_step0_energies.json")))
    namd = namd["restart_step_500000"]["energies"]
    pos, vel = _restart_state()
    slot = _cabi.energy_slot
    par0, _, packed0 = _chamber_system()
    par, _, packed = _chamber_system(charmm_impropers=True, urey_bradley=True)
    _, e0 = oracle.forces(packed0, pos)
    _, e = oracle.forces(packed, pos)
    ub = e[0, slot("bonds")] - e0[0, slot("bonds")]
    assert abs(e0[0, slot("bonds")] - namd["BOND"]) < 2e-4
    assert abs(e[0, slot("angles")] + ub - namd["ANGLE"]) < 2e-4
    assert abs(e[0, slot("dihedrals")] - namd["DIHED"]) < 2e-4
    assert abs(e[0, slot("impropers")] - namd["IMPRP"]) < 2e-4
    par.charges = par.charges * 0
    _, ev = oracle.forces(_cabi.PackedTopology(par, _cabi.make_pair_index(par.natoms, par.get_exclusions())), pos)
    assert abs(ev[0, slot("lj")] - namd["VDW"]) < 0.1  # no switching function here
    m = par.masses.numpy().ravel()
    ekin = 0.5 * (m[:, None] * vel[0] ** 2).sum()
    assert abs(ekin - namd["KINETIC"]) < 2e-4
    assert abs(2.0 * ekin / ((3 * 104 - 53) * BOLTZMAN) - namd["TEMP"]) < 2e-3


@pytest.mark.parametrize("mode", list(MODES))
def test_oracle_matches_reference_fp64_above_128_atoms(mode):
    """tests/golden/ala10x3.npz (tools/make_golden_large.py): the unmodified reference on a 312-atom solute -- three
    deca-alanine chains in one topology -- pins the oracle (and Parameters.tiled, make_pair_index) where the large-solute
    CUDA kernels run."""
    g = load_golden("ala10x3")
    assert g.par.natoms == 312
    mine = _cabi.make_pair_index(312, g.par.get_exclusions(("bonds", "angles", "1-4")))
    np.testing.assert_array_equal(mine, g.ava_idx)
    f, e = oracle.forces(g.packed, g["pos"], **MODES[mode])
    ref_f, ref_e = g[f"ref_{mode}_f64_forces"], g[f"ref_{mode}_f64_energies"]
    np.testing.assert_allclose(e, ref_e, rtol=1e-11, atol=1e-10)
    assert np.abs(f - ref_f).max() <= 1e-11 * np.abs(ref_f).max()
