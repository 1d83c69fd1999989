"""Parity of the CUDA path (through the drop-in classes -> custom op -> C-ABI) against the oracle and the
golden fixtures produced by the reference's own forces.py.

Tolerances (BASELINE.json north_star, calibrated in SURVEY.md 8c):
  * integer artefacts (pair list, exclusion mask)             bit-exact
  * forces   max|dF| / max|F| per replica                     <= 1e-5
  * energies per term and total vs fp64 reference              <= 1e-6 relative, every slot (the kernels evaluate
    energies in FP64 from the fp32 positions; forces stay fp32).  Achieved values: profiles/*parity_report.md.
"""
import os

import numpy as np
import pytest
import torch

import oracle
from conftest import MODES, SYSTEMS, load_golden
from molecular_dynamics_b200 import Forces, Integrator, SMDBias, System, _cabi
from molecular_dynamics_b200.parameters import Parameters

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TIMEFACTOR = 48.88821
F_TOL = 1e-5
E_TOL = 1e-6


def gpu_par(g):
    return Parameters.from_arrays(g.par.to_arrays(), device=DEV)


def run_gpu(g, pos, smd=None, centers=None, **kw):
    par = gpu_par(g)
    ext = None
    if smd is not None:
        ext = SMDBias([smd["atom1"]], [smd["atom2"]], fk=smd["fk"], cent_0=smd["center0"])
        if centers is not None:
            ext.center = torch.tensor(centers, device=DEV)
    F = Forces(par, external=ext, **kw)
    p = torch.tensor(np.asarray(pos, dtype=np.float32), device=DEV)
    R, N, _ = p.shape
    s = System(N, R, torch.float, DEV)
    ene = F.compute_batch(p, s.box, s.forces)
    torch.cuda.synchronize()
    return s.forces.cpu().numpy().astype(np.float64), ene.cpu().numpy(), F


def check_energies(e, e_ref, what=""):
    """north_star: every per-term energy and the total within 1e-6 RELATIVE of the fp64 reference (the reference's own
    arithmetic in torch.double on the same fp32-rounded positions and parameters).  The kernels evaluate every term in
    FP64 (energy_pass), so the literal contract holds -- no re-denomination by the magnitude of the cancelling pair
    terms.  atol only covers slots that are exactly zero in the reference ("1-4", disabled terms)."""
    e, e_ref = np.asarray(e), np.asarray(e_ref)
    for k, name in enumerate(_cabi.ENERGY_SLOTS):
        err = np.abs(e[..., k] - e_ref[..., k])
        assert (err <= E_TOL * np.abs(e_ref[..., k]) + 1e-9).all(), \
            f"{name}: {(err / (np.abs(e_ref[..., k]) + 1e-30)).max():.3e} {what}"
    tot, tot_ref = e.sum(-1), e_ref.sum(-1)
    assert (np.abs(tot - tot_ref) <= E_TOL * np.abs(tot_ref) + 1e-9).all(), f"total {what}"


def check(g, pos, f, e, f_ref, e_ref, mode):
    ferr = np.abs(f - f_ref).reshape(len(f), -1).max(1) / np.abs(f_ref).reshape(len(f), -1).max(1)
    assert ferr.max() <= F_TOL, f"force error {ferr.max():.3e}"
    check_energies(e, e_ref, mode)
    return ferr.max()


@pytest.mark.parametrize("system", SYSTEMS)
@pytest.mark.parametrize("mode", list(MODES))
def test_forces_and_energies_vs_reference_fp64(system, mode):
    g = load_golden(system)
    f, e, F = run_gpu(g, g["pos"], **MODES[mode])
    check(g, g["pos"], f, e, g[f"ref_{mode}_f64_forces"], g[f"ref_{mode}_f64_energies"], mode)
    # integer artefacts: Forces.ava_idx equals the reference's (forces.py:346-355)
    np.testing.assert_array_equal(F.ava_idx.cpu().numpy(), g.ava_idx)


def test_known_answer_vector_on_gpu():
    g = load_golden("bb40")
    f, e, _ = run_gpu(g, g["pos"], **MODES["rf"])
    np.testing.assert_allclose(e[0], g["kat_energies"], rtol=3e-6, atol=1e-9)
    np.testing.assert_allclose(f[0], g["kat_forces"], rtol=2e-4, atol=2e-3)


def test_exclusion_mask_bit_exact():
    for system in SYSTEMS:
        g = load_golden(system)
        F = Forces(gpu_par(g))
        mask = F.exclusion_mask(DEV)
        N = g.par.natoms
        allowed = np.zeros((N, 128), dtype=bool)
        allowed[g.ava_idx[:, 0], g.ava_idx[:, 1]] = True
        allowed[g.ava_idx[:, 1], g.ava_idx[:, 0]] = True
        bits = ((mask[:, :, None] >> np.arange(32, dtype=np.uint32)) & 1).reshape(N, 128).astype(bool)
        np.testing.assert_array_equal(bits, ~allowed)


def test_batched_frames_bb40_both_modes():
    g = load_golden("bb40")
    for mode in MODES:
        f, e, _ = run_gpu(g, g["frames"], **MODES[mode])
        check(g, g["frames"], f, e, g[f"frames_ref_{mode}_f64_forces"], g[f"frames_ref_{mode}_f64_energies"], mode)


def test_large_batch_vs_oracle_ala10():
    """4 096 perturbed ALA10 conformers (bench recipe: base frame r mod 101 + 0.05 A noise)."""
    g = load_golden("ala10")
    rng = np.random.default_rng(1234)
    base = g["base_frames"]
    R = 4096
    pos = (base[np.arange(R) % len(base)] + 0.05 * rng.standard_normal((R, 104, 3))).astype(np.float32)
    for mode in MODES:
        f, e, _ = run_gpu(g, pos, **MODES[mode])
        f_ref, e_ref = oracle.forces(g.packed, pos, **MODES[mode])
        check(g, pos, f, e, f_ref, e_ref, mode)


def test_smd_bias_fused():
    g = load_golden("ala10")
    a1, a2, fk, c = g["smd"]
    smd = dict(atom1=int(a1), atom2=int(a2), fk=float(fk), center0=float(c))
    f, e, _ = run_gpu(g, g["pos"], smd=smd)
    check(g, g["pos"], f, e, g["ref_smd_plain_f64_forces"], g["ref_smd_plain_f64_energies"], "plain+smd")
    cen = np.linspace(12, 34, len(g["pos"])).astype(np.float32)
    f, e, _ = run_gpu(g, g["pos"], smd=smd, centers=cen)
    f_ref, e_ref = oracle.forces(g.packed, g["pos"], smd=smd, smd_centers=cen)
    check(g, g["pos"], f, e, f_ref, e_ref, "plain+smd centres")


def test_term_subsets():
    g = load_golden("ala10_ff14sb")
    for terms in (("bonds",), ("angles", "bonds"), ("dihedrals", "1-4", "lj"), ("electrostatics",),
                  ("lj", "impropers"), ("dihedrals", "1-4", "electrostatics", "lj")):
        f, e, _ = run_gpu(g, g["pos"], terms=terms)
        f_ref, e_ref = oracle.forces(g.packed, g["pos"], terms=terms)
        check(g, g["pos"], f, e, f_ref, e_ref, str(terms))


def test_custom_op_registration_matches_the_direct_ctypes_call(monkeypatch):
    """Forces calls the ctypes function directly in eager mode; the same call is registered as the custom op
    b200md::forces_compute for traced graphs.  Both must fill the same bits."""
    g = load_golden("ala10")
    par = Parameters.from_arrays(g.par.to_arrays(), device=DEV)
    F = Forces(par, cutoff=9.0, rfa=True, switch_dist=7.5, external=SMDBias([3], [98], fk=1.0, cent_0=14.0))
    p = torch.tensor(g["pos"][:7].astype(np.float32), device=DEV)
    b = torch.zeros((7, 3, 3), device=DEV)
    f1 = torch.zeros_like(p)
    e1 = F.compute_batch(p, b, f1)
    f2, e2 = torch.zeros_like(p), torch.zeros((7, _cabi.NUM_ENERGIES), dtype=torch.float64, device=DEV)
    status = torch.zeros(7, dtype=torch.int32, device=DEV)
    (smd_on, a1, a2, fk, c0), cen = F._smd_args(7, torch.device(DEV), None)
    torch.ops.b200md.forces_compute(p, b, f2, e2, status, F._handle(torch.device(DEV)), *F._opts(), smd_on, a1, a2, fk, c0, cen)
    assert torch.equal(f1, f2) and torch.equal(e1, e2) and int(status.max()) == 0

    # the classes pick the custom ops when a graph is being compiled: same bits from Integrator.step either way
    def run():
        s = System(104, 7, torch.float, DEV)
        s.pos[:] = p
        F.compute_batch(s.pos, s.box, s.forces)
        integ = Integrator(s, F, timestep=1.0, device=DEV, gamma=1.0, T=300.0, seed=3)
        integ.energy_every = 5
        integ.step(niter=12)
        return s.pos.clone(), s.vel.clone(), integ.observables.clone()

    direct = run()
    monkeypatch.setattr(torch.compiler, "is_compiling", lambda: True)
    via_op = run()
    for a, c in zip(direct, via_op):
        assert torch.equal(a, c)


def test_reference_shaped_compute_api():
    g = load_golden("bb40")
    par = gpu_par(g)
    s = System(40, 1, torch.float, DEV)
    s.set_positions(torch.tensor(g["pos"][0].reshape(40, 3, 1), device=DEV))  # permuted view (Q17)
    F = Forces(par, cutoff=9, rfa=True, switch_dist=7.5)
    pot = F.compute(s.pos, s.box, s.forces, returnDetails=True)
    assert list(pot[0].keys()) == ["electrostatics", "lj", "bonds", "angles", "dihedrals", "1-4", "impropers", "external"]
    assert pot[0]["1-4"].item() == 0.0 and pot[0]["bonds"].shape == (1,)
    np.testing.assert_allclose([v.item() for v in pot[0].values()], g["kat_energies"], rtol=3e-6, atol=1e-9)
    tot = F.compute(s.pos, s.box, s.forces)
    assert abs(float(tot[0]) - g["kat_energies"].sum()) < 1e-3


def test_errors_match_reference():
    g = load_golden("diala")
    par = gpu_par(g)
    with pytest.raises(ValueError, match="not implemented"):
        Forces(par, terms=("bonds", "magic"))
    with pytest.raises(RuntimeError, match="1-4"):
        Forces(par, terms=("1-4", "lj"))
    F = Forces(par)
    pos = torch.tensor(g["pos"], device=DEV)
    pos[2, 5, 1] = float("nan")
    s = System(22, len(pos), torch.float, DEV)
    with pytest.raises(RuntimeError, match="Found NaN coordinates."):
        F.compute(pos, s.box, s.forces)
    with pytest.raises(RuntimeError, match="CUDA"):
        F.compute_batch(pos.cpu(), s.box.cpu(), s.forces.cpu())


def test_determinism_bitwise():
    g = load_golden("ala10")
    pos = np.repeat(g["pos"], 64, axis=0)
    f1, e1, _ = run_gpu(g, pos)
    f2, e2, _ = run_gpu(g, pos)
    np.testing.assert_array_equal(f1, f2)
    np.testing.assert_array_equal(e1, e2)
    # identical replicas give identical results wherever they sit in the batch
    np.testing.assert_array_equal(f1[:8], f1[8:16])


def test_autograd_matches_forces():
    g = load_golden("ala10")
    par = gpu_par(g)
    F = Forces(par)
    pos = torch.tensor(g["pos"][:2], device=DEV, requires_grad=True)
    s = System(104, 2, torch.float, DEV)
    pot = F.compute(pos, s.box, s.forces, returnDetails=True)
    loss = sum(sum(p.values()) for p in pot).sum()
    loss.backward()
    np.testing.assert_allclose(pos.grad.cpu().numpy(), -s.forces.cpu().numpy(), rtol=1e-6, atol=1e-5)
    # non-uniform term weights -> per-term force launches
    pos2 = torch.tensor(g["pos"][:2], device=DEV, requires_grad=True)
    pot = F.compute(pos2, s.box, s.forces, returnDetails=True)
    loss = sum(2.0 * p["bonds"] + 0.5 * p["lj"] + p["electrostatics"] for p in pot).sum()
    loss.backward()
    fb, _ = oracle.forces(g.packed, g["pos"][:2], terms=("bonds",))
    fl, _ = oracle.forces(g.packed, g["pos"][:2], terms=("lj", "dihedrals", "1-4"))
    fd, _ = oracle.forces(g.packed, g["pos"][:2], terms=("dihedrals",))
    fe, _ = oracle.forces(g.packed, g["pos"][:2], terms=("electrostatics", "dihedrals", "1-4"))
    want = -(2.0 * fb + 0.5 * (fl - fd) + (fe - fd))
    got = pos2.grad.cpu().numpy()
    assert np.abs(got - want).max() <= 2e-5 * np.abs(want).max()
    # the same graph edge above 128 atoms (three chains as one 312-atom topology: the large-solute kernels, term subsets
    # without pair terms included)
    gl = load_golden("ala10x3")
    FL = Forces(gpu_par(gl), cutoff=9.0, rfa=True, switch_dist=7.5)
    p3 = torch.tensor(gl["pos"][:2], device=DEV, requires_grad=True)
    s3 = System(312, 2, torch.float, DEV)
    pot = FL.compute(p3, s3.box, s3.forces, returnDetails=True)
    loss = sum(3.0 * p["angles"] + 0.25 * p["lj"] + p["electrostatics"] + p["bonds"] for p in pot).sum()
    loss.backward()
    m = dict(cutoff=9.0, rfa=True, switch_dist=7.5)
    fa, _ = oracle.forces(gl.packed, gl["pos"][:2], terms=("angles",), **m)
    fb, _ = oracle.forces(gl.packed, gl["pos"][:2], terms=("bonds",), **m)
    fl, _ = oracle.forces(gl.packed, gl["pos"][:2], terms=("lj", "dihedrals", "1-4"), **m)
    fd, _ = oracle.forces(gl.packed, gl["pos"][:2], terms=("dihedrals",), **m)
    fe, _ = oracle.forces(gl.packed, gl["pos"][:2], terms=("electrostatics", "dihedrals", "1-4"), **m)
    want = -(3.0 * fa + fb + 0.25 * (fl - fd) + (fe - fd))
    assert np.abs(p3.grad.cpu().numpy() - want).max() <= 2e-5 * np.abs(want).max()


# ---------------------------------------------------------------------------------------------------
# fused integrator
# ---------------------------------------------------------------------------------------------------
def make_system(g, pos, vel):
    par = gpu_par(g)
    R, N, _ = pos.shape
    s = System(N, R, torch.float, DEV)
    s.pos[:] = torch.tensor(pos, device=DEV)
    s.vel[:] = torch.tensor(vel, device=DEV)
    return par, s


@pytest.mark.parametrize("system", ["bb40", "ala10"])
def test_nve_short_horizon_vs_oracle_and_upstream_loop(system):
    g = load_golden(system)
    n = len(g["vv_vel0"])
    pos0 = (g["frames"] if system == "bb40" else g["pos"])[:n]
    par, s = make_system(g, pos0, g["vv_vel0"])
    F = Forces(par)
    F.compute_batch(s.pos, s.box, s.forces)
    integ = Integrator(s, F, timestep=float(g["vv_dt_fs"]), device=DEV)
    integ.energy_every = 1
    nsteps = int(g["vv_nsteps"])
    ekin, pot, T = integ.step(niter=nsteps)
    # vs the restated upstream loop driving the reference forces.py in fp64 (golden)
    assert np.abs(s.pos.cpu().numpy() - g["vv_pos"]).max() < 2e-4
    assert np.abs(s.vel.cpu().numpy() - g["vv_vel"]).max() < 2e-4 * np.abs(g["vv_vel"]).max() * 50
    obs = integ.observables.cpu().numpy()
    epot = obs[:, :, :8].sum(-1)
    np.testing.assert_allclose(epot, g["vv_epot"], rtol=2e-5, atol=2e-3)
    np.testing.assert_allclose(obs[:, :, _cabi.OBS_EKIN], g["vv_ekin"], rtol=2e-4)
    np.testing.assert_allclose(ekin, g["vv_ekin"][-1], rtol=2e-4)
    assert abs(float(pot[0]) - g["vv_epot"][-1, 0]) < 5e-3


def test_langevin_smd_short_horizon_vs_oracle():
    g = load_golden("ala10")
    pos0 = g["pos"][:4]
    vel0 = (0.1 * np.random.default_rng(5).standard_normal(pos0.shape)).astype(np.float32)
    par, s = make_system(g, pos0, vel0)
    rate = 22.0 / 5000
    ext = SMDBias([3], [98], fk=1.0, cent_0=14.0, rate_per_step=rate)
    F = Forces(par, external=ext)
    F.compute_batch(s.pos, s.box, s.forces)
    f0 = s.forces.cpu().numpy().astype(np.float64)
    integ = Integrator(s, F, timestep=1.0, device=DEV, gamma=1.0, T=300.0, seed=42)
    integ.energy_every = 5
    integ.replica0 = 1000
    integ.step(niter=30)
    out = oracle.integrate(g.packed, pos0, vel0, f0, nsteps=30, dt=np.float32(integ.dt), gamma=np.float32(integ.gamma),
                           temperature=300.0, seed=42, replica0=1000, energy_every=5,
                           smd=dict(atom1=3, atom2=98, fk=1.0, center0=14.0, center_rate=np.float32(rate)))
    assert np.abs(s.pos.cpu().numpy() - out["pos"]).max() < 3e-4
    obs = integ.observables.cpu().numpy()
    np.testing.assert_allclose(obs[:, :, _cabi.OBS_E2E_DIST], out["observables"][:, :, _cabi.OBS_E2E_DIST], rtol=1e-5)
    np.testing.assert_allclose(obs[:, :, 7], out["observables"][:, :, 7], rtol=2e-3, atol=1e-4)
    np.testing.assert_allclose(obs[:, :, _cabi.OBS_EKIN], out["observables"][:, :, _cabi.OBS_EKIN], rtol=1e-3)
    np.testing.assert_allclose(obs[:, :, _cabi.OBS_SMD_WORK], out["observables"][:, :, _cabi.OBS_SMD_WORK], rtol=1e-4,
                               atol=1e-7)


def test_smd_work_accumulates_across_calls_and_feeds_jarzynski(tmp_path):
    """W(t) per replica in one launch == the same run split over three step() calls (the state is
    bitwise; the work is an FP64 sum of the same terms regrouped at the call boundaries), the running sum of the
    recorded distances reproduces it, and the ensemble estimate / colvars file come out of it."""
    from molecular_dynamics_b200 import jarzynski_free_energy, read_colvars_traj, write_colvars_traj

    g = load_golden("ala10")
    R = 64
    pos0 = np.repeat(g["pos"][:1], R, 0)
    vel0 = (0.1 * np.random.default_rng(9).standard_normal(pos0.shape)).astype(np.float32)
    rate, fk, c0 = 0.002, 7.2, 13.0

    def run(chunks):
        par, s = make_system(g, pos0, vel0)
        F = Forces(par, external=SMDBias([3], [98], fk=fk, cent_0=c0, rate_per_step=rate))
        F.compute_batch(s.pos, s.box, s.forces)
        integ = Integrator(s, F, timestep=1.0, device=DEV, gamma=1.0, T=300.0, seed=7)
        integ.energy_every = 1
        obs = []
        for n in chunks:
            integ.step(niter=n)
            obs.append(integ.observables.cpu().numpy())
        return np.concatenate(obs, 0), integ

    one, integ = run([60])
    three, integ3 = run([20, 25, 15])
    np.testing.assert_array_equal(one[:, :, _cabi.OBS_E2E_DIST], three[:, :, _cabi.OBS_E2E_DIST])
    np.testing.assert_allclose(one[:, :, _cabi.OBS_SMD_WORK], three[:, :, _cabi.OBS_SMD_WORK], rtol=1e-13, atol=1e-15)
    np.testing.assert_array_equal(integ.smd_work.cpu().numpy(), one[-1, :, _cabi.OBS_SMD_WORK])
    centers = np.float32(c0) + np.float64(np.float32(rate)) * np.arange(1, 61)[:, None]
    centers = centers.astype(np.float32).astype(np.float64)  # the kernel hands the centre over as a float
    expect = np.cumsum(np.float32(fk) * (centers - one[:, :, _cabi.OBS_E2E_DIST]) * np.float64(np.float32(rate)), axis=0)
    np.testing.assert_allclose(one[:, :, _cabi.OBS_SMD_WORK], expect, rtol=1e-5, atol=1e-7)
    assert np.ptp(one[-1, :, _cabi.OBS_SMD_WORK]) > 0  # replicas diverge under the thermostat
    r = jarzynski_free_energy(integ.observables[:, :, _cabi.OBS_SMD_WORK], 300.0)
    assert r["dF"].shape == (60,) and bool((r["dF"] <= r["mean_work"] + 1e-12).all())
    path = str(tmp_path / "smd_out.colvars.traj")
    write_colvars_traj(path, np.arange(1, 61), one[:, 0, _cabi.OBS_E2E_DIST], centers[:, 0], one[:, 0, _cabi.OBS_SMD_WORK])
    np.testing.assert_allclose(read_colvars_traj(path)[3], one[:, 0, _cabi.OBS_SMD_WORK], rtol=1e-13)


def test_integrator_chunking_and_sharding_are_bitwise_invariant():
    """step(20) == step(10)+step(10); replicas [0,8) in one call == two shards with replica0 offsets."""
    g = load_golden("ala10")
    pos0 = g["pos"]
    vel0 = np.zeros_like(pos0)

    def run(chunks, shards):
        outs = []
        for lo, hi in shards:
            par, s = make_system(g, pos0[lo:hi], vel0[lo:hi])
            F = Forces(par)
            F.compute_batch(s.pos, s.box, s.forces)
            integ = Integrator(s, F, timestep=1.0, device=DEV, gamma=1.0, T=300.0, seed=7)
            integ.replica0 = lo
            for c in chunks:
                integ.step(niter=c)
            outs.append(s.pos.cpu().numpy())
        return np.concatenate(outs)

    a = run([20], [(0, 8)])
    b = run([10, 10], [(0, 8)])
    c = run([20], [(0, 3), (3, 8)])
    np.testing.assert_array_equal(a, b)
    np.testing.assert_array_equal(a, c)


def test_snapshots_and_reports_layout():
    g = load_golden("diala")
    pos0 = g["pos"]
    par, s = make_system(g, pos0, np.zeros_like(pos0))
    F = Forces(par)
    F.compute_batch(s.pos, s.box, s.forces)
    integ = Integrator(s, F, timestep=0.5, device=DEV)
    integ.energy_every, integ.snapshot_every = 4, 8
    integ.step(niter=16)
    assert integ.observables.shape == (4, 4, _cabi.NUM_OBSERVABLES) and integ.snapshots.shape == (2, 4, 22, 3)
    np.testing.assert_array_equal(integ.snapshots[-1].cpu().numpy(), s.pos.cpu().numpy())
    f0, _ = oracle.forces(g.packed, pos0)
    out = oracle.integrate(g.packed, pos0, np.zeros_like(pos0), f0, nsteps=16, dt=np.float32(integ.dt), energy_every=4,
                           snapshot_every=8)
    assert np.abs(integ.snapshots.cpu().numpy() - out["snapshots"]).max() < 1e-4
    np.testing.assert_allclose(integ.observables.cpu().numpy()[:, :, :8].sum(-1), out["observables"][:, :, :8].sum(-1),
                               rtol=1e-5, atol=1e-3)


def test_nve_energy_drift_100k_steps_matches_fp64_oracle():
    """BASELINE.json configs[2] at its own size: 8 192 replicas x 100 000 fused NVE steps (dt = 1 fs, bench conformers)
    on the GPU, and the FP64 oracle over the SAME 100 000 steps on the first 16 of those replicas.  Trajectories diverge
    chaotically after ~1e3 steps, so "matches" is a statement about ensemble statistics: the drift of E_pot + E_kin per
    replica (least-squares slope x run length) and the RMS fluctuation about the fit.  Tolerance = three standard
    errors of the difference of the two ensemble means (sigma from the ensembles themselves), not a fixed number."""
    import bench
    import sys as _sys
    _sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    from nve_drift import compare_ensembles, drift_stats

    g = load_golden("ala10")
    R, RO, NSTEPS, EVERY = 8192, 16, 100000, 200
    pos, vel = bench.synth_inputs(bench.load_ala10(), R)
    par, s = make_system(g, pos, vel)
    F = Forces(par)
    F.compute_batch(s.pos, s.box, s.forces)
    f0 = s.forces[:RO].cpu().numpy().astype(np.float64)
    integ = Integrator(s, F, timestep=1.0, device=DEV)
    integ.energy_every = EVERY
    chunks = []
    for _ in range(5):
        integ.step(niter=NSTEPS // 5)
        o = integ.observables
        chunks.append((o[:, :, :8].sum(-1) + o[:, :, _cabi.OBS_EKIN]).cpu().numpy())
    etot = np.concatenate(chunks, 0)
    assert etot.shape == (NSTEPS // EVERY, R) and np.isfinite(etot).all()
    out = oracle.integrate(g.packed, pos[:RO], vel[:RO], f0, nsteps=NSTEPS, dt=np.float32(integ.dt), energy_every=EVERY)
    eo = out["observables"][:, :, :8].sum(-1) + out["observables"][:, :, _cabi.OBS_EKIN]
    assert np.abs(etot[0, :RO] - eo[0]).max() < 0.05  # still the same trajectories after 200 steps
    cmp = compare_ensembles(drift_stats(etot), drift_stats(eo))
    assert abs(cmp["drift_diff"]) < 3.0 * cmp["drift_diff_se"], cmp
    assert abs(cmp["rms_diff"]) < 3.0 * cmp["rms_diff_se"], cmp
    # and the size of it: the ensemble drifts by well under 1e-3 of E_tot (~550 kcal/mol) over 100 ps
    assert abs(cmp["gpu_drift_mean"]) < 0.5 and cmp["gpu_rms_mean"] < 1.0, cmp


@pytest.mark.parametrize("system", ["bb40", "ala10"])
@pytest.mark.parametrize("case", ["rep_plain", "rep_cut", "cg_plain", "cg_rf", "combo_plain", "combo_cut"])
def test_repulsion_terms_vs_reference_fp64(system, case):
    """SURVEY 8 row a12 (forces.py:416-448): CUDA path against the unmodified reference run in fp64
    (tests/golden/repulsion.npz, tools/make_golden_repulsion.py)."""
    from test_oracle import REPULSION_CASES

    g = load_golden(system)
    with np.load(os.path.join(os.path.dirname(__file__), "golden", "repulsion.npz")) as z:
        r = {k: z[k] for k in z.files if k.startswith(system)}
    terms, kw = REPULSION_CASES[case]
    F = Forces(gpu_par(g), terms=terms, **kw)
    pos = torch.tensor(r[f"{system}_pos"], device=DEV)
    s = System(pos.shape[1], pos.shape[0], torch.float, DEV)
    pot = F.compute(pos, s.box, s.forces, returnDetails=True)
    fr = r[f"{system}_{case}_f64_forces"]
    f = s.forces.cpu().numpy()
    assert (np.abs(f - fr).reshape(len(f), -1).max(1) / np.abs(fr).reshape(len(f), -1).max(1)).max() <= 1e-5
    assert set(pot[0]) == set(terms) | {"external"}
    for t in terms:
        got = np.array([float(p[t]) for p in pot])
        np.testing.assert_allclose(got, r[f"{system}_{case}_f64_E_{t}"], rtol=1e-6, atol=1e-9)


@pytest.mark.parametrize("system", ["bb40", "ala10"])
@pytest.mark.parametrize("mode", sorted(MODES))
def test_minimum_image_vs_reference_fp64(system, mode):
    """SURVEY 8 row a3 (wrap_dist, forces.py:358-363): per-replica orthorhombic boxes smaller than the molecule (one
    replica of the batch in vacuum) against the unmodified reference run in fp64 (tests/golden/pbc.npz)."""
    g = load_golden(system)
    with np.load(os.path.join(os.path.dirname(__file__), "golden", "pbc.npz")) as z:
        r = {k: z[k] for k in z.files if k.startswith(system)}
    F = Forces(gpu_par(g), **MODES[mode])
    pos = torch.tensor(r[f"{system}_pos"], device=DEV)
    box = torch.tensor(r[f"{system}_box"], device=DEV)
    frc = torch.zeros_like(pos)
    ene = F.compute_batch(pos, box, frc).cpu().numpy()
    fr, er = r[f"{system}_{mode}_f64_forces"], r[f"{system}_{mode}_f64_energies"]
    f = frc.cpu().numpy()
    assert (np.abs(f - fr).reshape(len(f), -1).max(1) / np.abs(fr).reshape(len(f), -1).max(1)).max() <= 1e-5
    check_energies(ene, er, f"minimum image, {system} {mode}")


def test_integrator_with_box_short_horizon_vs_oracle():
    g = load_golden("ala10")
    pos0 = g["pos"][:4]
    vel0 = (0.1 * np.random.default_rng(11).standard_normal(pos0.shape)).astype(np.float32)
    par, s = make_system(g, pos0, vel0)
    box = np.zeros((4, 3, 3), dtype=np.float32)
    for r, b in enumerate(([27.0, 24.0, 22.0], [0.0, 0.0, 0.0], [23.0, 30.0, 25.0], [25.0, 25.0, 25.0])):
        box[r, 0, 0], box[r, 1, 1], box[r, 2, 2] = b
    s.box[:] = torch.tensor(box, device=DEV)
    F = Forces(par, cutoff=9.0, rfa=True, switch_dist=7.5)
    F.compute_batch(s.pos, s.box, s.forces)
    f0 = s.forces.cpu().numpy().astype(np.float64)
    integ = Integrator(s, F, timestep=1.0, device=DEV)
    integ.energy_every = 10
    integ.step(niter=30)
    out = oracle.integrate(g.packed, pos0, vel0, f0, nsteps=30, dt=np.float32(integ.dt), energy_every=10, box=box,
                           cutoff=9.0, rfa=True, switch_dist=7.5)
    assert np.abs(s.pos.cpu().numpy() - out["pos"]).max() < 3e-4
    obs = integ.observables.cpu().numpy()
    np.testing.assert_allclose(obs[:, :, :8].sum(-1), out["observables"][:, :, :8].sum(-1), rtol=2e-5, atol=2e-3)


def test_checkpoint_resume_is_bitwise(tmp_path):
    """a Langevin + SMD run resumed from a checkpoint continues bit-identically (counter-based RNG keyed by
    the global step, deterministic reductions)."""
    from molecular_dynamics_b200 import load_checkpoint, save_checkpoint

    g = load_golden("ala10")
    pos0 = g["pos"][:4]
    vel0 = (0.1 * np.random.default_rng(3).standard_normal(pos0.shape)).astype(np.float32)

    def fresh():
        par, s = make_system(g, pos0, vel0)
        F = Forces(par, external=SMDBias([3], [98], fk=1.0, cent_0=14.0, rate_per_step=1e-3))
        F.compute_batch(s.pos, s.box, s.forces)
        return s, Integrator(s, F, timestep=1.0, device=DEV, gamma=1.0, T=300.0, seed=7)

    s1, i1 = fresh()
    i1.step(niter=60)
    s2, i2 = fresh()
    i2.step(niter=25)
    save_checkpoint(str(tmp_path / "ck.pt"), s2, i2)
    s3, i3 = fresh()
    load_checkpoint(str(tmp_path / "ck.pt"), s3, i3)
    assert i3.step_count == 25
    i3.step(niter=35)
    assert torch.equal(s1.pos, s3.pos) and torch.equal(s1.vel, s3.vel) and torch.equal(s1.forces, s3.forces)


def test_minimizers_relax_perturbed_conformers():
    """upstream's minimize_bfgs call on one replica, and the batched FIRE minimiser on 32."""
    import bench
    from molecular_dynamics_b200 import minimize_bfgs, minimize_fire

    g = load_golden("ala10")
    pos, vel = bench.synth_inputs(bench.load_ala10(), 32)
    par, s = make_system(g, pos, vel)
    F = Forces(par)
    e0 = F.compute_batch(s.pos, s.box, s.forces).sum(1).cpu().numpy()
    out = minimize_fire(s, F, steps=400, fmax=5.0)
    e1 = out["epot"].cpu().numpy()
    assert (e1 < e0 - 50.0).all(), (e0, e1)            # the 0.05 A perturbation alone stores ~100 kcal/mol in the bonds
    assert float(out["fmax"].max()) < 60.0             # from several hundred kcal/mol/A
    assert not torch.isnan(s.pos).any()
    par1, s1 = make_system(g, pos[:1], vel[:1])
    F1 = Forces(par1)
    e0 = float(F1.compute_batch(s1.pos, s1.box, s1.forces).sum())
    minimize_bfgs(s1, F1, fmax=0.5, steps=200)
    e1 = float(F1.compute_batch(s1.pos, s1.box, s1.forces).sum())
    assert e1 < e0 - 100.0 and float(s1.forces.abs().max()) < 20.0
    with pytest.raises(RuntimeError):
        minimize_bfgs(s, F)


def test_tutorial_loop_shapes():
    """The tutorial's production loop (tutorial.ipynb cell 18) runs unchanged: Epot + Ekin, Wrapper, LogWriter."""
    from molecular_dynamics_b200 import Wrapper

    g = load_golden("bb40")
    par, s = make_system(g, g["frames"][:1], np.zeros_like(g["frames"][:1]))
    F = Forces(par, cutoff=9.0, rfa=True, switch_dist=7.5)
    F.compute(s.pos, s.box, s.forces)
    integ = Integrator(s, F, timestep=1.0, device=DEV, gamma=0.1, T=300.0)
    wrapper = Wrapper(par.natoms if hasattr(par, "natoms") else len(par.masses), par.bonds, DEV)
    Ekin, Epot, T = integ.step(niter=10)
    wrapper.wrap(s.pos, s.box)
    etot = Epot + Ekin
    assert isinstance(etot, np.ndarray) and etot.shape == (1,) and np.isfinite(etot).all()
    assert len(Epot) == 1 and float(Epot[0]) == float(np.asarray(Epot)[0])


class _Chamber:
    """ALA10 with every CHAMBER term family switched on: 22 harmonic CHARMM impropers (general torsion
    path, periodicity 0) and 65 Urey-Bradley 1-3 terms as extra bond rows."""

    def __init__(self):
        from molecular_dynamics_b200.parameters import parameters_from_prmtop
        from molecular_dynamics_b200.topology_io import read_pdb_coords

        gold = os.path.join(os.path.dirname(__file__), "golden")
        self.par = parameters_from_prmtop(os.path.join(gold, "full_da-1.3.prmtop"), precision=torch.float64,
                                          charmm_impropers=True, urey_bradley=True)
        base = read_pdb_coords(os.path.join(gold, "smd_ini.pdb")).astype(np.float64)
        rng = np.random.default_rng(21)
        self.pos = (base[None] + 0.04 * rng.standard_normal((6,) + base.shape)).astype(np.float32)
        self.pos[0] = base
        self.ava_idx = _cabi.make_pair_index(self.par.natoms, self.par.get_exclusions())
        self.packed = _cabi.PackedTopology(self.par, self.ava_idx)


@pytest.mark.parametrize("mode", sorted(MODES))
def test_chamber_impropers_and_urey_bradley_vs_oracle(mode):
    g = _Chamber()
    f_ref, e_ref = oracle.forces(g.packed, g.pos.astype(np.float64), **MODES[mode])
    f, e, F = run_gpu(g, g.pos, **MODES[mode])
    check(g, g.pos, f, e, f_ref, e_ref, mode)
    assert e[0, _cabi.energy_slot("impropers")] > 1.0  # the 22 impropers are evaluated, not dropped
    np.testing.assert_array_equal(F.ava_idx.cpu().numpy(), g.ava_idx)


def test_chamber_fused_integrator_short_horizon_vs_oracle():
    g = _Chamber()
    pos0 = g.pos[:3]
    vel0 = (0.05 * np.random.default_rng(2).standard_normal(pos0.shape)).astype(np.float32)
    par, s = make_system(g, pos0, vel0)
    F = Forces(par)
    F.compute_batch(s.pos, s.box, s.forces)
    f0 = s.forces.cpu().numpy().astype(np.float64)
    integ = Integrator(s, F, timestep=1.0, device=DEV)
    integ.energy_every = 10
    integ.step(niter=40)
    out = oracle.integrate(g.packed, pos0, vel0, f0, nsteps=40, dt=np.float32(integ.dt), energy_every=10)
    assert np.abs(s.pos.cpu().numpy() - out["pos"]).max() < 3e-4
    obs = integ.observables.cpu().numpy()
    np.testing.assert_allclose(obs[:, :, :8].sum(-1), out["observables"][:, :, :8].sum(-1), rtol=2e-5, atol=2e-3)
    k = _cabi.energy_slot("impropers")
    np.testing.assert_allclose(obs[:, :, k], out["observables"][:, :, k], rtol=2e-3, atol=1e-3)


class _Restart:
    """ALA10 from the CHAMBER prmtop as the reference loads it, started from the reference's own NAMD restart of
    replica 3 (rigidBonds ALL: all 53 X-H bonds at their lengths, 300 K velocities)."""

    def __init__(self, R=4):
        from molecular_dynamics_b200.parameters import parameters_from_prmtop
        from molecular_dynamics_b200.topology_io import read_namd_bin

        gold = os.path.join(os.path.dirname(__file__), "golden")
        self.par = parameters_from_prmtop(os.path.join(gold, "full_da-1.3.prmtop"), precision=torch.float64)
        x = read_namd_bin(os.path.join(gold, "smd_out.restart.coor"))
        v = read_namd_bin(os.path.join(gold, "smd_out.restart.vel"))
        rng = np.random.default_rng(4)
        self.pos = np.repeat(x[None], R, 0).astype(np.float32)
        self.vel = (v[None] * (1.0 + 0.2 * rng.standard_normal((R, 1, 1)))).astype(np.float32)
        self.vel[0] = v
        self.ava_idx = _cabi.make_pair_index(self.par.natoms, self.par.get_exclusions())
        self.packed = _cabi.PackedTopology(self.par, self.ava_idx)
        b, m = self.par.bonds.numpy(), self.par.masses.numpy().ravel()
        xh = (m[b[:, 0]] < 1.5) | (m[b[:, 1]] < 1.5)
        self.xh = b[xh]
        self.req = self.par.bond_params.numpy()[xh, 1].astype(np.float32).astype(np.float64)

    def bond_error(self, pos):
        """max |r - req| over the rigid bonds, in units of the fp32 spacing of the largest coordinate (the restart is
        not centred: coordinates reach 40 A, where a float resolves 3.8e-6 A)."""
        p = np.asarray(pos, dtype=np.float64)
        err = np.abs(np.linalg.norm(p[:, self.xh[:, 0]] - p[:, self.xh[:, 1]], axis=-1) - self.req).max()
        return err / float(np.spacing(np.float32(np.abs(p).max())))


@pytest.mark.parametrize("mode", sorted(MODES))
def test_rigid_bonds_short_horizon_vs_oracle(mode):
    """SURVEY 8f N4 (constraints): SHAKE / RATTLE in the fused kernel against the FP64 oracle, 2 fs steps."""
    g = _Restart()
    par, s = make_system(g, g.pos, g.vel)
    F = Forces(par, **MODES[mode])
    F.compute_batch(s.pos, s.box, s.forces)
    f0 = s.forces.cpu().numpy().astype(np.float64)
    integ = Integrator(s, F, timestep=2.0, device=DEV, rigid_bonds=True)
    integ.energy_every = 10
    ekin, pot, T = integ.step(niter=40)
    assert integ.num_constraints == 53
    out = oracle.integrate(g.packed, g.pos, g.vel, f0, nsteps=40, dt=np.float32(integ.dt), energy_every=10,
                           rigid_bonds=True, **MODES[mode])
    assert np.abs(s.pos.cpu().numpy() - out["pos"]).max() < 5e-4
    assert np.abs(s.vel.cpu().numpy() - out["vel"]).max() < 5e-3 * np.abs(out["vel"]).max()
    assert g.bond_error(s.pos.cpu().numpy()) < 3.0  # a few spacings of the stored fp32 coordinates
    obs = integ.observables.cpu().numpy()
    np.testing.assert_allclose(obs[:, :, _cabi.OBS_EKIN], out["observables"][:, :, _cabi.OBS_EKIN], rtol=1e-3)
    np.testing.assert_allclose(obs[:, :, :8].sum(-1), out["observables"][:, :, :8].sum(-1), rtol=2e-5, atol=5e-3)
    # temperature over 3 N - 53 degrees of freedom
    np.testing.assert_allclose(T, 2.0 * ekin / ((3 * 104 - 53) * 0.001987191), rtol=1e-6)
    # velocities carry no component along the rigid bonds
    v, x = s.vel.cpu().numpy().astype(np.float64), s.pos.cpu().numpy().astype(np.float64)
    rv = ((x[:, g.xh[:, 0]] - x[:, g.xh[:, 1]]) * (v[:, g.xh[:, 0]] - v[:, g.xh[:, 1]])).sum(-1)
    assert np.abs(rv).max() < 2e-5


def test_rigid_bonds_2fs_langevin_and_energy_conservation():
    g = _Restart(R=256)
    # NVE, 2 fs, 5 000 steps = 10 ps: bonds held, total energy without drift beyond the integrator's own fluctuation
    par, s = make_system(g, g.pos, g.vel)
    F = Forces(par)
    F.compute_batch(s.pos, s.box, s.forces)
    integ = Integrator(s, F, timestep=2.0, device=DEV, rigid_bonds=True)
    integ.energy_every = 50
    integ.step(niter=5000)
    assert g.bond_error(s.pos.cpu().numpy()) < 3.0
    obs = integ.observables.cpu().numpy()
    etot = obs[:, :, :8].sum(-1) + obs[:, :, _cabi.OBS_EKIN]
    drift = np.abs(etot[-20:].mean(0) - etot[:20].mean(0))
    assert np.median(drift) < 0.5 and drift.max() < 2.0, (np.median(drift), drift.max())
    # Langevin at 300 K, 2 fs: the thermostat temperature is reached over 3 N - 53 degrees of freedom
    par, s = make_system(g, g.pos, g.vel)
    F = Forces(par)
    F.compute_batch(s.pos, s.box, s.forces)
    integ = Integrator(s, F, timestep=2.0, device=DEV, gamma=5.0, T=300.0, seed=3, rigid_bonds=True)
    integ.energy_every = 50
    integ.step(niter=1000)
    temps = []
    for _ in range(4):
        temps.append(integ.step(niter=500)[2])
    assert abs(np.mean(temps) - 300.0) < 8.0, np.mean(temps)
    assert g.bond_error(s.pos.cpu().numpy()) < 3.0


def test_rigid_bonds_is_a_noop_without_hydrogens_and_refuses_unsupported_groups():
    g = load_golden("bb40")
    pos0 = g["frames"][:4]
    vel0 = (0.05 * np.random.default_rng(1).standard_normal(pos0.shape)).astype(np.float32)
    res = []
    for rigid in (False, True):
        par, s = make_system(g, pos0, vel0)
        F = Forces(par)
        F.compute_batch(s.pos, s.box, s.forces)
        integ = Integrator(s, F, timestep=1.0, device=DEV, rigid_bonds=rigid)
        integ.step(niter=20)
        res.append(s.pos.cpu().numpy())
    np.testing.assert_array_equal(res[0], res[1])
    # a hydrogen hanging on a hydrogen-mass atom is not a group the kernel holds: refused, not mis-integrated
    g = _Restart(R=2)
    m = g.par.masses.numpy().ravel()
    parent = int(g.xh[0, 0] if m[g.xh[0, 0]] > 1.5 else g.xh[0, 1])
    g.par.masses[parent] = 1.008
    par, s = make_system(g, g.pos, g.vel)
    F = Forces(par)
    F.compute_batch(s.pos, s.box, s.forces)
    integ = Integrator(s, F, timestep=2.0, device=DEV, rigid_bonds=True)
    with pytest.raises(NotImplementedError, match="rigid_bonds"):
        integ.step(niter=2)


def _wrap_angle(a):
    return (a + np.pi) % (2 * np.pi) - np.pi


@pytest.mark.parametrize("system", ["ala10", "bb40", "ala10x3"])
def test_dihedral_angles_and_backward_vs_oracle(system):
    """Energy.calc_phi batched (cgan.py:166-207) and its analytic backward (ala10x3: 312 atoms, the
    kernels for solutes above 128 atoms)."""
    g = load_golden(system)
    pos = (g["frames"] if system == "bb40" else g["pos"])[:12].astype(np.float32)
    par = gpu_par(g)
    F = Forces(par)
    p = torch.tensor(pos, device=DEV, requires_grad=True)
    phi = F.dihedral_angles(p)
    ref = oracle.dihedral_angles(g.packed, pos.astype(np.float64))
    assert phi.shape == ref.shape == (len(pos), len(g.par.dihedrals))
    assert np.abs(_wrap_angle(phi.detach().cpu().numpy() - ref)).max() < 2e-5
    # backward of L = sum w phi against central differences of the FP64 oracle
    rng = np.random.default_rng(0)
    w = rng.standard_normal(ref.shape)
    (phi * torch.tensor(w, device=DEV, dtype=torch.float32)).sum().backward()
    grad = p.grad.cpu().numpy().astype(np.float64)
    h = 1e-5
    for _ in range(40):
        r, a, c = rng.integers(len(pos)), rng.integers(pos.shape[1]), rng.integers(3)
        pp, pm = pos.astype(np.float64), pos.astype(np.float64)
        pp[r, a, c] += h
        pm[r, a, c] -= h
        d = _wrap_angle(oracle.dihedral_angles(g.packed, pp[r:r + 1]) - oracle.dihedral_angles(g.packed, pm[r:r + 1]))
        fd = (w[r] * d[0]).sum() / (2 * h)
        assert abs(grad[r, a, c] - fd) < 2e-4 * max(1.0, np.abs(grad[r]).max()), (r, a, c, grad[r, a, c], fd)
    # minimum image: shifting atoms by box vectors leaves the angles unchanged
    box = np.zeros((len(pos), 3, 3), dtype=np.float32)
    box[:, 0, 0], box[:, 1, 1], box[:, 2, 2] = 60.0, 70.0, 80.0
    shifted = pos.copy()
    shifted[:, ::3, 0] += 60.0
    shifted[:, 1::4, 2] -= 80.0
    phib = F.dihedral_angles(torch.tensor(shifted, device=DEV), torch.tensor(box, device=DEV))
    assert np.abs(_wrap_angle(phib.cpu().numpy() - ref)).max() < 1e-4


def test_batched_energy_adapter_matches_per_sample_and_backpropagates():
    """The ML callers' loss (cgan.py:345-360): sum over the batch of (sum of terms + 200 * E2End) / B, here from two
    launches forward and two backward instead of a rebuilt Energy object per sample."""
    from molecular_dynamics_b200 import Energy

    g = load_golden("ala10")
    pos = g["pos"][:16].astype(np.float32)
    B = len(pos)
    colvar = {"name": "E2End Harm", "fk": 1.0, "cent_0": 12.0, "cent_1": 34.0, "T": 1.0, "group1": [3], "group2": [98]}
    en = Energy(parameters=gpu_par(g), colvar=colvar, device=DEV)
    times = np.linspace(0.0, 1.0, B).astype(np.float32)
    x = torch.tensor(pos, device=DEV, requires_grad=True)
    lib = _cabi.load_library()
    n0 = lib.b200md_launch_count()
    ene, phis = en.calc_energy_batch(x, torch.tensor(times))
    assert lib.b200md_launch_count() - n0 == 3  # forces, FP64 energies, dihedral angles
    total = sum(v for k, v in ene.items() if k != "E2End Harm").sum() / B + 200.0 * ene["E2End Harm"].sum() / B
    n0 = lib.b200md_launch_count()
    total.backward()
    assert lib.b200md_launch_count() - n0 == 1  # {restraint} alone: columns with equal weights form a group, the largest group comes from the forward's total force
    # oracle: the adapter's settings (cutoff 9, reaction field, switch 7.5), per-sample centres
    mode = dict(cutoff=9, rfa=True, switch_dist=7.5)
    cen = (12.0 + 22.0 * times).astype(np.float32)
    smd = dict(atom1=3, atom2=98, fk=1.0, center0=0.0)
    f_all, e_all = oracle.forces(g.packed, pos, smd=smd, smd_centers=cen, **mode)
    f_terms, e_terms = oracle.forces(g.packed, pos, **mode)
    want = e_terms.sum() / B + 200.0 * e_all[:, 7].sum() / B
    assert abs(float(total.detach()) - want) < 2e-5 * abs(want)
    grad_ref = -(f_terms + 200.0 * (f_all - f_terms)) / B
    err = np.abs(x.grad.cpu().numpy() - grad_ref).max() / np.abs(grad_ref).max()
    assert err < 2e-5, err
    np.testing.assert_allclose(ene["E2End Harm"].detach().cpu().numpy(), e_all[:, 7], rtol=2e-5, atol=1e-6)
    # per-sample adapter call: same numbers, reference-shaped return value
    pot, phi0 = en.calc_energy(pos[B - 1].reshape(104, 3, 1), float(times[B - 1]))
    assert set(pot[0]) == set(ene) and phi0.shape == (len(g.par.dihedrals),)
    for k in ene:
        np.testing.assert_allclose(float(pot[0][k].detach()), float(ene[k][B - 1].detach()), rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(phi0.cpu().numpy(), phis[B - 1].detach().cpu().numpy(), atol=1e-6)
    ref_phi = oracle.dihedral_angles(g.packed, pos.astype(np.float64))
    assert np.abs(_wrap_angle(phis.detach().cpu().numpy() - ref_phi)).max() < 2e-5


class _RandomMolecule:
    """Seeded random topology for the packing paths no fixed fixture reaches: atom counts around the 32-lane block
    edges and at the 128-atom limit, branched bonds, several terms per dihedral with periodicities on both sides of
    the Fourier-series limit (6) and non-integer ones, harmonic (periodicity 0) slices, zero-k dihedrals that reduce
    to their 1-4 pair, random LJ tables.  The oracle restates forces.py for arbitrary parameter objects, so it is the
    reference here."""

    def __init__(self, natoms, seed, R=3, long_bonds=()):
        rng = np.random.default_rng(seed)
        N = natoms
        # start structure: a helix (no collinear triples, which make torsion gradients singular), jittered
        t = np.arange(N) * rng.uniform(0.9, 1.1)
        a_h, b_h = rng.uniform(1.5, 1.8), rng.uniform(0.8, 1.0)
        base = np.stack([a_h * np.cos(t), a_h * np.sin(t), b_h * t], 1)
        self.pos = (base[None] + 0.08 * rng.standard_normal((R, N, 3))).astype(np.float32)
        ntypes = int(rng.integers(1, 7))
        sig, eps = rng.uniform(1.5, 3.4, ntypes), rng.uniform(0.02, 0.25, ntypes)
        s = 0.5 * (sig[:, None] + sig[None, :])
        e = np.sqrt(eps[:, None] * eps[None, :])
        arrs = {"mapped_atom_types": rng.integers(0, ntypes, N), "A": 4 * e * s ** 12, "B": 4 * e * s ** 6,
                "charges": rng.uniform(-0.6, 0.6, N), "masses": rng.choice([1.008, 12.011, 14.007, 15.999], N).reshape(N, 1)}
        bonds = [(i, i + 1) for i in range(N - 1)]
        for _ in range(N // 6):  # branches / rings between lattice neighbours further along the snake
            i = int(rng.integers(0, N - 3))
            j = int(min(N - 1, i + rng.integers(2, 6)))
            if (i, j) not in bonds:
                bonds.append((i, j))
        bonds = np.array(sorted(set(bonds) | set(long_bonds)))  # long_bonds: cross-links far along the chain (disulfide-like)
        dist0 = np.linalg.norm(base[bonds[:, 0]] - base[bonds[:, 1]], axis=1)
        arrs["bonds"] = bonds
        arrs["bond_params"] = np.stack([rng.uniform(100, 450, len(bonds)), dist0 + rng.uniform(-0.1, 0.1, len(bonds))], 1)
        for lb in long_bonds:  # hundreds of A long on this helix: a soft spring, so that fp32 rounding of the length (3e-5 A)
            arrs["bond_params"][(bonds == np.array(lb)).all(1), 0] = 2.0  # stays far below the force tolerance
        if N >= 3:
            ang = np.array([(i, i + 1, i + 2) for i in range(N - 2)])
            arrs["angles"] = ang
            arrs["angle_params"] = np.stack([rng.uniform(20, 90, len(ang)), rng.uniform(1.5, 2.3, len(ang))], 1)
        if N >= 4:
            dih = np.array([(i, i + 1, i + 2, i + 3) for i in range(N - 3)])
            nd = len(dih)
            arrs["dihedrals"] = dih
            nsl = int(rng.integers(1, 4))
            arrs["dihedral_nslices"] = np.array(nsl)
            for j in range(nsl):
                idx = np.sort(rng.choice(nd, size=max(1, nd // (j + 1)), replace=False)) if j else np.arange(nd)
                per = rng.choice([1, 2, 3, 4, 6], len(idx)).astype(np.float64)
                k = rng.uniform(0.0, 2.5, len(idx))
                if j == 0:
                    k[rng.random(len(idx)) < 0.25] = 0.0  # zero-k dihedrals: only their 1-4 pair survives
                if j == 1:
                    per[rng.random(len(idx)) < 0.3] = rng.choice([7.0, 8.0, 2.5])  # beyond the series / non-integer
                if j == 2 and seed % 2:
                    per[:] = 0.0  # a CHARMM harmonic slice (forces.py:570-576)
                    k *= 10.0
                arrs[f"dihedral_idx_{j}"] = idx
                arrs[f"dihedral_params_{j}"] = np.stack([k, rng.uniform(-np.pi, np.pi, len(idx)), per], 1)
            ex = np.zeros((N, N), bool)
            ex[bonds[:, 0], bonds[:, 1]] = ex[bonds[:, 1], bonds[:, 0]] = True
            ex[ang[:, 0], ang[:, 2]] = ex[ang[:, 2], ang[:, 0]] = True
            keep = ~ex[dih[:, 0], dih[:, 3]]
            if keep.any():
                i14 = dih[keep][:, [0, 3]]
                arrs["idx14"] = i14
                arrs["nonbonded_14_params"] = np.stack([rng.uniform(1e2, 5e3, len(i14)), rng.uniform(1, 50, len(i14)),
                                                        rng.choice([1.0, 2.0], len(i14)), rng.choice([1.0, 1.2], len(i14))], 1)
            nimp = N // 5
            if nimp:
                c = rng.integers(1, N - 2, nimp)
                imp = np.stack([c, c - 1, c + 1, np.minimum(c + 2, N - 1)], 1)
                arrs["impropers"] = imp
                arrs["improper_nslices"] = np.array(1)
                arrs["improper_idx_0"] = np.arange(nimp)
                harmonic = seed % 3 == 0
                arrs["improper_params_0"] = np.stack([rng.uniform(1, 40, nimp), rng.uniform(-0.4, 0.4, nimp) if harmonic
                                                      else rng.choice([0.0, np.pi], nimp),
                                                      np.zeros(nimp) if harmonic else np.full(nimp, 2.0)], 1)
        self.par = Parameters.from_arrays(arrs, precision=torch.float64)
        self.ava_idx = _cabi.make_pair_index(N, self.par.get_exclusions())
        self.packed = _cabi.PackedTopology(self.par, self.ava_idx)


@pytest.mark.parametrize("natoms,seed", [(2, 20), (3, 21), (4, 0), (9, 1), (31, 2), (32, 3), (33, 4), (57, 5), (64, 6), (65, 7), (96, 8), (97, 9),
                                         (104, 10), (127, 11), (128, 12), (128, 13)])
def test_random_topologies_vs_oracle(natoms, seed):
    g = _RandomMolecule(natoms, seed)
    for mode in ("plain", "rf"):
        f_ref, e_ref = oracle.forces(g.packed, g.pos, **MODES[mode])
        f, e, F = run_gpu(g, g.pos, **MODES[mode])
        ferr = np.abs(f - f_ref).reshape(len(f), -1).max(1) / np.abs(f_ref).reshape(len(f), -1).max(1)
        assert ferr.max() < 3e-5, (mode, ferr.max())
        check_energies(e, e_ref, f"random topology {natoms} atoms, {mode}")
        np.testing.assert_array_equal(F.ava_idx.cpu().numpy().reshape(-1, 2), g.ava_idx.reshape(-1, 2))
    # the fused integrator on the same topology (fixed-point bonded accumulation, static schedule, bank spreading)
    vel0 = (0.02 * np.random.default_rng(seed).standard_normal(g.pos.shape)).astype(np.float32)
    par, s = make_system(g, g.pos, vel0)
    F = Forces(par)
    F.compute_batch(s.pos, s.box, s.forces)
    f0 = s.forces.cpu().numpy().astype(np.float64)
    integ = Integrator(s, F, timestep=0.5, device=DEV)
    integ.energy_every = 5
    integ.step(niter=10)
    out = oracle.integrate(g.packed, g.pos, vel0, f0, nsteps=10, dt=np.float32(integ.dt), energy_every=5)
    assert np.abs(s.pos.cpu().numpy() - out["pos"]).max() < 2e-4
    obs = integ.observables.cpu().numpy()
    ref = out["observables"]
    sc = np.abs(ref[:, :, :8]).sum(-1) + 1.0
    assert (np.abs(obs[:, :, :8].sum(-1) - ref[:, :, :8].sum(-1)) <= 2e-4 * sc).all()


def test_apply_constraints_projects_a_perturbed_start_like_the_oracle():
    g = _Restart(R=8)
    rng = np.random.default_rng(6)
    m = g.par.masses.numpy().ravel()
    hyd = m < 1.5
    bad = g.pos.copy()
    bad[:, hyd] += (0.05 * rng.standard_normal(bad[:, hyd].shape)).astype(np.float32)  # hydrogens off their bond lengths
    ek0 = 0.5 * (m[None, :, None] * g.vel.astype(np.float64) ** 2).sum((1, 2))

    def first_step(project):
        par, s = make_system(g, bad, g.vel)
        F = Forces(par)
        integ = Integrator(s, F, timestep=2.0, device=DEV, rigid_bonds=True)
        if project:
            integ.apply_constraints()
        state = s.pos.cpu().numpy(), s.vel.cpu().numpy()
        F.compute_batch(s.pos, s.box, s.forces)
        ekin, _, _ = integ.step(niter=1)
        return state, ekin, integ, s

    assert g.bond_error(bad) > 1000.0
    (p, v), ekin, integ, s = first_step(True)
    assert g.bond_error(p) < 3.0
    po, vo = oracle.constrain(g.packed, bad, g.vel)
    assert np.abs(p - po).max() < 2e-5 and np.abs(v - vo).max() < 2e-5 * max(1.0, np.abs(vo).max())
    # the first rigid step of an unprojected start does the same projection itself (SHAKE, and RATTLE takes the
    # bond-parallel velocity correction out again): both starts continue with the same kinetic energy
    _, ekin_unprojected, _, _ = first_step(False)
    assert np.abs(ekin - ek0).max() < 0.2 * ek0.mean(), (ekin - ek0)
    assert np.abs(ekin - ekin_unprojected).max() < 0.03 * ek0.mean()
    # idempotent
    before = s.pos.clone()
    integ.apply_constraints()
    assert float((s.pos - before).abs().max()) < 2e-5


def test_dihedral_angles_kernel_against_the_reference_data_sets_phi_psi():
    """The batched calc_phi kernel against the phi / psi angles the reference's data set stores beside its backbone
    frames (tests/golden/bb40_phi_psi.npy; the oracle is pinned on the same fixture in tests/test_oracle.py)."""
    from test_oracle import _phi_psi_fixture

    g, frames, rows, want = _phi_psi_fixture()
    F = Forces(gpu_par(g))
    phi = F.dihedral_angles(torch.tensor(frames, device=DEV, dtype=torch.float32))
    got = np.degrees(phi.cpu().numpy().astype(np.float64))[:, rows]
    diff = (got - want + 180.0) % 360.0 - 180.0
    assert np.abs(diff).max() < 0.02, np.abs(diff).max()


def test_chamber_energies_on_gpu_against_namd_at_the_restart_frame():
    """The CUDA path end to end (prmtop loader -> Forces -> kernel) against NAMD's own energies of the reference's
    restart frame (see tests/test_oracle.py::test_chamber_prmtop_route_pinned_by_namd_energies_at_the_restart_frame)."""
    import json

    from molecular_dynamics_b200.parameters import parameters_from_prmtop
    from molecular_dynamics_b200.topology_io import read_namd_bin

    gold = os.path.join(os.path.dirname(__file__), "golden")
    namd = json.load(open(os.path.join(gold, "namd_step0_energies.json")))["restart_step_500000"]["energies"]
    x = torch.tensor(read_namd_bin(os.path.join(gold, "smd_out.restart.coor"))[None], dtype=torch.float32, device=DEV)
    slot = _cabi.energy_slot
    res = {}
    for name, kw in (("plain", {}), ("full", dict(charmm_impropers=True, urey_bradley=True))):
        par = parameters_from_prmtop(os.path.join(gold, "full_da-1.3.prmtop"), device=DEV, **kw)
        s = System(104, 1, torch.float, DEV)
        res[name] = Forces(par).compute_batch(x, s.box, s.forces).cpu().numpy()[0]
    ub = res["full"][slot("bonds")] - res["plain"][slot("bonds")]
    assert abs(res["plain"][slot("bonds")] - namd["BOND"]) < 2e-3
    assert abs(res["full"][slot("angles")] + ub - namd["ANGLE"]) < 2e-3
    assert abs(res["full"][slot("dihedrals")] - namd["DIHED"]) < 2e-3
    assert abs(res["full"][slot("impropers")] - namd["IMPRP"]) < 2e-3


# ---- solutes above 128 atoms: the large-solute kernels (csrc/b200md_large.cuh) --------------------------------------
@pytest.mark.parametrize("natoms,seed", [(129, 30), (160, 31), (257, 32), (300, 33), (700, 35), (1000, 34)])
def test_large_solute_forces_vs_oracle(natoms, seed):
    """SURVEY 8f N4 (larger solutes, e.g. the adk system named at V_Calculations/Test-10_diALA/potcalc.py:114): above 128
    atoms the same Forces.compute runs as bonded-window + half-shell pair-rectangle kernels; same tolerances as the on-chip path."""
    g = _RandomMolecule(natoms, seed, R=2)
    for mode in ("plain", "rf"):
        f_ref, e_ref = oracle.forces(g.packed, g.pos, **MODES[mode])
        f, e, F = run_gpu(g, g.pos, **MODES[mode])
        ferr = np.abs(f - f_ref).reshape(len(f), -1).max(1) / np.abs(f_ref).reshape(len(f), -1).max(1)
        assert ferr.max() < 1e-5, (mode, ferr.max())
        check_energies(e, e_ref, f"large random topology {natoms} atoms, {mode}")
        # every kind runs in FP64: the slots agree with the oracle far below the contract
        for slot in range(8):
            np.testing.assert_allclose(e[:, slot], e_ref[:, slot], rtol=1e-9, atol=1e-9)
        np.testing.assert_array_equal(F.ava_idx.cpu().numpy().reshape(-1, 2), g.ava_idx.reshape(-1, 2))
    # integer artefact: the exclusion bit matrix is the complement of the reference's pair list
    m = F.exclusion_mask(DEV)
    assert m.shape == (natoms, (natoms + 31) // 32)
    bits = ((m[:, :, None] >> np.arange(32, dtype=np.uint32)) & 1).reshape(natoms, -1)[:, :natoms].astype(bool)
    allowed = np.zeros((natoms, natoms), bool)
    allowed[g.ava_idx[:, 0], g.ava_idx[:, 1]] = allowed[g.ava_idx[:, 1], g.ava_idx[:, 0]] = True
    np.testing.assert_array_equal(~bits, allowed)


def test_large_solute_cross_links_far_along_the_chain_vs_oracle():
    """Bonded windows stage the positions their items reference in shared memory when those span at most 320 atoms; a
    cross-link far along the chain (here atoms 5-650 and 100-400 of 700) widens the span of its windows beyond that, and they
    read positions from HBM instead (GlobalPos in large_bonded_forces_kernel).  Same results either way; in cutoff mode the
    cross-link bonds are longer than the cutoff and drop out (forces.py:117-125)."""
    g = _RandomMolecule(700, 36, R=2, long_bonds=((5, 650), (100, 400)))
    for mode in ("plain", "rf"):
        f_ref, e_ref = oracle.forces(g.packed, g.pos, **MODES[mode])
        f, e, _ = run_gpu(g, g.pos, **MODES[mode])
        ferr = np.abs(f - f_ref).reshape(len(f), -1).max(1) / np.abs(f_ref).reshape(len(f), -1).max(1)
        assert ferr.max() < 1e-5, (mode, ferr.max())
        check_energies(e, e_ref, f"large solute with cross-links, {mode}")
    assert np.abs(f_ref[:, 5]).max() > 0 and np.abs(f[:, 650] - f_ref[:, 650]).max() < 1e-5 * np.abs(f_ref).max()


def test_large_solute_term_subsets_box_and_smd_vs_oracle():
    g = _RandomMolecule(200, 40, R=3)
    # per-replica boxes (one replica in vacuum), cutoff / reaction field / switch
    box = np.zeros((3, 3, 3), dtype=np.float32)
    for r, b in enumerate(([31.0, 28.0, 40.0], [0.0, 0.0, 0.0], [26.0, 33.0, 37.0])):
        box[r, 0, 0], box[r, 1, 1], box[r, 2, 2] = b
    par = gpu_par(g)
    p = torch.tensor(g.pos, device=DEV)
    s = System(200, 3, torch.float, DEV)
    s.box[:] = torch.tensor(box, device=DEV)
    F = Forces(par, **MODES["rf"])
    e = F.compute_batch(p, s.box, s.forces).cpu().numpy()
    f_ref, e_ref = oracle.forces(g.packed, g.pos, box=box.reshape(3, 9), **MODES["rf"])
    f = s.forces.cpu().numpy()
    assert np.abs(f - f_ref).max() / np.abs(f_ref).max() < 1e-5
    check_energies(e, e_ref, "large solute, per-replica boxes")
    # term subsets + the SMD restraint with per-replica centres
    smd = dict(atom1=3, atom2=190, fk=1.5, center0=0.0)
    cen = np.array([150.0, 160.0, 170.0], dtype=np.float32)
    for terms in (("bonds", "angles"), ("lj",), ("electrostatics", "dihedrals", "1-4"), ("repulsion", "impropers")):
        f_ref, e_ref = oracle.forces(g.packed, g.pos, smd=smd, smd_centers=cen, terms=terms)
        f, e, _ = run_gpu(g, g.pos, smd=smd, centers=cen, terms=terms)
        assert np.abs(f - f_ref).max() / np.abs(f_ref).max() < 1e-5, terms
        check_energies(e, e_ref, f"large solute {terms}")


def test_large_solute_integrator_vs_oracle_and_bitwise_determinism():
    g = _RandomMolecule(300, 41, R=3)
    vel0 = (0.02 * np.random.default_rng(41).standard_normal(g.pos.shape)).astype(np.float32)
    rate = 1e-3

    def run(replica0=500, sl=slice(None)):
        par, s = make_system(g, g.pos[sl], vel0[sl])
        F = Forces(par, external=SMDBias([3], [290], fk=1.0, cent_0=240.0, rate_per_step=rate))
        F.compute_batch(s.pos, s.box, s.forces)
        f0 = s.forces.cpu().numpy().astype(np.float64)
        integ = Integrator(s, F, timestep=0.5, device=DEV, gamma=1.0, T=300.0, seed=9)
        integ.energy_every = 5
        integ.snapshot_every = 10
        integ.replica0 = replica0
        integ.step(niter=20)
        return s, integ, f0

    s, integ, f0 = run()
    out = oracle.integrate(g.packed, g.pos, vel0, f0, nsteps=20, dt=np.float32(integ.dt), gamma=np.float32(integ.gamma),
                           temperature=300.0, seed=9, replica0=500, energy_every=5, snapshot_every=10,
                           smd=dict(atom1=3, atom2=290, fk=1.0, center0=240.0, center_rate=np.float32(rate)))
    assert np.abs(s.pos.cpu().numpy() - out["pos"]).max() < 3e-4
    assert np.abs(s.vel.cpu().numpy() - out["vel"]).max() < 3e-4 * max(1.0, np.abs(out["vel"]).max())
    obs, ref = integ.observables.cpu().numpy(), out["observables"]
    sc = np.abs(ref[:, :, :8]).sum(-1) + 1.0
    assert (np.abs(obs[:, :, :8].sum(-1) - ref[:, :, :8].sum(-1)) <= 2e-4 * sc).all()
    np.testing.assert_allclose(obs[:, :, _cabi.OBS_EKIN], ref[:, :, _cabi.OBS_EKIN], rtol=1e-3)
    np.testing.assert_allclose(obs[:, :, _cabi.OBS_E2E_DIST], ref[:, :, _cabi.OBS_E2E_DIST], rtol=1e-5)
    np.testing.assert_allclose(obs[:, :, _cabi.OBS_SMD_WORK], ref[:, :, _cabi.OBS_SMD_WORK], rtol=1e-4, atol=1e-7)
    assert np.abs(integ.snapshots.cpu().numpy() - out["snapshots"]).max() < 3e-4
    # same run again, and replica 2 alone as its own shard: bit-identical (private force slots and partial rows, fixed-order
    # reductions, RNG keyed by the global replica index)
    s2, integ2, _ = run()
    assert torch.equal(s.pos, s2.pos) and torch.equal(s.vel, s2.vel) and torch.equal(integ.observables, integ2.observables)
    s3, integ3, _ = run(replica0=502, sl=slice(2, 3))
    assert torch.equal(s.pos[2:3], s3.pos) and torch.equal(integ.observables[:, 2:3], integ3.observables)


@pytest.mark.parametrize("mode", ["plain", "rf"])
def test_large_solute_persistent_steps_kernel_is_bitwise_the_per_step_launches(mode, monkeypatch):
    """Few replicas of a large solute: runs of steps between report steps are ONE cooperative launch (large_steps_kernel:
    forces, grid barrier, half-kicks, grid barrier per step); with many replicas every step is 3 launches.  Same unit code,
    same summation orders: every bit of the trajectory, the records and the snapshots agrees -- Langevin noise, a moving
    SMD restraint, snapshot steps (which end a run) and a periodic replica included."""
    g = _RandomMolecule(400, 43, R=4)
    vel0 = (0.02 * np.random.default_rng(43).standard_normal(g.pos.shape)).astype(np.float32)
    out = {}
    for flag in ("0", "1"):
        monkeypatch.setenv("B200MD_LARGE_PERSISTENT", flag)
        par, s = make_system(g, g.pos, vel0)
        s.box[1, 0, 0], s.box[1, 1, 1], s.box[1, 2, 2] = 60.0, 60.0, 900.0
        F = Forces(par, external=SMDBias([3], [390], fk=1.0, cent_0=300.0, rate_per_step=1e-3), **MODES[mode])
        F.compute_batch(s.pos, s.box, s.forces)
        integ = Integrator(s, F, timestep=0.5, device=DEV, gamma=1.0, T=300.0, seed=5)
        integ.energy_every, integ.snapshot_every = 6, 7
        lib = _cabi.load_library()
        n0 = lib.b200md_launch_count()
        integ.step(niter=23)
        out[flag] = (s.pos.clone(), s.vel.clone(), s.forces.clone(), integ.observables.clone(), integ.snapshots.clone(),
                     lib.b200md_launch_count() - n0)
    for a, b in zip(out["0"][:5], out["1"][:5]):
        assert torch.isfinite(b).all() and torch.equal(a, b)
    assert out["1"][5] < out["0"][5] / 2  # 23 steps: 4 report steps + 7 runs (+ first half-kicks) against 3 launches a step


def test_large_solute_more_replicas_than_grid_rows():
    """The large-solute grids carry the replica in blockIdx.y (at most 65 535 rows): beyond that a CTA loops over replicas.
    66 000 replicas of a 129-atom molecule, copies of two conformers: every copy must carry the bits of its original,
    through Forces.compute and a few NVE steps (per-step launches: the units of a step exceed the resident grid)."""
    g = _RandomMolecule(129, 44, R=2)
    R = 66000
    par = gpu_par(g)
    idx = torch.arange(R, device=DEV) % 2
    s = System(129, R, torch.float, DEV)
    s.pos[:] = torch.tensor(g.pos, device=DEV)[idx]
    s.vel[:] = torch.tensor((0.01 * np.random.default_rng(44).standard_normal(g.pos.shape)).astype(np.float32), device=DEV)[idx]
    F = Forces(par, **MODES["rf"])
    e = F.compute_batch(s.pos, s.box, s.forces)
    f_ref, e_ref = oracle.forces(g.packed, g.pos, **MODES["rf"])
    assert np.abs(s.forces[:2].cpu().numpy() - f_ref).max() / np.abs(f_ref).max() < 1e-5
    check_energies(e[:2].cpu().numpy(), e_ref, "large solute, 66 000 replicas")
    for t in (s.forces, e):
        assert torch.equal(t, t[:2][idx])
    integ = Integrator(s, F, timestep=1.0, device=DEV)
    integ.energy_every = 2
    integ.step(niter=3)
    for t in (s.pos, s.vel, s.forces):
        assert torch.isfinite(t).all() and torch.equal(t, t[:2][idx])
    assert torch.equal(integ.observables, integ.observables[:, :2][:, idx])


def test_large_solute_minimise_then_equilibrate():
    """The reference's own use of a large solute (potcalc.py / tutorial.ipynb on a prmtop): minimise, draw
    Maxwell-Boltzmann velocities, run Langevin dynamics -- on a 3-chain deca-alanine bundle (312 atoms) behind the same
    classes: FIRE lowers the energy of perturbed conformers, the thermostat then holds ~300 K."""
    from molecular_dynamics_b200 import maxwell_boltzmann, minimize_fire

    g = load_golden("ala10")
    copies, R = 3, 8
    parL = Parameters.from_arrays(g.par.tiled(copies).to_arrays(), device=DEV)
    shift = np.repeat(np.arange(copies), 104)[None, :, None] * np.array([0.0, 12.0, 12.0])
    rng = np.random.default_rng(8)
    pos = (g["base_frames"][:R * copies].reshape(R, copies * 104, 3) + shift + 0.05 * rng.standard_normal((R, copies * 104, 3))).astype(np.float32)
    s = System(copies * 104, R, torch.float, DEV)
    s.pos[:] = torch.tensor(pos, device=DEV)
    F = Forces(parL, cutoff=9.0, rfa=True, switch_dist=7.5)
    e0 = F.compute_batch(s.pos, s.box, s.forces).sum(1).cpu().numpy()
    out = minimize_fire(s, F, steps=300, fmax=5.0)
    e1 = out["epot"].cpu().numpy()
    assert (e1 < e0 - 150.0).all(), (e0, e1)
    assert not torch.isnan(s.pos).any()
    s.set_velocities(maxwell_boltzmann(parL.masses, T=300.0, replicas=R))
    F.compute_batch(s.pos, s.box, s.forces)
    integ = Integrator(s, F, timestep=1.0, device=DEV, gamma=10.0, T=300.0, seed=12)  # 0.1 ps coupling: from a minimum, half
    integ.energy_every = 50                                                            # of the drawn kinetic energy goes into
    integ.step(niter=1000)                                                             # the potential within a few periods
    ek = integ.observables[-8:, :, _cabi.OBS_EKIN].mean().item()
    temp = 2.0 * ek / (3 * copies * 104 * 0.001987191)  # Ekin = 3/2 N kB T, kB in kcal/mol/K
    assert 270.0 < temp < 330.0, temp  # (300.5 K measured: deterministic seeds)
    assert torch.isfinite(s.pos).all()


def test_large_solute_energy_adapter_forward_backward():
    """The batched Energy adapter (cgan.py:137-244, 345-360) on a solute above 128 atoms: per-term energies, the restraint
    with one centre per sample, torsion angles, and the gradient of the ML callers' loss, against the oracle."""
    from molecular_dynamics_b200 import Energy

    g = load_golden("ala10x3")
    pos = g["pos"].astype(np.float32)
    B = len(pos)
    colvar = {"name": "E2End Harm", "fk": 1.0, "cent_0": 12.0, "cent_1": 34.0, "T": 1.0, "group1": [3], "group2": [306]}
    en = Energy(parameters=gpu_par(g), colvar=colvar, device=DEV)
    times = np.linspace(0.0, 1.0, B).astype(np.float32)
    x = torch.tensor(pos, device=DEV, requires_grad=True)
    ene, phis = en.calc_energy_batch(x, torch.tensor(times))
    total = sum(v for k, v in ene.items() if k != "E2End Harm").sum() / B + 200.0 * ene["E2End Harm"].sum() / B \
        + (phis * phis).sum() * 1e-3
    total.backward()
    mode = dict(cutoff=9, rfa=True, switch_dist=7.5)
    cen = (12.0 + 22.0 * times).astype(np.float32)
    smd = dict(atom1=3, atom2=306, fk=1.0, center0=0.0)
    f_all, e_all = oracle.forces(g.packed, pos, smd=smd, smd_centers=cen, **mode)
    f_terms, e_terms = oracle.forces(g.packed, pos, **mode)
    ref_phi = oracle.dihedral_angles(g.packed, pos.astype(np.float64))
    assert np.abs(_wrap_angle(phis.detach().cpu().numpy() - ref_phi)).max() < 2e-5
    want = e_terms.sum() / B + 200.0 * e_all[:, 7].sum() / B + (ref_phi ** 2).sum() * 1e-3
    assert abs(float(total.detach()) - want) < 2e-5 * abs(want)
    # energy part of the gradient from the oracle's forces; the angle part from the kernel's own backward, checked against
    # the oracle's central differences in test_dihedral_angles_and_backward_vs_oracle
    x2 = torch.tensor(pos, device=DEV, requires_grad=True)
    ph2 = en.forces.dihedral_angles(x2)
    ((ph2 * ph2).sum() * 1e-3).backward()
    grad_ref = -(f_terms + 200.0 * (f_all - f_terms)) / B + x2.grad.cpu().numpy()
    err = np.abs(x.grad.cpu().numpy() - grad_ref).max() / np.abs(grad_ref).max()
    assert err < 2e-5, err


@pytest.mark.parametrize("mode", sorted(MODES))
def test_large_solute_rigid_bonds_vs_oracle(mode):
    """SHAKE / RATTLE above 128 atoms (one thread per hydrogen group, from HBM): three copies of the reference's restart
    frame as one 312-atom topology, 2 fs steps, against the FP64 oracle; apply_constraints on a perturbed start."""
    g = _Restart(R=3)
    copies = 3
    parL = g.par.tiled(copies)
    N = 104 * copies
    shift = np.repeat(np.arange(copies), 104)[None, :, None] * np.array([0.0, 15.0, 15.0])
    pos = (np.tile(g.pos, (1, copies, 1)) + shift).astype(np.float32)
    vel = np.tile(g.vel, (1, copies, 1)).astype(np.float32)
    packed = _cabi.PackedTopology(parL, _cabi.make_pair_index(N, parL.get_exclusions()))
    xh = np.concatenate([g.xh + 104 * k for k in range(copies)])
    req = np.tile(g.req, copies)

    def bond_error(p):
        p = np.asarray(p, dtype=np.float64)
        err = np.abs(np.linalg.norm(p[:, xh[:, 0]] - p[:, xh[:, 1]], axis=-1) - req).max()
        return err / float(np.spacing(np.float32(np.abs(p).max())))

    parG = Parameters.from_arrays(parL.to_arrays(), device=DEV)
    s = System(N, 3, torch.float, DEV)
    s.pos[:] = torch.tensor(pos, device=DEV)
    s.vel[:] = torch.tensor(vel, device=DEV)
    F = Forces(parG, **MODES[mode])
    F.compute_batch(s.pos, s.box, s.forces)
    f0 = s.forces.cpu().numpy().astype(np.float64)
    integ = Integrator(s, F, timestep=2.0, device=DEV, rigid_bonds=True)
    integ.energy_every = 10
    ekin, pot, T = integ.step(niter=40)
    assert integ.num_constraints == 53 * copies
    out = oracle.integrate(packed, pos, vel, f0, nsteps=40, dt=np.float32(integ.dt), energy_every=10, rigid_bonds=True, **MODES[mode])
    assert np.abs(s.pos.cpu().numpy() - out["pos"]).max() < 5e-4
    assert np.abs(s.vel.cpu().numpy() - out["vel"]).max() < 5e-3 * np.abs(out["vel"]).max()
    assert bond_error(s.pos.cpu().numpy()) < 3.0
    obs = integ.observables.cpu().numpy()
    np.testing.assert_allclose(obs[:, :, _cabi.OBS_EKIN], out["observables"][:, :, _cabi.OBS_EKIN], rtol=1e-3)
    np.testing.assert_allclose(obs[:, :, :8].sum(-1), out["observables"][:, :, :8].sum(-1), rtol=2e-5, atol=1.5e-2)
    np.testing.assert_allclose(T, 2.0 * ekin / ((3 * N - 53 * copies) * 0.001987191), rtol=1e-6)
    v, x = s.vel.cpu().numpy().astype(np.float64), s.pos.cpu().numpy().astype(np.float64)
    rv = ((x[:, xh[:, 0]] - x[:, xh[:, 1]]) * (v[:, xh[:, 0]] - v[:, xh[:, 1]])).sum(-1)
    assert np.abs(rv).max() < 2e-5
    if mode == "plain":  # the projection of a perturbed start
        rng = np.random.default_rng(6)
        hyd = parL.masses.numpy().ravel() < 1.5
        bad = pos.copy()
        bad[:, hyd] += (0.05 * rng.standard_normal(bad[:, hyd].shape)).astype(np.float32)
        s.pos[:] = torch.tensor(bad, device=DEV)
        s.vel[:] = torch.tensor(vel, device=DEV)
        assert bond_error(bad) > 1000.0
        integ.apply_constraints()
        po, vo = oracle.constrain(packed, bad, vel)
        assert bond_error(s.pos.cpu().numpy()) < 3.0
        assert np.abs(s.pos.cpu().numpy() - po).max() < 2e-5
        assert np.abs(s.vel.cpu().numpy() - vo).max() < 2e-5 * max(1.0, np.abs(vo).max())


def test_large_solute_unsupported_options_raise():
    g = _RandomMolecule(140, 42, R=1)
    par, s = make_system(g, g.pos, np.zeros_like(g.pos))
    F = Forces(par)
    F.compute_batch(s.pos, s.box, s.forces)
    # rigid bonds above 128 atoms: same refusal of groups the solvers cannot hold (this random molecule hangs hydrogen-mass
    # atoms on each other)
    m = g.par.masses.numpy().ravel()
    if ((m < 3.5).sum() > 0):
        try:
            Integrator(s, F, timestep=1.0, device=DEV, rigid_bonds=True).step(niter=1)
        except NotImplementedError as e:
            assert "rigid_bonds" in str(e)


def test_large_solute_bundle_of_ala10_chains_vs_oracle():
    """A compact large solute from real conformers: 6 deca-alanine chains 14 A apart as one 624-atom topology
    (Parameters.tiled), plain and cutoff / reaction field / switch (where whole j tiles beyond the cutoff are dropped by
    their bounding boxes, which must not change a bit of the result), against the oracle."""
    g = load_golden("ala10")
    copies, R = 6, 3
    parL = g.par.tiled(copies)
    N = 104 * copies
    shift = np.repeat(np.arange(copies), 104)[None, :, None] * np.array([0.0, 10.0, 10.0])  # closest inter-chain pair > 5 A
    frames = g["base_frames"][:R * copies].reshape(R, copies * 104, 3)
    pos = (frames + shift).astype(np.float32)
    idx = _cabi.make_pair_index(N, parL.get_exclusions())
    packed = _cabi.PackedTopology(parL, idx)
    parG = Parameters.from_arrays(parL.to_arrays(), device=DEV)
    for mode in ("plain", "rf"):
        F = Forces(parG, **MODES[mode])
        s = System(N, R, torch.float, DEV)
        e = F.compute_batch(torch.tensor(pos, device=DEV), s.box, s.forces).cpu().numpy()
        f = s.forces.cpu().numpy()
        f_ref, e_ref = oracle.forces(packed, pos, **MODES[mode])
        assert np.abs(f - f_ref).max() / np.abs(f_ref).max() < 1e-5, mode
        np.testing.assert_allclose(e, e_ref, rtol=1e-9, atol=1e-9)  # FP64 energies, pair terms included
    # the same conformers far apart in a box: minimum image brings them back into range, no tile may be dropped
    box = np.zeros((R, 3, 3), dtype=np.float32)
    box[:, 0, 0], box[:, 1, 1], box[:, 2, 2] = 80.0, 120.0, 120.0
    posb = pos.copy()
    posb[:, 104:208, 0] += 160.0  # chain 1 shifted by two box lengths (box y, z = 80: the 50 A bundle stays clear of its images)
    F = Forces(parG, **MODES["rf"])
    s = System(N, R, torch.float, DEV)
    s.box[:] = torch.tensor(box, device=DEV)
    F.compute_batch(torch.tensor(posb, device=DEV), s.box, s.forces)
    f_ref, _ = oracle.forces(packed, posb, box=box.reshape(R, 9), **MODES["rf"])
    assert np.abs(s.forces.cpu().numpy() - f_ref).max() / np.abs(f_ref).max() < 1e-5
    # a box that makes the first and the last chain neighbours THROUGH the boundary (60 A in y, z: chain 5 at (50, 50) is
    # 10 A from chain 0's image), long in x: the periodic tile-box test (minimum image of the box centres) keeps those
    # tile pairs and drops the far ones; one replica stays in vacuum
    box[:, 0, 0], box[:, 1, 1], box[:, 2, 2] = 200.0, 60.0, 60.0
    box[1] = 0.0
    s.box[:] = torch.tensor(box, device=DEV)
    e = F.compute_batch(torch.tensor(pos, device=DEV), s.box, s.forces).cpu().numpy()
    f_ref, e_ref = oracle.forces(packed, pos, box=box.reshape(R, 9), **MODES["rf"])
    assert np.abs(s.forces.cpu().numpy() - f_ref).max() / np.abs(f_ref).max() < 1e-5
    np.testing.assert_allclose(e, e_ref, rtol=1e-9, atol=1e-9)
    f_vac, _ = oracle.forces(packed, pos[:1], **MODES["rf"])
    assert np.abs(f_ref[0] - f_vac[0]).max() > 1e-3  # the boundary neighbours do interact


def test_large_solute_benchmark_bundle_far_apart_chains_vs_oracle():
    """The topology bench.py and tools/profile_large.py time: 10 deca-alanine chains 40 A apart as one 1 040-atom topology
    (33 tiles -> 4 atom blocks of 9, 9, 9, 6 tiles, 10 pair rectangles), here on a 4 x 3 grid around the origin (on the
    bench's line the coordinates reach 390 A, where the fp32 state alone costs 1.5e-5 A per step).  In cutoff mode some
    rectangles lie wholly beyond the cutoff and write zero rows without staging anything; tiles that straddle two chains
    have 40 A boxes.  Forces, energies, and a few NVE steps of the integrator (fused half-kicks, partial rows summed in the
    half-kick kernel) against the oracle."""
    g = load_golden("ala10")
    copies, R = 10, 2
    parL = g.par.tiled(copies)
    N = 104 * copies
    k = np.repeat(np.arange(copies), 104)
    shift = np.stack([40.0 * (k % 4) - 60.0, 40.0 * (k // 4) - 40.0, np.zeros(N)], 1)[None]
    pos = g["base_frames"][:R * copies].reshape(R, N, 3) + shift
    pos = (pos - pos.mean(1, keepdims=True) + shift.mean(1, keepdims=True)).astype(np.float32)
    packed = _cabi.PackedTopology(parL, _cabi.make_pair_index(N, parL.get_exclusions()))
    parG = Parameters.from_arrays(parL.to_arrays(), device=DEV)
    for mode in ("plain", "rf"):
        F = Forces(parG, **MODES[mode])
        s = System(N, R, torch.float, DEV)
        s.pos[:] = torch.tensor(pos, device=DEV)
        e = F.compute_batch(s.pos, s.box, s.forces).cpu().numpy()
        f_ref, e_ref = oracle.forces(packed, pos, **MODES[mode])
        assert np.abs(s.forces.cpu().numpy() - f_ref).max() / np.abs(f_ref).max() < 1e-5, mode
        np.testing.assert_allclose(e, e_ref, rtol=1e-9, atol=1e-9)
        f_only = torch.zeros_like(s.forces)
        F.compute_batch(s.pos, s.box, f_only, energies=False)
        assert torch.equal(f_only, s.forces)  # forces do not depend on whether energies are recorded
        # 12 NVE steps, energies every 5 (report steps run the FP64 kernels beside the force launches)
        vel0 = (0.01 * np.random.default_rng(3).standard_normal(pos.shape)).astype(np.float32)
        s.vel[:] = torch.tensor(vel0, device=DEV)
        integ = Integrator(s, F, timestep=1.0, device=DEV)
        integ.energy_every = 5
        integ.step(niter=12)
        o = oracle.integrate(packed, pos, vel0, f_ref, nsteps=12, dt=np.float32(integ.dt), energy_every=5, **MODES[mode])
        # fp32 state at coordinates up to ~80 A (half an ulp = 4e-6 A per step, times bond constants of ~1e3 in the forces)
        assert np.abs(s.pos.cpu().numpy() - o["pos"]).max() < 1e-4, mode
        assert np.abs(s.vel.cpu().numpy() - o["vel"]).max() < 1e-3 * max(1.0, np.abs(o["vel"]).max()), mode
        obs, ref = integ.observables.cpu().numpy(), o["observables"]
        assert obs.shape[0] == 3  # steps 5, 10 and the last one
        sc = np.abs(ref[:, :, :8]).sum(-1) + 1.0
        assert (np.abs(obs[:, :, :8].sum(-1) - ref[:, :, :8].sum(-1)) <= 1e-5 * sc).all(), mode
        np.testing.assert_allclose(obs[:, :, _cabi.OBS_EKIN], ref[:, :, _cabi.OBS_EKIN], rtol=1e-4)


def test_large_solute_results_do_not_depend_on_the_batch():
    """The launch geometry per replica (i-blocks, bonded CTAs, order of the energy partials) is a function of the topology
    only: a replica's forces and energies are bit-identical whatever batch it is evaluated in."""
    g = _RandomMolecule(1000, 50, R=2)
    par = gpu_par(g)
    F = Forces(par)
    rng = np.random.default_rng(50)
    pos = (g.pos[rng.integers(0, 2, 48)] + 0.01 * rng.standard_normal((48, 1000, 3))).astype(np.float32)
    p = torch.tensor(pos, device=DEV)
    s = System(1000, 48, torch.float, DEV)
    e = F.compute_batch(p, s.box, s.forces)
    for sl in (slice(0, 1), slice(17, 18), slice(40, 43)):
        n = sl.stop - sl.start
        s1 = System(1000, n, torch.float, DEV)
        e1 = F.compute_batch(p[sl].contiguous(), s1.box, s1.forces)
        assert torch.equal(s1.forces, s.forces[sl]) and torch.equal(e1, e[sl])
    f_ref, e_ref = oracle.forces(g.packed, pos[:2])
    assert np.abs(s.forces[:2].cpu().numpy() - f_ref).max() / np.abs(f_ref).max() < 1e-5


@pytest.mark.parametrize("mode", list(MODES))
def test_large_solute_vs_reference_fp64(mode):
    """The large-solute kernels against the unmodified reference itself (tests/golden/ala10x3.npz, 312 atoms, fp64 run of
    TorchMD_Fix/forces.py on the same fp32 positions): same tolerances as the 104-atom fixtures."""
    g = load_golden("ala10x3")
    f, e, F = run_gpu(g, g["pos"], **MODES[mode])
    check(g, g["pos"], f, e, g[f"ref_{mode}_f64_forces"], g[f"ref_{mode}_f64_energies"], mode)
    np.testing.assert_array_equal(F.ava_idx.cpu().numpy(), g.ava_idx)


# ---- round 2: time slicing, force-only launches, boundary gaps ---------------------------------------------------
def _slicing_run(g, R, niter, slice_len, rigid=False):
    rng = np.random.default_rng(11)
    pos0 = (g["pos"][rng.integers(0, len(g["pos"]), R)] + 0.02 * rng.standard_normal((R, 104, 3))).astype(np.float32)
    vel0 = (0.01 * rng.standard_normal(pos0.shape)).astype(np.float32)
    par, s = make_system(g, pos0, vel0)
    F = Forces(par, external=SMDBias([3], [98], fk=1.0, cent_0=14.0, rate_per_step=1e-3))
    F.compute_batch(s.pos, s.box, s.forces)
    integ = Integrator(s, F, timestep=1.0, device=DEV, gamma=1.0, T=300.0, seed=5, rigid_bonds=rigid)
    integ.energy_every, integ.snapshot_every = 6, 9
    if rigid:
        integ.apply_constraints()
    old = os.environ.get("B200MD_SLICE")
    try:
        if slice_len is None:
            os.environ.pop("B200MD_SLICE", None)
        else:
            os.environ["B200MD_SLICE"] = str(slice_len)
        integ.step(niter=niter)
    finally:
        if old is None:
            os.environ.pop("B200MD_SLICE", None)
        else:
            os.environ["B200MD_SLICE"] = old
    torch.cuda.synchronize()
    return [t.cpu().numpy() for t in (s.pos, s.vel, s.forces, integ.observables, integ.snapshots)]


@pytest.mark.parametrize("R,rigid", [(40, False), (700, False), (40, True)])
def test_time_slicing_is_bitwise_invariant(R, rigid):
    """The fused integrator hands out (slice of steps, replica) units from a ticket counter and parks a replica's state
    in HBM between its slices (csrc/b200md.cu, integrate_kernel).  A step's arithmetic does not depend on the slicing:
    positions, velocities, forces, every observable record (SMD work carried across slices included) and the snapshots
    are bit-identical for one slice per launch, the default, and slices of 1 and 7 steps -- also with more replicas
    than resident CTAs and with the rigid-bond kernel."""
    g = load_golden("ala10")
    niter = 37
    ref = _slicing_run(g, R, niter, niter, rigid)
    assert np.isfinite(ref[0]).all()
    for sl in (None, 1, 7):
        out = _slicing_run(g, R, niter, sl, rigid)
        for a, b in zip(ref, out):
            np.testing.assert_array_equal(a, b)


def test_force_only_launch_matches_and_backward_uses_it():
    """compute_batch(..., energies=False) skips the FP64 energy pass (NULL energies in the C-ABI): same force bits."""
    g = load_golden("ala10")
    for mode in MODES:
        F = Forces(gpu_par(g), **MODES[mode])
        p = torch.tensor(g["pos"], device=DEV)
        s1, s2 = System(104, len(p), torch.float, DEV), System(104, len(p), torch.float, DEV)
        e = F.compute_batch(p, s1.box, s1.forces)
        assert F.compute_batch(p, s2.box, s2.forces, energies=False) is None
        assert torch.equal(s1.forces, s2.forces) and e.shape == (len(p), 8)


def test_compute_batch_host_pipelines_copies_and_matches_compute_batch():
    """Forces.compute_batch_host: host buffers in and out, the batch cut into chunks over three streams.  Same kernels on the
    same inputs: bit-identical to compute_batch, for ragged chunk sizes, with a box, force-only, and twice in a row."""
    g = load_golden("ala10")
    rng = np.random.default_rng(3)
    R = 1001
    pos = (g["pos"][rng.integers(0, len(g["pos"]), R)] + 0.02 * rng.standard_normal((R, 104, 3))).astype(np.float32)
    hp = torch.tensor(pos).pin_memory()
    box = torch.zeros((R, 3, 3))
    box[::3, 0, 0], box[::3, 1, 1], box[::3, 2, 2] = 41.0, 38.0, 44.0
    for mode, b in (("plain", None), ("rf", box)):
        F = Forces(gpu_par(g), **MODES[mode])
        dp = torch.tensor(pos, device=DEV)
        db = torch.zeros((R, 3, 3), device=DEV) if b is None else b.to(DEV)
        f_ref = torch.empty_like(dp)
        e_ref = F.compute_batch(dp, db, f_ref)
        for chunks in (1, 7, 8):
            f, e = F.compute_batch_host(hp, box=b, chunks=chunks, device=DEV)
            assert f.is_pinned() and torch.equal(f, f_ref.cpu()) and torch.equal(e, e_ref.cpu()), (mode, chunks)
        f2, e2 = F.compute_batch_host(hp, box=b, energies=False, device=DEV)
        assert e2 is None and torch.equal(f2, f_ref.cpu())
    bad = hp.clone()
    bad[500, 3, 1] = float("nan")
    with pytest.raises(RuntimeError, match="NaN"):
        F.compute_batch_host(bad, device=DEV)
    # a solute above 128 atoms through the same call (large-solute kernels, their second stream and workspace under the
    # chunk stream): 3 deca-alanine chains as one 312-atom topology
    copies, RL = 3, 40
    parL = Parameters.from_arrays(g.par.tiled(copies).to_arrays(), device=DEV)
    shift = np.repeat(np.arange(copies), 104)[None, :, None] * np.array([0.0, 11.0, 11.0])
    posL = (pos[:RL * copies].reshape(RL, copies * 104, 3) + shift).astype(np.float32)
    FL = Forces(parL, **MODES["rf"])
    dpL = torch.tensor(posL, device=DEV)
    fL_ref = torch.empty_like(dpL)
    eL_ref = FL.compute_batch(dpL, torch.zeros((RL, 3, 3), device=DEV), fL_ref)
    for chunks in (1, 3):
        fL, eL = FL.compute_batch_host(torch.tensor(posL).pin_memory(), chunks=chunks, device=DEV)
        assert torch.equal(fL, fL_ref.cpu()) and torch.equal(eL, eL_ref.cpu()), chunks


def test_report_after_the_last_step_when_niter_is_not_a_multiple():
    """Integrator.step returns the energies of the END state (upstream's contract): with energy_every = 4 and 10 steps the
    records are steps 4, 8 and 10."""
    g = load_golden("diala")
    pos0 = g["pos"]
    par, s = make_system(g, pos0, np.zeros_like(pos0))
    F = Forces(par)
    F.compute_batch(s.pos, s.box, s.forces)
    f0 = s.forces.cpu().numpy().astype(np.float64)
    integ = Integrator(s, F, timestep=0.5, device=DEV)
    integ.energy_every = 4
    ekin, pot, T = integ.step(niter=10)
    assert integ.observables.shape[0] == 3
    out = oracle.integrate(g.packed, pos0, np.zeros_like(pos0), f0, nsteps=10, dt=np.float32(integ.dt), energy_every=4)
    ref = out["observables"]
    assert ref.shape[0] == 3
    np.testing.assert_allclose(integ.observables.cpu().numpy()[:, :, :8].sum(-1), ref[:, :, :8].sum(-1), rtol=2e-5, atol=2e-3)
    e_end = F.compute_batch(s.pos, s.box, torch.empty_like(s.forces)).sum(1).cpu().numpy()
    np.testing.assert_allclose(np.asarray(pot), e_end, rtol=1e-6)


def test_ctypes_only_host_entry_point_vs_reference_fp64():
    """The binding INTEGRATION.md promises a non-torch host: b200md_topology_create -> b200md_forces_compute_host with
    plain numpy buffers through ctypes (no torch tensor, op or stream on the call path), compared with the stored fp64
    output of the reference's forces.py on bb40.npz."""
    import ctypes as C

    g = load_golden("bb40")
    lib = _cabi.load_library()
    pos = np.ascontiguousarray(g["pos"], dtype=np.float32)
    R, N, _ = pos.shape
    h = lib.b200md_topology_create(C.byref(g.packed.desc), 0)
    assert h, _cabi.last_error()
    try:
        for mode, kw in MODES.items():
            fo = _cabi.ForceOpts(sum(_cabi.TERM_BITS[t] for t in ("electrostatics", "lj", "bonds", "angles", "dihedrals", "1-4", "impropers")),
                                 float(kw.get("cutoff", 0.0)), int(kw.get("rfa", False)), 78.5, float(kw.get("switch_dist", 0.0)))
            so = _cabi.SmdOpts(0, 0, 0, 0.0, 0.0, 0.0, None)
            frc = np.full((R, N, 3), np.nan, dtype=np.float32)
            ene = np.full((R, 8), np.nan, dtype=np.float64)
            status = np.full(R, 99, dtype=np.uint32)
            rc = lib.b200md_forces_compute_host(C.c_void_p(h), C.byref(fo), C.byref(so), R, pos.ctypes.data_as(C.c_void_p), None,
                                                frc.ctypes.data_as(C.c_void_p), ene.ctypes.data_as(C.c_void_p),
                                                status.ctypes.data_as(C.c_void_p))
            assert rc == _cabi.OK, _cabi.last_error()
            assert (status == 0).all()
            check(g, pos, frc.astype(np.float64), ene, g[f"ref_{mode}_f64_forces"], g[f"ref_{mode}_f64_energies"], f"host {mode}")
        bad = pos.copy()
        bad[0, 0, 0] = np.nan
        rc = lib.b200md_forces_compute_host(C.c_void_p(h), C.byref(fo), C.byref(so), R, bad.ctypes.data_as(C.c_void_p), None,
                                            frc.ctypes.data_as(C.c_void_p), ene.ctypes.data_as(C.c_void_p), None)
        assert rc == _cabi.ERR_NAN_COORDINATES and "NaN" in _cabi.last_error()
    finally:
        lib.b200md_topology_destroy(C.c_void_p(h))


class _ToyExternal:
    """forces.py:318-323: anything with calculate(pos, box) -> (energy (R,), force (R, N, 3)).  A harmonic tether of every
    atom to the origin, written with torch ops so that it also carries an autograd edge."""

    def __init__(self, k):
        self.k = k

    def calculate(self, pos, box):
        return 0.5 * self.k * (pos * pos).sum(dim=(1, 2)), -self.k * pos


def test_generic_external_object_in_compute_autograd_and_integrator():
    g = load_golden("bb40")
    pos0 = np.ascontiguousarray(g["frames"][:3], dtype=np.float32)
    k = 0.05
    # compute: energy in the "external" column, force added (forces.py:318-323)
    F = Forces(gpu_par(g), external=_ToyExternal(k))
    F0 = Forces(gpu_par(g))
    p = torch.tensor(pos0, device=DEV)
    s, s0 = System(40, 3, torch.float, DEV), System(40, 3, torch.float, DEV)
    e, e0 = F.compute_batch(p, s.box, s.forces), F0.compute_batch(p, s0.box, s0.forces)
    np.testing.assert_allclose(e[:, 7].cpu().numpy(), 0.5 * k * (pos0.astype(np.float64) ** 2).sum((1, 2)), rtol=1e-6)
    np.testing.assert_allclose((s.forces - s0.forces).cpu().numpy(), -k * pos0, rtol=1e-4, atol=1e-5)
    # autograd with equal weights on every term: the external gradient is counted once
    pg = p.clone().requires_grad_(True)
    tot = F.compute(pg, s.box, torch.zeros_like(s.forces), explicit_forces=False)
    tot.backward()
    np.testing.assert_allclose(pg.grad.cpu().numpy(), -s.forces.cpu().numpy(), rtol=1e-5, atol=1e-4)
    # Integrator.step: launches per step with the hook in between, against the oracle driven the same way on the host
    vel0 = np.zeros_like(pos0)
    par, sy = make_system(g, pos0, vel0)
    Fi = Forces(par, external=_ToyExternal(k))
    Fi.compute_batch(sy.pos, sy.box, sy.forces)
    integ = Integrator(sy, Fi, timestep=0.5, device=DEV)
    ekin, pot, T = integ.step(niter=12)
    x = pos0.astype(np.float64)
    v = np.zeros_like(x)
    f, _ = oracle.forces(g.packed, x.astype(np.float32))
    f = f - k * x
    m = g.par.masses.numpy().astype(np.float64).reshape(1, -1, 1)
    dt = float(np.float32(integ.dt))
    for _ in range(12):
        v += 0.5 * dt * f / m
        x += dt * v
        f, en = oracle.forces(g.packed, x)
        f = f - k * x
        v += 0.5 * dt * f / m
    assert np.abs(sy.pos.cpu().numpy() - x).max() < 2e-4
    np.testing.assert_allclose(np.asarray(pot), en.sum(1) + 0.5 * k * (x ** 2).sum((1, 2)), rtol=2e-5)
    np.testing.assert_allclose(ekin, (0.5 * m * v * v).sum((1, 2)), rtol=2e-3)


def test_two_forwards_before_one_backward_keep_their_own_smd_centres():
    """the backward of the batched Energy adapter must use the restraint centres of ITS forward, not whatever
    bias.center holds when backward runs (real + fake batch, gradient accumulation; different batch sizes)."""
    from molecular_dynamics_b200 import Energy

    g = load_golden("ala10")
    colvar = {"name": "E2End Harm", "fk": 1.0, "cent_0": 12.0, "cent_1": 34.0, "T": 1.0, "group1": [3], "group2": [98]}
    en = Energy(parameters=gpu_par(g), colvar=colvar, device=DEV)
    xa = torch.tensor(g["pos"][:4], device=DEV)
    xb = torch.tensor(g["pos"][:3] + 0.01, device=DEV)
    ta, tb = torch.tensor([0.1, 0.3, 0.5, 0.7], device=DEV), torch.tensor([0.9, 0.2, 0.6], device=DEV)

    def loss(ene):
        return sum(v for k, v in ene.items() if k != "E2End Harm").sum() + 200.0 * ene["E2End Harm"].sum()

    # separately
    grads = []
    for x, t in ((xa, ta), (xb, tb)):
        xg = x.clone().requires_grad_(True)
        loss(en.calc_energy_batch(xg, t)[0]).backward()
        grads.append(xg.grad.clone())
    # both forwards first, one backward
    xga, xgb = xa.clone().requires_grad_(True), xb.clone().requires_grad_(True)
    la = loss(en.calc_energy_batch(xga, ta)[0])
    lb = loss(en.calc_energy_batch(xgb, tb)[0])
    (la + lb).backward()
    assert torch.equal(xga.grad, grads[0]) and torch.equal(xgb.grad, grads[1])


def test_double_precision_tensors_are_accepted():
    """systems.py:32-36: the reference classes take precision=torch.double.  The drop-in accepts float64 positions / forces
    (kernels compute forces in fp32, energies in FP64) and hands results back in the caller's dtype."""
    g = load_golden("ala10")
    F = Forces(gpu_par(g), **MODES["rf"])
    s = System(104, len(g["pos"]), torch.double, DEV)
    s.pos[:] = torch.tensor(g["pos"], device=DEV, dtype=torch.double)
    pot = F.compute(s.pos, s.box, s.forces, returnDetails=True)
    assert s.forces.dtype == torch.float64 and pot[0]["lj"].dtype == torch.float64
    e = np.array([[float(p[k]) for k in _cabi.ENERGY_SLOTS] for p in pot])
    check(g, g["pos"], s.forces.cpu().numpy(), e, g["ref_rf_f64_forces"], g["ref_rf_f64_energies"], "double tensors")
    phi = F.dihedral_angles(s.pos)
    assert phi.dtype == torch.float64


def test_library_calls_leave_the_current_device_alone():
    """an entry point runs on the topology's device and restores the caller's current device."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    g = load_golden("bb40")
    par = Parameters.from_arrays(g.par.to_arrays(), device="cuda:1")
    F = Forces(par)
    p = torch.tensor(g["pos"], device="cuda:1")
    s = System(40, len(p), torch.float, "cuda:1")
    torch.cuda.set_device(0)
    e = F.compute_batch(p, s.box, s.forces)
    assert torch.cuda.current_device() == 0
    integ = Integrator(s, F, timestep=0.5, device="cuda:1")
    s.pos[:] = p
    integ.step(niter=3)
    assert torch.cuda.current_device() == 0
    f_ref, e_ref = oracle.forces(g.packed, g["pos"])
    check_energies(e.cpu().numpy(), e_ref, "cuda:1")
    # a solute above 128 atoms on the other device: the handle's second stream, workspace and cooperative launch live there
    gl = load_golden("ala10x3")
    res = []
    for dev in ("cuda:1", "cuda:0"):
        FL = Forces(Parameters.from_arrays(gl.par.to_arrays(), device=dev), **MODES["rf"])
        sl = System(312, len(gl["pos"]), torch.float, dev)
        sl.pos[:] = torch.tensor(gl["pos"], device=dev)
        torch.cuda.set_device(0)
        el = FL.compute_batch(sl.pos, sl.box, sl.forces)
        integ = Integrator(sl, FL, timestep=1.0, device=dev, gamma=1.0, T=300.0, seed=2)
        integ.energy_every = 4
        integ.step(niter=9)
        assert torch.cuda.current_device() == 0
        res.append((el.cpu(), sl.pos.cpu(), integ.observables.cpu()))
    for a, b in zip(*res):
        assert torch.equal(a, b)


def test_fire_update_kernel_vs_restatement():
    """the fused FIRE update kernel (b200md_fire_update) against its FP64 restatement (oracle/fire_reference.py),
    iteration by iteration on the same force field: 25 iterations from perturbed ALA10 conformers, including uphill
    steps (velocity reset, time-step cut) and capped displacements."""
    import ctypes as C

    import bench
    from oracle.fire_reference import fire_update

    g = load_golden("ala10")
    R = 8
    pos0, _ = bench.synth_inputs(bench.load_ala10(), R)
    par, s = make_system(g, pos0, np.zeros_like(pos0))
    F = Forces(par)
    lib = _cabi.load_library()
    opts = _cabi.FireOpts(0.2, 0.1, 1.1, 0.5, 0.1, 0.99, 5)
    vel = torch.zeros_like(s.pos)
    state = torch.zeros((R, 4), dtype=torch.float32, device=DEV)
    state[:, 0], state[:, 1] = 0.05, 0.1
    x, v = pos0.astype(np.float64), np.zeros(pos0.shape)
    dt, alpha, npos = np.full(R, 0.05), np.full(R, 0.1), np.zeros(R)
    sr = System(104, R, torch.float, DEV)
    ups = 0
    for it in range(25):
        F.compute_batch(s.pos, s.box, s.forces, energies=False)
        rc = lib.b200md_fire_update(0, R, 104, C.c_void_p(s.pos.data_ptr()), C.c_void_p(vel.data_ptr()), C.c_void_p(s.forces.data_ptr()),
                                    C.c_void_p(state.data_ptr()), C.byref(opts), C.c_void_p(torch.cuda.current_stream().cuda_stream))
        assert rc == _cabi.OK, _cabi.last_error()
        sr.pos[:] = torch.tensor(x.astype(np.float32), device=DEV)
        F.compute_batch(sr.pos, sr.box, sr.forces, energies=False)
        x, v, dt, alpha, npos = fire_update(x, v, sr.forces.cpu().numpy().astype(np.float64), dt, alpha, npos)
        st = state.cpu().numpy()
        ups += int((npos == 0).sum())
        np.testing.assert_allclose(st[:, 0], dt, rtol=1e-5)
        np.testing.assert_allclose(st[:, 1], alpha, rtol=1e-5)
        np.testing.assert_array_equal(st[:, 2], npos)
        assert np.abs(s.pos.cpu().numpy() - x).max() < 2e-4, it
    assert ups > 0  # the comparison saw uphill resets
