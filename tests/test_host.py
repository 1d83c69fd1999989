"""Host-side logic and the C-ABI surface (no GPU needed)."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from conftest import ROOT, SYSTEMS, load_golden
from molecular_dynamics_b200 import Forces, System, _cabi


def test_library_loads_and_exports_every_declared_symbol():
    import __graft_entry__ as entry

    entry.build()
    header = open(os.path.join(ROOT, "include", "b200md.h")).read()
    declared = set(re.findall(r"\b(b200md_[a-z0-9_]+)\s*\(", header))
    declared -= {"b200md_topology_desc", "b200md_force_opts", "b200md_smd_opts", "b200md_integrate_opts"}
    assert declared == set(_cabi.EXPORTED_SYMBOLS)
    lib = ctypes.CDLL(_cabi.LIB_PATH)
    for sym in declared:
        assert hasattr(lib, sym), sym
    assert b"sm_100a" in _cabi.load_library().b200md_version()


def test_struct_layouts_match_header_sizes(tmp_path):
    """sizeof of every struct as gcc sees include/b200md.h equals the ctypes mirror."""
    import subprocess

    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "b200md.h"\nint main(){printf("%zu %zu %zu %zu\\n",'
                   "sizeof(b200md_topology_desc),sizeof(b200md_force_opts),sizeof(b200md_smd_opts),"
                   "sizeof(b200md_integrate_opts));return 0;}\n")
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    sizes = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    mirror = [ctypes.sizeof(c) for c in (_cabi.TopologyDesc, _cabi.ForceOpts, _cabi.SmdOpts, _cabi.IntegrateOpts)]
    assert sizes == mirror


def test_product_path_has_no_cpu_fallback():
    g = load_golden("diala")
    F = Forces(g.par)
    s = System(22, 1, torch.float, "cpu")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        F.compute(s.pos, s.box, s.forces)
    # nothing under molecular_dynamics_b200/ imports the oracle
    pkg = os.path.join(ROOT, "molecular_dynamics_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert not re.search(r"^\s*(import|from)\s+oracle", src, re.M), fn


def test_forces_constructor_contract():
    g = load_golden("bb40")
    F = Forces(g.par, cutoff=9, rfa=True, switch_dist=7.5)
    assert F.natoms == 40 and F.require_distances and F.cutoff == 9 and F.rfa and F.switch_dist == 7.5
    assert F.energies == ["electrostatics", "lj", "bonds", "angles", "dihedrals", "1-4", "impropers"]
    assert F.ava_idx.dtype == torch.int64 and tuple(F.ava_idx.shape) == (648, 2)
    np.testing.assert_array_equal(F.ava_idx.numpy(), g.ava_idx)
    assert Forces(g.par, terms=("bonds",)).ava_idx is None
    with pytest.raises(ValueError):
        Forces(g.par, terms=("bonds", "nope"))
    with pytest.raises(RuntimeError):
        Forces(g.par, terms=("1-4", "lj"))


def test_torsion_slice_flattening_keeps_reference_branch_rule():
    """forces.py:565: a slice is AMBER iff ALL its periodicities are > 0, else every term is harmonic."""
    sl = [{"idx": torch.tensor([0, 1]), "params": torch.tensor([[1.0, 0.0, 2.0], [2.0, 0.1, 0.0]])},
          {"idx": torch.tensor([1]), "params": torch.tensor([[0.5, 3.14, 3.0]])}]
    owner, prm, harm = _cabi._flatten_torsion_slices(sl)
    np.testing.assert_array_equal(owner, [0, 1, 1])
    np.testing.assert_array_equal(harm, [1, 1, 0])
    assert prm.shape == (3, 3)


def test_system_setters_follow_reference():
    s = System(5, 2, torch.float, "cpu")
    assert s.natoms == 5 and s.nreplicas == 2 and s.box.shape == (2, 3, 3)
    src = torch.arange(30, dtype=torch.float32).reshape(5, 3, 2).requires_grad_(True)
    s.set_positions(src)
    assert s.pos.shape == (2, 5, 3) and s.pos.requires_grad and not s.pos.is_contiguous()
    with pytest.raises(RuntimeError):
        s.set_positions(torch.zeros(5, 4, 2))
    s.set_box(np.array([10.0, 11.0, 12.0]))
    np.testing.assert_array_equal(s.box[1].numpy(), np.diag([10.0, 11.0, 12.0]))
    with pytest.raises(RuntimeError):
        s.set_velocities(torch.zeros(1, 5, 3))
    s.set_velocities(torch.ones(2, 5, 3, dtype=torch.float64))
    assert s.vel.dtype == torch.float32 and float(s.vel.sum()) == 30.0
    s.set_forces(np.ones((2, 5, 3)))
    assert float(s.forces.sum()) == 30.0


def test_packed_topology_counts():
    want = {"ala10": (104, 12, 103, 183, 250, 260, 250, 4820), "bb40": (40, 4, 39, 47, 46, 46, 46, 648)}
    for name, w in want.items():
        d = load_golden(name).packed.desc
        assert (d.natoms, d.ntypes, d.nbonds, d.nangles, d.ndihedrals, d.ndihedral_terms, d.n14, d.npairs) == w


def test_logwriter_and_trajectory_writer(tmp_path):
    import csv

    import torch

    from molecular_dynamics_b200 import LogWriter, TrajectoryWriter

    log = LogWriter(path=str(tmp_path / "logs"), keys=("iter", "ns", "epot", "ekin", "etot", "T"), name="monitor.csv")
    log.write_row({"iter": 10, "ns": 1e-5, "epot": [torch.tensor(-3.5)], "ekin": np.array([2.0]), "etot": np.array([-1.5]), "T": np.array([300.0])})
    log.write_row({"iter": 20, "ns": 2e-5, "epot": torch.tensor([1.0, 2.0]), "ekin": np.array([2.0, 3.0]), "etot": np.array([3.0, 5.0]), "T": 1})
    rows = list(csv.reader(open(tmp_path / "logs" / "monitor.csv")))
    assert rows[0] == ["iter", "ns", "epot", "ekin", "etot", "T"]
    assert rows[1][0] == "10" and float(rows[1][2]) == -3.5 and float(rows[1][5]) == 300.0
    assert rows[2][2] == "1.0;2.0"
    tw = TrajectoryWriter(str(tmp_path / "traj.npy"))
    for k in range(3):
        tw.append(torch.full((1, 5, 3), float(k)))
    assert tw.save() == [str(tmp_path / "traj.npy")]
    assert np.load(tmp_path / "traj.npy").shape == (3, 5, 3)  # (frames, natoms, 3): the reference's trajectory layout
    tw = TrajectoryWriter(str(tmp_path / "t2.npy"), layout="tutorial")
    tw.append(np.zeros((2, 5, 3)))
    tw.append(np.ones((2, 5, 3)))
    tw.save()
    assert np.load(tmp_path / "t2.npy").shape == (2, 5, 3, 2)


def test_wrapper_moves_whole_groups_into_the_box():
    import torch

    from molecular_dynamics_b200 import Wrapper

    bonds = np.array([[0, 1], [1, 2], [4, 5]])
    w = Wrapper(7, bonds)
    assert sorted(len(g) for g in w.groups) == [2, 3] and sorted(w.nongrouped.tolist()) == [3, 6]
    pos = torch.tensor([[[10.5, 0.1, 0.1], [11.0, 0.2, 0.1], [11.5, 0.3, 0.1], [-0.5, 25.0, 3.0],
                         [0.2, 0.2, -7.9], [0.4, 0.2, -8.1], [3.0, 3.0, 3.0]]])
    ref = pos.clone()
    box = torch.zeros(1, 3, 3)
    w.wrap(pos, box)
    assert torch.equal(pos, ref)  # vacuum: no-op
    box[0, 0, 0], box[0, 1, 1], box[0, 2, 2] = 10.0, 10.0, 10.0
    w.wrap(pos, box)
    np.testing.assert_allclose(pos[0, :3, 0], [0.5, 1.0, 1.5], atol=1e-6)  # group moved as a whole (centroid 11 -> 1)
    np.testing.assert_allclose(pos[0, 3], [9.5, 5.0, 3.0], atol=1e-6)
    np.testing.assert_allclose(pos[0, 4:6, 2], [2.1, 1.9], atol=1e-5)
    np.testing.assert_allclose(pos[0, 1] - pos[0, 0], ref[0, 1] - ref[0, 0], atol=1e-6)  # intra-group geometry kept


def test_colvars_traj_round_trip_and_reference_layout(tmp_path):
    import os

    from molecular_dynamics_b200 import read_colvars_traj, write_colvars_traj

    ref = os.path.join(os.path.dirname(__file__), "golden", "smd_out_head.colvars.traj")
    steps, d, c, w = read_colvars_traj(ref)
    out = str(tmp_path / "mine.colvars.traj")
    write_colvars_traj(out, steps, d, c, w)
    s2, d2, c2, w2 = read_colvars_traj(out)
    np.testing.assert_array_equal(s2, steps)
    for a, b in ((d, d2), (c, c2), (w, w2)):
        np.testing.assert_array_equal(a, b)  # %.14e both ways: the text round-trips exactly
    ref_lines, my_lines = open(ref).read().splitlines(), open(out).read().splitlines()
    assert my_lines[0].split() == ref_lines[0].split()
    assert [ln.split() for ln in my_lines[1:6]] == [ln.split() for ln in ref_lines[1:6]]
    with pytest.raises(ValueError):
        write_colvars_traj(out, steps, d[:-1], c, w)


def test_jarzynski_estimator_single_rank_limits():
    import torch

    from molecular_dynamics_b200 import BOLTZMAN, jarzynski_free_energy

    kT = BOLTZMAN * 300.0
    # identical work on every replica: dF = W exactly, zero variance
    r = jarzynski_free_energy(torch.full((4, 100), 3.25, dtype=torch.float64), 300.0)
    np.testing.assert_allclose(r["dF"].numpy(), 3.25, rtol=1e-12)
    np.testing.assert_allclose(r["dF_cumulant2"].numpy(), 3.25, rtol=1e-12)
    # Gaussian work: dF -> mean - var / 2kT; large offsets must not overflow (shifted exponentials)
    g = torch.Generator().manual_seed(0)
    w = 500.0 + 0.3 * torch.randn((200000,), generator=g, dtype=torch.float64)
    r = jarzynski_free_energy(w, 300.0)
    assert abs(float(r["dF"]) - (500.0 - 0.09 / (2 * kT))) < 5e-3
    assert abs(float(r["dF_cumulant2"]) - (500.0 - 0.09 / (2 * kT))) < 5e-3
    assert float(r["dF"]) <= float(r["mean_work"])  # Jensen: dF <= <W>
    ref = -kT * (torch.logsumexp(-w / kT, 0) - np.log(w.numel()))
    np.testing.assert_allclose(float(r["dF"]), float(ref), rtol=1e-12)


def test_namd_restart_files_against_the_chamber_prmtop():
    """read_namd_bin / read_xsc on the reference's own restart of replica 3 (copied by tools/make_golden_chamber.py).
    The run used rigidBonds ALL, so every X-H distance in the .coor file equals the prmtop's equilibrium length: one
    check that pins the binary reader, the prmtop bond table and its atom indexing together.  The .vel file is in
    A / 48.88821 fs, the integrator's unit: the kinetic temperature over the unconstrained degrees of freedom comes
    out at the thermostat's 300 K within the fluctuation of 259 degrees of freedom."""
    import os

    from molecular_dynamics_b200 import BOLTZMAN
    from molecular_dynamics_b200.parameters import parameters_from_prmtop
    from molecular_dynamics_b200.topology_io import read_namd_bin, read_xsc

    gold = os.path.join(os.path.dirname(__file__), "golden")
    x = read_namd_bin(os.path.join(gold, "smd_out.restart.coor"))
    v = read_namd_bin(os.path.join(gold, "smd_out.restart.vel"))
    xsc = read_xsc(os.path.join(gold, "smd_out.restart.xsc"))
    assert x.shape == v.shape == (104, 3) and xsc["step"] == 500000 and not xsc["box"].any()
    par = parameters_from_prmtop(os.path.join(gold, "full_da-1.3.prmtop"), precision=torch.float64)
    b, bp, m = par.bonds.numpy(), par.bond_params.numpy(), par.masses.numpy().ravel()
    r = np.linalg.norm(x[b[:, 0]] - x[b[:, 1]], axis=1)
    xh = (m[b[:, 0]] < 1.5) | (m[b[:, 1]] < 1.5)
    assert xh.sum() == 53 and np.abs(r[xh] - bp[xh, 1]).max() < 1e-6
    assert np.abs(r[~xh] - bp[~xh, 1]).max() < 0.2  # the heavy-atom bonds fluctuate thermally
    ekin = 0.5 * (m[:, None] * v ** 2).sum()
    temp = 2.0 * ekin / ((3 * 104 - 53) * BOLTZMAN)
    assert 200.0 < temp < 400.0
    with pytest.raises(ValueError):
        read_namd_bin(os.path.join(gold, "smd_out.restart.xsc"))


def test_parameters_tiled_is_copies_of_the_molecule():
    """Parameters.tiled (bundle of chains in one topology, used by the large-solute bench): bonded energies of three
    far-apart copies are three times the molecule's, the pair list grows by the inter-molecular pairs."""
    import oracle
    from molecular_dynamics_b200 import Parameters, _cabi

    with np.load(os.path.join(os.path.dirname(__file__), "golden", "ala10.npz")) as z:
        g = {k: z[k] for k in z.files}
    par = Parameters.from_arrays({k[4:]: v for k, v in g.items() if k.startswith("par_")})
    p3 = par.tiled(3)
    assert len(p3.masses) == 312 and len(p3.bonds) == 3 * len(par.bonds) and len(p3.idx14) == 3 * len(par.idx14)
    idx1 = _cabi.make_pair_index(104, par.get_exclusions())
    idx3 = _cabi.make_pair_index(312, p3.get_exclusions())
    assert len(idx3) == 3 * len(idx1) + 3 * 104 * 104
    pos = g["pos"][:1]
    pos3 = np.concatenate([pos + np.float32(64.0 * c) for c in range(3)], 1)
    _, e1 = oracle.forces(_cabi.PackedTopology(par, idx1), pos)
    _, e3 = oracle.forces(_cabi.PackedTopology(p3, idx3), pos3)
    np.testing.assert_allclose(e3[0, 2:5], 3 * e1[0, 2:5], rtol=2e-5)
