"""b200md_integrate_noalloc -- the entry point that allocates nothing (include/b200md.h; no counterpart in the reference)
-- against b200md_integrate, both called through the C-ABI with ctypes.

Collected last on purpose: a failure here hides no other test.  Passes alone in a fresh
process and after the rest of the suite (it builds its own topology handle and buffers; the only shared state is the
library's fixed device scratch array, which every call re-arms).
"""
import ctypes as C

import numpy as np
import pytest
import torch

from conftest import load_golden
from molecular_dynamics_b200 import Forces, SMDBias, _cabi
from molecular_dynamics_b200.parameters import Parameters

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
E_TOL = 1e-6  # the energy tolerance of tests/test_gpu_parity.py


def _vp(t):
    return C.c_void_p(t.data_ptr())


def _call(fn, handle, state, R, nsteps, step0, energy_every, snapshot_every, smd, langevin):
    """One integrate call through the C-ABI on copies of `state`; returns (pos, vel, forces, observables, snapshots, status)."""
    pos, vel, frc = (t[:R].clone().contiguous() for t in state)
    N = pos.shape[1]
    nrep = -(-nsteps // energy_every)
    obs = torch.full((nrep, R, _cabi.NUM_OBSERVABLES), float("nan"), dtype=torch.float64, device=DEV)
    snaps = torch.full((nsteps // snapshot_every, R, N, 3), float("nan"), dtype=torch.float32, device=DEV)
    status = torch.full((R,), 99, dtype=torch.int32, device=DEV)
    terms = sum(_cabi.TERM_BITS[t] for t in ("electrostatics", "lj", "bonds", "angles", "dihedrals", "1-4", "impropers"))
    fo = _cabi.ForceOpts(terms, 0.0, 0, 78.5, 0.0)
    so = _cabi.SmdOpts(1, 3, 98, 1.0, 14.0, 1e-3, None) if smd else _cabi.SmdOpts(0, 0, 0, 0.0, 0.0, 0.0, None)
    gamma, temp = (1.0 / (1000.0 / 48.88821), 300.0) if langevin else (0.0, 0.0)
    io = _cabi.IntegrateOpts(1.0 / 48.88821, gamma, temp, 5, step0, 0, nsteps, energy_every, snapshot_every, 0)
    stream = C.c_void_p(torch.cuda.current_stream(torch.device(DEV)).cuda_stream)
    rc = fn(C.c_void_p(handle), C.byref(fo), C.byref(so), C.byref(io), R, _vp(pos), _vp(vel), _vp(frc), None, _vp(obs),
            _vp(snaps), _vp(status), stream)
    assert rc == _cabi.OK, _cabi.last_error()
    torch.cuda.synchronize()
    return pos, vel, frc, obs, snaps, status


def test_integrate_noalloc_matches_integrate_on_two_calls_in_a_row():
    g = load_golden("ala10")
    lib = _cabi.load_library()
    par = Parameters.from_arrays(g.par.to_arrays(), device=DEV)
    F = Forces(par, external=SMDBias([3], [98], fk=1.0, cent_0=14.0))
    handle = F._handle(torch.device(DEV))
    R = 900  # more replicas than resident CTAs (740): units of several slices are in flight for one replica
    rng = np.random.default_rng(21)
    pos0 = (g["pos"][rng.integers(0, len(g["pos"]), R)] + 0.02 * rng.standard_normal((R, 104, 3))).astype(np.float32)
    pos = torch.tensor(pos0, device=DEV)
    vel = torch.tensor((0.01 * rng.standard_normal(pos0.shape)).astype(np.float32), device=DEV)
    frc = torch.zeros_like(pos)
    F.compute_batch(pos, torch.zeros((R, 3, 3), device=DEV), frc, energies=False)
    state = (pos, vel, frc)

    # call 1: Langevin + moving restraint (the carried-work slot of the scratch array is in use), 900 replicas, 37 steps;
    # call 2: NVE, fewer replicas, continuing from call 1's end state -- the array still holds call 1's slice counts
    plans = [dict(R=R, nsteps=37, step0=0, energy_every=6, snapshot_every=9, smd=True, langevin=True),
             dict(R=300, nsteps=41, step0=37, energy_every=10, snapshot_every=20, smd=False, langevin=False)]
    for k, plan in enumerate(plans):
        ref = _call(lib.b200md_integrate, handle, state, **plan)
        out = _call(lib.b200md_integrate_noalloc, handle, state, **plan)
        for name, a, b in zip(("pos", "vel", "forces"), ref[:3], out[:3]):
            assert torch.isfinite(b).all(), (k, name)
            assert torch.equal(a, b), (k, name)  # the same arithmetic whatever the scratch: bit-identical
        assert torch.equal(ref[4], out[4]) and torch.equal(ref[5], out[5]), k  # snapshots, status flags
        assert int(out[5].max()) == 0
        o_ref, o = ref[3].cpu().numpy(), out[3].cpu().numpy()
        assert np.isfinite(o).all()
        np.testing.assert_allclose(o, o_ref, rtol=E_TOL, atol=1e-9)
        state = tuple(t.clone() for t in out[:3])  # the second call starts where the first one ended

    # sizes beyond the fixed array are refused with a clear error, before anything is launched
    fo = _cabi.ForceOpts(_cabi.TERM_BITS["bonds"], 0.0, 0, 78.5, 0.0)
    so = _cabi.SmdOpts(0, 0, 0, 0.0, 0.0, 0.0, None)
    io = _cabi.IntegrateOpts(1.0 / 48.88821, 0.0, 0.0, 0, 0, 0, 4, 0, 0, 0)
    rc = lib.b200md_integrate_noalloc(C.c_void_p(handle), C.byref(fo), C.byref(so), C.byref(io), _cabi.NOALLOC_MAX_REPLICAS + 1,
                                      _vp(pos), _vp(vel), _vp(frc), None, None, None, None, None)
    assert rc == _cabi.ERR_UNSUPPORTED
    assert "B200MD_NOALLOC_MAX_REPLICAS" in _cabi.last_error() and str(_cabi.NOALLOC_MAX_REPLICAS) in _cabi.last_error()
