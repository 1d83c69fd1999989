"""CPU oracle for the force-field + integrator path -- TEST INFRASTRUCTURE ONLY.

Only tests/, ``__graft_entry__.smoke()`` and bench.py's ``cpu_baseline`` / ``--impl reference`` legs
may import this package, and only as the checker.  ``molecular_dynamics_b200`` never imports it.

Two restatements live here:

* ``md_oracle.c``  -- FP64 scalar C (OpenMP over replicas); fast enough for 65 536 replicas and
  100 k-step drift runs.  Wrapped by :func:`forces`, :func:`integrate`, :func:`normal3`.
* ``torchmd_port.py`` -- torch-CPU restatement that keeps the reference's op-per-term, Python
  loop-per-replica cost profile; used as the timed CPU baseline ("port").

Pinning: tests/test_oracle.py checks both against the notebook known-answer vector
(V_Calculations/Test-1/tutorial.ipynb:231-271 of the reference) and against tests/golden/*.npz, which
tools/make_golden.py produced by running the reference's unmodified TorchMD_Fix/forces.py.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from molecular_dynamics_b200 import _cabi

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "md_oracle.c")
_SO = os.path.join(_HERE, "_build", "libmd_oracle.so")
_LIB = None


def build(force: bool = False) -> str:
    """gcc -O2 -fopenmp md_oracle.c -> oracle/_build/libmd_oracle.so (rebuilt when the source is newer)."""
    hdr = os.path.join(_HERE, "..", "include", "b200md.h")
    if (not force and os.path.exists(_SO)
            and os.path.getmtime(_SO) >= max(os.path.getmtime(_SRC), os.path.getmtime(hdr))):
        return _SO
    os.makedirs(os.path.dirname(_SO), exist_ok=True)
    cmd = ["gcc", "-O2", "-fopenmp", "-fPIC", "-shared", "-std=gnu11", "-o", _SO, _SRC, "-lm"]
    subprocess.run(cmd, check=True)
    return _SO


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        vp = C.c_void_p
        L.oracle_forces.restype = C.c_int
        L.oracle_forces.argtypes = [C.POINTER(_cabi.TopologyDesc), C.POINTER(_cabi.ForceOpts),
                                    C.POINTER(_cabi.SmdOpts), C.c_int64, vp, vp, vp, vp, vp]
        L.oracle_integrate.restype = C.c_int
        L.oracle_integrate.argtypes = [C.POINTER(_cabi.TopologyDesc), C.POINTER(_cabi.ForceOpts),
                                       C.POINTER(_cabi.SmdOpts), C.POINTER(_cabi.IntegrateOpts), C.c_int64,
                                       vp, vp, vp, vp, vp, vp]
        L.oracle_dihedral_angles.restype = C.c_int
        L.oracle_dihedral_angles.argtypes = [C.POINTER(_cabi.TopologyDesc), C.c_int64, vp, vp, vp]
        L.oracle_constrain.restype = C.c_int
        L.oracle_constrain.argtypes = [C.POINTER(_cabi.TopologyDesc), C.c_int64, vp, vp]
        L.oracle_normal3.restype = None
        L.oracle_normal3.argtypes = [C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint64, vp]
        L.oracle_philox4x32_10.restype = None
        L.oracle_philox4x32_10.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, vp]
        _LIB = L
    return _LIB


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def force_opts(terms=None, cutoff=None, rfa=False, solventDielectric=78.5, switch_dist=None):
    if terms is None:
        terms = ("electrostatics", "lj", "bonds", "angles", "dihedrals", "1-4", "impropers")
    mask = 0
    for t in terms:
        mask |= _cabi.TERM_BITS[t.lower()]
    return _cabi.ForceOpts(mask, float(cutoff) if cutoff is not None else 0.0, int(bool(rfa)),
                           float(solventDielectric), float(switch_dist) if switch_dist is not None else 0.0)


def smd_opts(smd=None):
    """smd: None or dict(atom1, atom2, fk, center0, center_rate=0)."""
    if smd is None:
        return _cabi.SmdOpts(0, 0, 0, 0.0, 0.0, 0.0, None)
    return _cabi.SmdOpts(1, int(smd["atom1"]), int(smd["atom2"]), float(smd["fk"]), float(smd["center0"]),
                         float(smd.get("center_rate", 0.0)), None)


def forces(packed: "_cabi.PackedTopology", pos, box=None, smd=None, smd_centers=None, **fopts):
    """Forces.compute in FP64 on fp32 inputs.  pos (R,N,3) -> (forces (R,N,3) f64, energies (R,8) f64)."""
    pos = np.ascontiguousarray(pos, dtype=np.float32)
    R, N, _ = pos.shape
    assert N == packed.desc.natoms
    box = None if box is None else np.ascontiguousarray(box, dtype=np.float32).reshape(R, 9)
    cen = None if smd_centers is None else np.ascontiguousarray(smd_centers, dtype=np.float32)
    f = np.zeros((R, N, 3), dtype=np.float64)
    e = np.zeros((R, _cabi.NUM_ENERGIES), dtype=np.float64)
    fo, so = force_opts(**fopts), smd_opts(smd)
    rc = lib().oracle_forces(C.byref(packed.desc), C.byref(fo), C.byref(so), R, _ptr(pos), _ptr(box), _ptr(cen),
                             _ptr(f), _ptr(e))
    if rc == _cabi.ERR_NAN_COORDINATES:
        raise RuntimeError("Found NaN coordinates.")
    if rc != 0:
        raise RuntimeError(f"oracle_forces failed with status {rc}")
    return f, e


def constrain(packed, pos, vel=None):
    """b200md_constrain in FP64: returns (pos, vel) on the rigid-bond constraint surface."""
    p = np.array(pos, dtype=np.float64, order="C")
    v = None if vel is None else np.array(vel, dtype=np.float64, order="C")
    rc = lib().oracle_constrain(C.byref(packed.desc), p.shape[0], _ptr(p), _ptr(v))
    if rc != 0:
        raise RuntimeError(f"oracle_constrain failed with status {rc}")
    return p, v


def dihedral_angles(packed, pos, box=None):
    """Energy.calc_phi in FP64: pos (R, N, 3) float64 -> (R, ndihedrals) float64."""
    p = np.ascontiguousarray(pos, dtype=np.float64)
    R, N, _ = p.shape
    box = None if box is None else np.ascontiguousarray(box, dtype=np.float32).reshape(R, 9)
    out = np.zeros((R, packed.desc.ndihedrals), dtype=np.float64)
    rc = lib().oracle_dihedral_angles(C.byref(packed.desc), R, _ptr(p), _ptr(box), _ptr(out))
    if rc != 0:
        raise RuntimeError(f"oracle_dihedral_angles failed with status {rc}")
    return out


def integrate(packed, pos, vel, forces_in, nsteps, dt, gamma=0.0, temperature=0.0, seed=0, step0=0, replica0=0,
              energy_every=0, snapshot_every=0, box=None, smd=None, rigid_bonds=False, **fopts):
    """Integrator.step(nsteps) in FP64.  Returns dict(pos, vel, forces, observables, snapshots)."""
    p = np.array(pos, dtype=np.float64, order="C")
    v = np.array(vel, dtype=np.float64, order="C")
    f = np.array(forces_in, dtype=np.float64, order="C")
    R, N, _ = p.shape
    box = None if box is None else np.ascontiguousarray(box, dtype=np.float32).reshape(R, 9)
    nrep = -(-nsteps // energy_every) if energy_every > 0 else 1
    nsnap = nsteps // snapshot_every if snapshot_every > 0 else 0
    obs = np.zeros((nrep, R, _cabi.NUM_OBSERVABLES), dtype=np.float64)
    snaps = np.zeros((nsnap, R, N, 3), dtype=np.float64) if nsnap else None
    io = _cabi.IntegrateOpts(float(dt), float(gamma), float(temperature), int(seed), int(step0), int(replica0),
                             int(nsteps), int(energy_every), int(snapshot_every), int(bool(rigid_bonds)))
    fo, so = force_opts(**fopts), smd_opts(smd)
    rc = lib().oracle_integrate(C.byref(packed.desc), C.byref(fo), C.byref(so), C.byref(io), R, _ptr(p), _ptr(v),
                                _ptr(f), _ptr(box), _ptr(obs), _ptr(snaps))
    if rc != 0:
        raise RuntimeError(f"oracle_integrate failed with status {rc}")
    return {"pos": p, "vel": v, "forces": f, "observables": obs, "snapshots": snaps}


def normal3(seed, replica, atom, step):
    out = np.zeros(3, dtype=np.float64)
    lib().oracle_normal3(int(seed), int(replica), int(atom), int(step), _ptr(out))
    return out


def philox(seed, c0, c1, c2, c3):
    out = np.zeros(4, dtype=np.uint32)
    lib().oracle_philox4x32_10(int(seed), c0, c1, c2, c3, _ptr(out))
    return out
