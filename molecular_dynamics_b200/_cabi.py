"""ctypes mirror of include/b200md.h and the host-side flattening of a ``parameters`` object.

This is the only place that knows the C struct layouts.  ``load_library()`` is the product path's
single entry to native code and raises if ``libb200md.so`` has not been built -- there is no CPU
fallback behind it.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np
import torch

NUM_ENERGIES = 8
NUM_OBSERVABLES = 11
OBS_EKIN = 8
OBS_E2E_DIST = 9
OBS_SMD_WORK = 10
MAX_ATOMS = 128  # one-CTA-per-replica kernels; above it (up to MAX_ATOMS_LARGE) the large-solute kernels run
MAX_ATOMS_LARGE = 32768
NOALLOC_MAX_REPLICAS = 65536  # b200md_integrate_noalloc: size of its fixed device scratch array

# dict order of the reference's result (forces.py:40-50, :100)
ENERGY_SLOTS = ("electrostatics", "lj", "bonds", "angles", "dihedrals", "1-4", "impropers", "external")


def energy_slot(term):
    """Column of the (R, 8) energy block a term is reported in: the one enabled pair term of the lj family
    (lj / repulsion / repulsioncg, forces.py:258-316) uses the "lj" column (include/b200md.h)."""
    return ENERGY_SLOTS.index("lj" if term in ("repulsion", "repulsioncg") else term)
TERM_BITS = {"electrostatics": 1 << 0, "lj": 1 << 1, "bonds": 1 << 2, "angles": 1 << 3, "dihedrals": 1 << 4,
             "1-4": 1 << 5, "impropers": 1 << 6, "repulsion": 1 << 7, "repulsioncg": 1 << 8}

OK, ERR_INVALID_ARGUMENT, ERR_TOO_MANY_ATOMS, ERR_CUDA, ERR_NAN_COORDINATES, ERR_UNSUPPORTED = range(6)
FLAG_NAN_POS = 1
FLAG_CONSTRAINT = 2
FLAG_FORCE_RANGE = 4

_i32p, _f32p, _f64p, _u32p = (C.POINTER(C.c_int32), C.POINTER(C.c_float), C.POINTER(C.c_double),
                              C.POINTER(C.c_uint32))


class TopologyDesc(C.Structure):
    _fields_ = [
        ("natoms", C.c_int32), ("ntypes", C.c_int32),
        ("atom_type", _i32p), ("charge", _f32p), ("mass", _f32p), ("lj_a", _f32p), ("lj_b", _f32p),
        ("nbonds", C.c_int32), ("bond_idx", _i32p), ("bond_prm", _f32p),
        ("nangles", C.c_int32), ("angle_idx", _i32p), ("angle_prm", _f32p),
        ("ndihedrals", C.c_int32), ("dihedral_idx", _i32p),
        ("ndihedral_terms", C.c_int32), ("dihedral_term_owner", _i32p), ("dihedral_term_prm", _f32p),
        ("dihedral_term_harmonic", _i32p),
        ("n14", C.c_int32), ("idx14", _i32p), ("prm14", _f32p),
        ("nimpropers", C.c_int32), ("improper_idx", _i32p),
        ("nimproper_terms", C.c_int32), ("improper_term_owner", _i32p), ("improper_term_prm", _f32p),
        ("improper_term_harmonic", _i32p),
        ("npairs", C.c_int32), ("pair_idx", _i32p),
    ]


class ForceOpts(C.Structure):
    _fields_ = [("terms", C.c_uint32), ("cutoff", C.c_float), ("rfa", C.c_int32),
                ("solvent_dielectric", C.c_float), ("switch_dist", C.c_float)]


class SmdOpts(C.Structure):
    _fields_ = [("enabled", C.c_int32), ("atom1", C.c_int32), ("atom2", C.c_int32), ("fk", C.c_float),
                ("center0", C.c_float), ("center_rate", C.c_float), ("centers", C.c_void_p)]


class IntegrateOpts(C.Structure):
    _fields_ = [("dt", C.c_float), ("gamma", C.c_float), ("temperature", C.c_float), ("seed", C.c_uint64),
                ("step0", C.c_uint64), ("replica0", C.c_int64), ("nsteps", C.c_int32),
                ("energy_every", C.c_int32), ("snapshot_every", C.c_int32), ("rigid_bonds", C.c_int32)]


class FireOpts(C.Structure):
    _fields_ = [("dt_max", C.c_float), ("max_move", C.c_float), ("f_inc", C.c_float), ("f_dec", C.c_float),
                ("alpha_start", C.c_float), ("f_alpha", C.c_float), ("n_min", C.c_int32)]


# ----------------------------------------------------------------------------------------------
# flattening
# ----------------------------------------------------------------------------------------------
def _np(t, dtype):
    if t is None:
        return None
    if isinstance(t, torch.Tensor):
        t = t.detach().cpu().numpy()
    return np.ascontiguousarray(np.asarray(t), dtype=dtype)


def _flatten_torsion_slices(slices):
    """par.dihedral_params (list of {idx, params}) -> owner[], prm[], harmonic[] (forces.py:559-576).

    ``harmonic`` is decided per *slice* exactly as the reference does: a slice takes the AMBER
    branch iff ``torch.all(per > 0)`` over the whole slice, otherwise every term of that slice is
    CHARMM-harmonic."""
    owner, prm, harm = [], [], []
    for sl in slices or []:
        idx = _np(sl["idx"], np.int32).reshape(-1)
        p = _np(sl["params"], np.float32).reshape(-1, 3)
        is_harmonic = 0 if bool(np.all(p[:, 2] > 0)) else 1
        owner.append(idx)
        prm.append(p)
        harm.append(np.full(len(idx), is_harmonic, dtype=np.int32))
    if not owner:
        return (np.zeros(0, np.int32), np.zeros((0, 3), np.float32), np.zeros(0, np.int32))
    return np.concatenate(owner), np.concatenate(prm), np.concatenate(harm)


def make_pair_index(natoms, excludepairs):
    """Forces._make_indeces (forces.py:346-355): all i<j not excluded, row-major, int64 (P,2)."""
    full = np.ones((natoms, natoms), dtype=bool)
    if len(excludepairs):
        ex = np.asarray(excludepairs)
        full[ex[:, 0], ex[:, 1]] = False
        full[ex[:, 1], ex[:, 0]] = False
    full = np.triu(full, 1)
    return np.stack(np.nonzero(full), axis=1).astype(np.int64)


class PackedTopology:
    """Owns the numpy buffers a TopologyDesc points into."""

    def __init__(self, par, pair_idx):
        a = self.arrays = {}
        natoms = len(par.masses)
        a["atom_type"] = (_np(par.mapped_atom_types, np.int32) if par.mapped_atom_types is not None
                          else np.zeros(natoms, np.int32))
        a["charge"] = _np(par.charges, np.float32) if par.charges is not None else np.zeros(natoms, np.float32)
        a["mass"] = _np(par.masses, np.float32).reshape(-1)
        if par.A is not None:
            a["lj_a"], a["lj_b"] = _np(par.A, np.float32), _np(par.B, np.float32)
            ntypes = a["lj_a"].shape[0]
        else:
            ntypes = int(a["atom_type"].max()) + 1 if natoms else 1
            a["lj_a"] = np.zeros((ntypes, ntypes), np.float32)
            a["lj_b"] = np.zeros((ntypes, ntypes), np.float32)

        def idx_prm(idx, prm, wi, wp):
            if idx is None:
                return np.zeros((0, wi), np.int32), np.zeros((0, wp), np.float32)
            return _np(idx, np.int32).reshape(-1, wi), _np(prm, np.float32).reshape(-1, wp)

        a["bond_idx"], a["bond_prm"] = idx_prm(par.bonds, par.bond_params, 2, 2)
        a["angle_idx"], a["angle_prm"] = idx_prm(par.angles, par.angle_params, 3, 2)
        a["idx14"], a["prm14"] = idx_prm(par.idx14, par.nonbonded_14_params, 2, 4)
        a["dihedral_idx"] = (_np(par.dihedrals, np.int32).reshape(-1, 4) if par.dihedrals is not None
                             else np.zeros((0, 4), np.int32))
        (a["dihedral_term_owner"], a["dihedral_term_prm"],
         a["dihedral_term_harmonic"]) = _flatten_torsion_slices(par.dihedral_params if par.dihedrals is not None else None)
        a["improper_idx"] = (_np(par.impropers, np.int32).reshape(-1, 4) if par.impropers is not None
                             else np.zeros((0, 4), np.int32))
        (a["improper_term_owner"], a["improper_term_prm"],
         a["improper_term_harmonic"]) = _flatten_torsion_slices(par.improper_params if par.impropers is not None else None)
        a["pair_idx"] = (_np(pair_idx, np.int32).reshape(-1, 2) if pair_idx is not None
                         else np.zeros((0, 2), np.int32))

        d = self.desc = TopologyDesc()
        d.natoms, d.ntypes = natoms, ntypes
        d.nbonds, d.nangles, d.ndihedrals = len(a["bond_idx"]), len(a["angle_idx"]), len(a["dihedral_idx"])
        d.ndihedral_terms, d.n14 = len(a["dihedral_term_owner"]), len(a["idx14"])
        d.nimpropers, d.nimproper_terms = len(a["improper_idx"]), len(a["improper_term_owner"])
        d.npairs = len(a["pair_idx"])
        ptr_t = {np.dtype(np.int32): _i32p, np.dtype(np.float32): _f32p}
        for name, arr in a.items():
            setattr(d, name, arr.ctypes.data_as(ptr_t[arr.dtype]))


# ----------------------------------------------------------------------------------------------
# library loading
# ----------------------------------------------------------------------------------------------
_LIB = None
# B200MD_LIB selects another build of the same library (kernel experiments); default: the in-tree build
LIB_PATH = os.environ.get("B200MD_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "libb200md.so")


def declare_signatures(lib):
    vp = C.c_void_p
    lib.b200md_topology_create.restype = vp
    lib.b200md_topology_create.argtypes = [C.POINTER(TopologyDesc), C.c_int]
    lib.b200md_topology_destroy.restype = None
    lib.b200md_topology_destroy.argtypes = [vp]
    lib.b200md_topology_exclusion_mask.restype = C.c_int
    lib.b200md_topology_exclusion_mask.argtypes = [vp, vp]
    lib.b200md_topology_natoms.restype = C.c_int
    lib.b200md_topology_natoms.argtypes = [vp]
    lib.b200md_topology_mask_words.restype = C.c_int
    lib.b200md_topology_mask_words.argtypes = [vp]
    lib.b200md_topology_ndihedrals.restype = C.c_int
    lib.b200md_topology_ndihedrals.argtypes = [vp]
    lib.b200md_dihedral_angles.restype = C.c_int
    lib.b200md_dihedral_angles.argtypes = [vp, C.c_int64, vp, vp, vp, vp]
    lib.b200md_dihedral_angles_backward.restype = C.c_int
    lib.b200md_dihedral_angles_backward.argtypes = [vp, C.c_int64, vp, vp, vp, vp, vp]
    lib.b200md_constrain.restype = C.c_int
    lib.b200md_constrain.argtypes = [vp, C.c_int64, vp, vp, vp, vp]
    lib.b200md_topology_num_constraints.restype = C.c_int
    lib.b200md_topology_num_constraints.argtypes = [vp]
    lib.b200md_forces_compute.restype = C.c_int
    lib.b200md_forces_compute.argtypes = [vp, C.POINTER(ForceOpts), C.POINTER(SmdOpts), C.c_int64, vp, vp, vp, vp,
                                          vp, vp]
    lib.b200md_integrate.restype = C.c_int
    lib.b200md_integrate.argtypes = [vp, C.POINTER(ForceOpts), C.POINTER(SmdOpts), C.POINTER(IntegrateOpts),
                                     C.c_int64, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.b200md_integrate_noalloc.restype = C.c_int
    lib.b200md_integrate_noalloc.argtypes = list(lib.b200md_integrate.argtypes)
    lib.b200md_forces_compute_host.restype = C.c_int
    lib.b200md_forces_compute_host.argtypes = [vp, C.POINTER(ForceOpts), C.POINTER(SmdOpts), C.c_int64, vp, vp, vp,
                                               vp, vp]
    lib.b200md_fire_update.restype = C.c_int
    lib.b200md_fire_update.argtypes = [C.c_int, C.c_int64, C.c_int32, vp, vp, vp, vp, C.POINTER(FireOpts), vp]
    lib.b200md_fp32_peak_tflops.restype = C.c_double
    lib.b200md_fp32_peak_tflops.argtypes = [C.c_int, C.c_int]
    lib.b200md_launch_count.restype = C.c_int64
    lib.b200md_launch_count.argtypes = []
    lib.b200md_last_error.restype = C.c_char_p
    lib.b200md_last_error.argtypes = []
    lib.b200md_version.restype = C.c_char_p
    lib.b200md_version.argtypes = []
    return lib


EXPORTED_SYMBOLS = (
    "b200md_topology_create", "b200md_topology_destroy", "b200md_topology_exclusion_mask",
    "b200md_topology_natoms", "b200md_topology_mask_words", "b200md_topology_num_constraints", "b200md_topology_ndihedrals",
    "b200md_dihedral_angles", "b200md_dihedral_angles_backward", "b200md_constrain", "b200md_forces_compute", "b200md_integrate", "b200md_integrate_noalloc", "b200md_forces_compute_host",
    "b200md_fire_update", "b200md_fp32_peak_tflops", "b200md_launch_count", "b200md_last_error", "b200md_version",
)


def load_library():
    """dlopen libb200md.so (built in-tree by __graft_entry__.build()).  No fallback."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: the CUDA extension has not been built "
                "(run `python -c 'import __graft_entry__ as g; g.build()'`). There is no CPU fallback.")
        _LIB = declare_signatures(C.CDLL(LIB_PATH))
    return _LIB


def last_error() -> str:
    return load_library().b200md_last_error().decode()
