"""``Parameters``: the duck-typed parameter object ``Forces`` reads.

The reference obtains it from the un-vendored ``torchmd.parameters.Parameters``; its attribute
contract is what ``TorchMD_Fix/forces.py`` reads:

    masses (N,1)            forces.py:60         bonds, bond_params          forces.py:111-116
    device                  forces.py:64         angles, angle_params        forces.py:134-139
    get_exclusions(types)   forces.py:64         dihedrals, dihedral_params  forces.py:147-158
    charges                 forces.py:215,273    idx14, nonbonded_14_params  forces.py:176-200
    mapped_atom_types, A, B forces.py:284-286    impropers, improper_params  forces.py:228-239

The builder semantics restate upstream TorchMD 1.0.x (call site in the reference:
V_Calculations/Test-6_full_system/potcalc.py:36-48); the PSF + YAML route is pinned by the
notebook known-answer vector (V_Calculations/Test-1/tutorial.ipynb:231), the prmtop route is not.
"""
from __future__ import annotations

import math

import numpy as np
import torch

from .topology_io import PrmtopForceField, Topology, YamlForceField, read_prmtop, read_psf

ALL_TERMS = ("bonds", "angles", "dihedrals", "impropers", "1-4", "electrostatics", "lj")


def lorentz_berthelot_AB(sigma: np.ndarray, epsilon: np.ndarray):
    """A = 4 eps sigma^12, B = 4 eps sigma^6 with sigma_ij = (s_i+s_j)/2, eps_ij = sqrt(e_i e_j)."""
    sig = 0.5 * (sigma[None, :] + sigma[:, None])
    eps = np.sqrt(epsilon[None, :] * epsilon[:, None])
    s6 = sig ** 6
    return eps * 4 * (s6 * s6), eps * 4 * s6


class Parameters:
    """Container of index and parameter tensors with the reference's attribute names."""

    _INDEX = ("mapped_atom_types", "bonds", "angles", "dihedrals", "idx14", "impropers")
    _REAL = ("A", "B", "charges", "masses", "bond_params", "angle_params", "nonbonded_14_params")

    def __init__(self, precision=torch.float, device="cpu"):
        self.A = self.B = None
        self.bonds = self.bond_params = None
        self.charges = self.masses = self.mapped_atom_types = None
        self.angles = self.angle_params = None
        self.dihedrals = self.dihedral_params = None
        self.idx14 = self.nonbonded_14_params = None
        self.impropers = self.improper_params = None
        self.precision = precision
        self.device = device

    # ---- same helpers as upstream ---------------------------------------------------------
    @property
    def natoms(self):
        return len(self.masses)

    def to_(self, device):
        for name in self._INDEX + self._REAL:
            v = getattr(self, name)
            if v is not None:
                setattr(self, name, v.to(device))
        for plist in (self.dihedral_params, self.improper_params):
            if plist is not None:
                for sl in plist:
                    sl["idx"] = sl["idx"].to(device)
                    sl["params"] = sl["params"].to(device)
        self.device = device
        return self

    def precision_(self, precision):
        for name in self._REAL:
            v = getattr(self, name)
            if v is not None:
                setattr(self, name, v.type(precision))
        for plist in (self.dihedral_params, self.improper_params):
            if plist is not None:
                for sl in plist:
                    sl["params"] = sl["params"].type(precision)
        self.precision = precision
        return self

    def get_exclusions(self, types=("bonds", "angles", "1-4"), fullarray=False):
        """List of excluded [i, j] pairs: bonds, angle ends, dihedral ends (read at forces.py:64)."""
        excl = []
        if self.bonds is not None and "bonds" in types:
            excl += self.bonds.cpu().numpy().tolist()
        if self.angles is not None and "angles" in types:
            excl += self.angles.cpu().numpy()[:, [0, 2]].tolist()
        if self.dihedrals is not None and "1-4" in types:
            excl += self.dihedrals.cpu().numpy()[:, [0, 3]].tolist()
        if fullarray:
            n = self.natoms
            full = np.zeros((n, n), dtype=bool)
            if len(excl):
                e = np.array(excl)
                full[e[:, 0], e[:, 1]] = True
                full[e[:, 1], e[:, 0]] = True
            return full
        return excl

    # ---- (de)serialisation: small .npz fixtures travel to the GPU box --------------------
    def to_arrays(self) -> dict:
        out = {}
        for name in self._INDEX + self._REAL:
            v = getattr(self, name)
            if v is not None:
                out[name] = v.detach().cpu().numpy()
        for label, plist in (("dihedral", self.dihedral_params), ("improper", self.improper_params)):
            if plist is not None:
                out[f"{label}_nslices"] = np.array(len(plist))
                for j, sl in enumerate(plist):
                    out[f"{label}_idx_{j}"] = sl["idx"].cpu().numpy()
                    out[f"{label}_params_{j}"] = sl["params"].cpu().numpy()
        return out

    @classmethod
    def from_arrays(cls, arrs, precision=torch.float, device="cpu"):
        par = cls(precision=precision, device=device)
        for name in cls._INDEX + cls._REAL:
            if name in arrs:
                setattr(par, name, torch.tensor(np.asarray(arrs[name])))
        for label in ("dihedral", "improper"):
            if f"{label}_nslices" in arrs:
                plist = []
                for j in range(int(arrs[f"{label}_nslices"])):
                    plist.append({"idx": torch.tensor(np.asarray(arrs[f"{label}_idx_{j}"])),
                                  "params": torch.tensor(np.asarray(arrs[f"{label}_params_{j}"]))})
                setattr(par, f"{label}_params", plist)
        par.precision_(precision)
        par.to_(device)
        return par

    def tiled(self, copies):
        """Parameters of ``copies`` identical molecules side by side in one topology (a bundle of chains): atom
        indices of copy c are shifted by c * natoms, per-atom and per-item parameters are repeated, the type-pair tables
        are shared.  Molecules of one topology interact through the non-bonded terms like any other atoms."""
        a = self.to_arrays()
        n = len(a["masses"])
        out = {}
        for k, v in a.items():
            if k in ("A", "B") or k.endswith("_nslices"):
                out[k] = v
            elif k in ("bonds", "angles", "dihedrals", "idx14", "impropers"):
                out[k] = np.concatenate([v + c * n for c in range(copies)])
            elif k.startswith(("dihedral_idx_", "improper_idx_")):
                rows = len(a["dihedrals" if k.startswith("dihedral") else "impropers"])
                out[k] = np.concatenate([v + c * rows for c in range(copies)])
            else:  # per-atom arrays, per-item parameter rows, per-slice parameter rows
                out[k] = np.concatenate([v] * copies)
        return Parameters.from_arrays(out, precision=self.precision, device=self.device)

    def save(self, path):
        np.savez_compressed(path, **self.to_arrays())

    @classmethod
    def load(cls, path, precision=torch.float, device="cpu"):
        with np.load(path) as z:
            return cls.from_arrays({k: z[k] for k in z.files}, precision=precision, device=device)


# ----------------------------------------------------------------------------------------------
# builders
# ----------------------------------------------------------------------------------------------
def _unique_topology(top: Topology):
    uqbonds = np.unique([sorted(b) for b in top.bonds], axis=0) if len(top.bonds) else None
    uqangles = (np.unique([a if a[0] < a[2] else a[::-1] for a in top.angles.tolist()], axis=0)
                if len(top.angles) else None)
    uqdih = (np.unique([d if d[0] < d[3] else d[::-1] for d in top.dihedrals.tolist()], axis=0)
             if len(top.dihedrals) else None)
    uqimp = np.unique(top.impropers, axis=0) if top.impropers is not None and len(top.impropers) else None
    return uqbonds, uqangles, uqdih, uqimp


def _slices_from_terms(term_lists):
    """term_lists[i] = [(k, phase, per), ...] -> list over term rank j of {idx, params}."""
    nslices = max(len(t) for t in term_lists)
    out = []
    for j in range(nslices):
        idx = [i for i, t in enumerate(term_lists) if len(t) > j]
        prm = [list(map(float, term_lists[i][j])) for i in idx]
        out.append({"idx": torch.tensor(idx, dtype=torch.int64), "params": torch.tensor(prm, dtype=torch.float64)})
    return out


def _select_14(par: Parameters, uqdih):
    excl = par.get_exclusions(types=("bonds", "angles"), fullarray=True)
    keep = ~excl[uqdih[:, 0], uqdih[:, 3]]
    dih14 = uqdih[keep]
    if not len(dih14):
        return None
    first = np.unique(dih14[:, [0, 3]], axis=0, return_index=True)[1]
    return dih14[first]


def _improper_centre(quad, bonds):
    nb = {int(a): set() for a in quad}
    for i, j in bonds:
        if int(i) in nb and int(j) in nb:
            nb[int(i)].add(int(j))
            nb[int(j)].add(int(i))
    for a, s in nb.items():
        if len(s) == 3:
            return a
    raise RuntimeError(f"no central atom found for improper {quad}")


def build_parameters_yaml(top: Topology, ff: YamlForceField, terms=ALL_TERMS, precision=torch.float,
                          device="cpu") -> Parameters:
    """PSF topology + YAML force field -> Parameters (the route the known-answer vector pins)."""
    terms = [t.lower() for t in terms]
    par = Parameters()
    uqtypes, inv = np.unique(top.atomtype.astype(str), return_inverse=True)
    par.mapped_atom_types = torch.tensor(inv.astype(np.int64))
    if "electrostatics" in terms:
        par.charges = torch.tensor(top.charge.astype(np.float64))  # PSF column, not the FF charge
    par.masses = torch.tensor([ff.mass(t) for t in top.atomtype], dtype=torch.float64).unsqueeze(1)
    if "lj" in terms:
        sig = np.array([ff.lj(t)[0] for t in uqtypes], dtype=np.float64)
        eps = np.array([ff.lj(t)[1] for t in uqtypes], dtype=np.float64)
        A, B = lorentz_berthelot_AB(sig, eps)
        par.A, par.B = torch.tensor(A), torch.tensor(B)
    uqbonds, uqangles, uqdih, uqimp = _unique_topology(top)
    tname = lambda rows: uqtypes[inv[rows]]
    if "bonds" in terms and uqbonds is not None:
        par.bonds = torch.tensor(uqbonds.astype(np.int64))
        par.bond_params = torch.tensor([ff.bond(*t) for t in tname(uqbonds)], dtype=torch.float64)
    if "angles" in terms and uqangles is not None:
        par.angles = torch.tensor(uqangles.astype(np.int64))
        par.angle_params = torch.tensor([ff.angle(*t) for t in tname(uqangles)], dtype=torch.float64)
    if "dihedrals" in terms and uqdih is not None:
        par.dihedrals = torch.tensor(uqdih.astype(np.int64))
        par.dihedral_params = _slices_from_terms([ff.dihedral_terms(*t) for t in tname(uqdih)])
    if "1-4" in terms and uqdih is not None:
        dih14 = _select_14(par, uqdih)
        if dih14 is not None:
            par.idx14 = torch.tensor(dih14[:, [0, 3]].astype(np.int64))
            rows = []
            for t in tname(dih14):
                scnb, scee, s1, e1, s4, e4 = ff.one_four(*t)
                sig, eps = 0.5 * (s1 + s4), math.sqrt(e1 * e4)
                s6 = sig ** 6
                rows.append([eps * 4 * s6 * s6, eps * 4 * s6, scnb, scee])
            par.nonbonded_14_params = torch.tensor(rows, dtype=torch.float64)
    if "impropers" in terms and uqimp is not None:
        par.impropers = torch.tensor(uqimp.astype(np.int64))
        prm = []
        for quad in uqimp:
            try:
                prm.append(ff.improper(*uqtypes[inv[quad]]))
            except RuntimeError:
                c = _improper_centre(quad, uqbonds)
                rest = sorted(int(a) for a in quad if int(a) != c)
                order = [rest[0], rest[1], c, rest[2]]
                prm.append(ff.improper(*uqtypes[inv[order]]))
        par.improper_params = [{"idx": torch.arange(len(uqimp), dtype=torch.int64),
                                "params": torch.tensor(prm, dtype=torch.float64)}]
    return par.precision_(precision).to_(device)


def build_parameters_prmtop(top: Topology, ff: PrmtopForceField, terms=ALL_TERMS, precision=torch.float,
                            device="cpu", charmm_impropers=False, urey_bradley=False) -> Parameters:
    """prmtop topology + its own parameter tables -> Parameters (parameters read by index).

    The two CHAMBER-only term families are OFF by default because the reference's route drops them silently
    (the 22 harmonic impropers and 65 Urey-Bradley terms of the ALA10 prmtop never reach
    ``Forces``); ``charmm_impropers=True`` adds them as a periodicity-0 improper slice (the reference's CHARMM
    branch, forces.py:570-576) and ``urey_bradley=True`` as extra rows of the bond list.  The exclusion set does
    not change: Urey-Bradley pairs are the end atoms of angles, already excluded as 1-3."""
    terms = [t.lower() for t in terms]
    par = Parameters()
    # LJ types: the prmtop's own ATOM_TYPE_INDEX (A/B tables are given over these)
    par.mapped_atom_types = torch.tensor(ff.type_index.copy())
    if "electrostatics" in terms:
        par.charges = torch.tensor(top.charge.astype(np.float64))
    par.masses = torch.tensor(top.masses.astype(np.float64)).unsqueeze(1)
    if "lj" in terms:
        A, B = ff.lj_tables()
        par.A, par.B = torch.tensor(A), torch.tensor(B)
    uqbonds, uqangles, uqdih, uqimp = _unique_topology(top)
    if "bonds" in terms and uqbonds is not None:
        par.bonds = torch.tensor(uqbonds.astype(np.int64))
        par.bond_params = torch.tensor([ff.bond(*r) for r in uqbonds], dtype=torch.float64)
    if "angles" in terms and uqangles is not None:
        par.angles = torch.tensor(uqangles.astype(np.int64))
        par.angle_params = torch.tensor([ff.angle(*r) for r in uqangles], dtype=torch.float64)
    if "dihedrals" in terms and uqdih is not None:
        par.dihedrals = torch.tensor(uqdih.astype(np.int64))
        par.dihedral_params = _slices_from_terms([ff.dihedral_terms(r) for r in uqdih])
    if "1-4" in terms and uqdih is not None:
        dih14 = _select_14(par, uqdih)
        if dih14 is not None:
            par.idx14 = torch.tensor(dih14[:, [0, 3]].astype(np.int64))
            A14, B14 = ff.lj_tables(fourteen=True)
            rows = []
            for r in dih14:
                ti, tj = ff.type_index[r[0]], ff.type_index[r[3]]
                scnb, scee = ff.dihedral_scale14(r)
                rows.append([A14[ti, tj], B14[ti, tj], scnb, scee])
            par.nonbonded_14_params = torch.tensor(rows, dtype=torch.float64)
    if "impropers" in terms and uqimp is not None:
        par.impropers = torch.tensor(uqimp.astype(np.int64))
        par.improper_params = [{"idx": torch.arange(len(uqimp), dtype=torch.int64),
                                "params": torch.tensor([ff.improper_term(r) for r in uqimp], dtype=torch.float64)}]
    if urey_bradley and "bonds" in terms:
        from .topology_io import chamber_urey_bradley

        pairs, prm = chamber_urey_bradley(ff.sec)
        if len(pairs):
            have = par.bonds is not None
            par.bonds = torch.cat([par.bonds, torch.tensor(pairs)]) if have else torch.tensor(pairs)
            par.bond_params = torch.cat([par.bond_params, torch.tensor(prm)]) if have else torch.tensor(prm)
    if charmm_impropers and "impropers" in terms:
        from .topology_io import chamber_impropers

        quads, prm = chamber_impropers(ff.sec)
        if len(quads):
            n0 = 0 if par.impropers is None else len(par.impropers)
            sl = {"idx": torch.arange(n0, n0 + len(quads), dtype=torch.int64), "params": torch.tensor(prm)}
            par.impropers = torch.tensor(quads) if n0 == 0 else torch.cat([par.impropers, torch.tensor(quads)])
            par.improper_params = [sl] if n0 == 0 else list(par.improper_params) + [sl]
    return par.precision_(precision).to_(device)


def parameters_from_psf_yaml(psf_path, yaml_path, **kw) -> Parameters:
    return build_parameters_yaml(read_psf(psf_path), YamlForceField(yaml_path), **kw)


def parameters_from_prmtop(prmtop_path, **kw) -> Parameters:
    top, ff = read_prmtop(prmtop_path)
    return build_parameters_prmtop(top, ff, **kw)
