"""File readers for the topology / coordinate formats the reference's force-field path consumes.

Host-side, one-time work.  The reference never vendors these
readers in a standalone form (it relies on moleculekit / parmed / torchmd, all absent), so they are
restated here from the format as used by the reference:

* X-PLOR PSF            -> reference V_Calculations/Test-6_full_system/readers.py:1402-1482
* AMBER / CHAMBER prmtop -> reference V_Calculations/Test-6_full_system/readers.py:1240-1399
  (including the CHAMBER charge-field patch at :1295-1303, i.e. 24-wide fields / 18.2223)
* YAML force field look-up with ``X`` wild-cards -> reference
  V_Calculations/Test-3_energy_module/ff_yaml.py:12-108
* PDB fixed columns, NAMD binary ``.coor`` / ``.vel`` (int32 N + N*3 float64)

Nothing here touches the GPU.
"""
from __future__ import annotations

import itertools
import math
import re
import struct
from dataclasses import dataclass, field

import numpy as np

__all__ = [
    "Topology",
    "read_psf",
    "read_pdb_coords",
    "read_namd_bin",
    "read_prmtop",
    "YamlForceField",
    "PrmtopForceField",
    "NAMD_VEL_TO_ANGSTROM_PER_PS",
]

# NAMD stores velocities in its internal unit; x 20.45482706 gives Angstrom / ps.
NAMD_VEL_TO_ANGSTROM_PER_PS = 20.45482706


@dataclass
class Topology:
    """Plain topology record: what ``moleculekit.Molecule`` exposes to ``torchmd.Parameters``."""

    name: list = field(default_factory=list)
    atomtype: np.ndarray | None = None  # (N,) str
    charge: np.ndarray | None = None  # (N,) f64
    masses: np.ndarray | None = None  # (N,) f64
    bonds: np.ndarray | None = None  # (nb, 2) int
    angles: np.ndarray | None = None  # (na, 3) int
    dihedrals: np.ndarray | None = None  # (nd, 4) int
    impropers: np.ndarray | None = None  # (ni, 4) int

    @property
    def natoms(self) -> int:
        return len(self.atomtype)


def _int_rows(tokens, width):
    arr = np.asarray(tokens, dtype=np.int64)
    if arr.size % width:
        raise ValueError(f"index section length {arr.size} is not a multiple of {width}")
    return arr.reshape(-1, width)


def read_psf(path: str) -> Topology:
    """X-PLOR PSF reader (sections NATOM / NBOND / NTHETA / NPHI / NIMPHI; 1-based -> 0-based)."""
    sections = {"!NATOM": "atoms", "!NBOND": "bonds", "!NTHETA": "angles", "!NPHI": "dihedrals",
                "!NIMPHI": "impropers"}
    width = {"bonds": 2, "angles": 3, "dihedrals": 4, "impropers": 4}
    mode = None
    names, types, charges, masses = [], [], [], []
    toks = {k: [] for k in width}
    with open(path) as fh:
        for line in fh:
            stripped = line.strip()
            if not stripped:
                mode = None
                continue
            hit = [v for k, v in sections.items() if k in line]
            if hit:
                mode = hit[0]
                continue
            if "!" in line and re.search(r"!N[A-Z]+", line):
                mode = None  # a section we do not consume (NDON, NACC, NNB, NGRP, NCRTERM ...)
                continue
            if mode == "atoms":
                f = stripped.split()
                # serial segid resid resname name type charge mass [...]
                names.append(f[4])
                types.append(f[5])
                charges.append(float(f[6]))
                masses.append(float(f[7]))
            elif mode in toks:
                toks[mode] += [int(x) - 1 for x in stripped.split()]
    top = Topology(name=names, atomtype=np.array(types, dtype=object),
                   charge=np.array(charges, dtype=np.float64), masses=np.array(masses, dtype=np.float64))
    for k, w in width.items():
        setattr(top, k, _int_rows(toks[k], w))
    return top


def read_pdb_coords(path: str) -> np.ndarray:
    """ATOM/HETATM fixed columns 31-38 / 39-46 / 47-54 -> (N, 3) float64."""
    xyz = []
    with open(path) as fh:
        for line in fh:
            if line.startswith(("ATOM", "HETATM")):
                xyz.append([float(line[30:38]), float(line[38:46]), float(line[46:54])])
    return np.asarray(xyz, dtype=np.float64)


def read_namd_bin(path: str) -> np.ndarray:
    """NAMD binary coordinate / velocity file: little-endian int32 N, then N*3 float64.  Velocities are in NAMD's
    internal unit, A per 48.88821 fs -- the same TIMEFACTOR the integrator uses, so a ``.vel`` file drops into
    ``System.vel`` unchanged."""
    with open(path, "rb") as fh:
        raw = fh.read()
    (n,) = struct.unpack("<i", raw[:4])
    if len(raw) != 4 + 24 * n:
        raise ValueError(f"{path}: size {len(raw)} does not match {n} atoms")
    return np.frombuffer(raw, dtype="<f8", offset=4).reshape(n, 3).copy()


def read_xsc(path: str) -> dict:
    """NAMD extended-system file -> {"step", "box" (3, 3) rows a, b, c (zeros in vacuum), "origin" (3,)}.  The
    ``#$LABELS`` line names the columns; the reference's vacuum runs write only ``step o_x o_y o_z``."""
    labels, values = None, None
    with open(path) as fh:
        for line in fh:
            if line.startswith("#$LABELS"):
                labels = line.split()[1:]
            elif line.strip() and not line.startswith("#"):
                values = [float(x) for x in line.split()]
    if labels is None or values is None or len(values) < len(labels):
        raise ValueError(f"{path}: not an xsc file")
    rec = dict(zip(labels, values))
    box = np.array([[rec.get(f"{v}_{c}", 0.0) for c in "xyz"] for v in "abc"], dtype=np.float64)
    return {"step": int(rec["step"]), "box": box, "origin": np.array([rec.get(f"o_{c}", 0.0) for c in "xyz"])}


# ----------------------------------------------------------------------------------------------
# prmtop
# ----------------------------------------------------------------------------------------------
_FMT = re.compile(r"%FORMAT\(\s*(\d*)\s*([aAiIeEfF])\s*(\d+)(?:\.(\d+))?\s*\)")


def _prmtop_sections(path: str) -> dict:
    """Split a prmtop into {FLAG: list of typed values} using the %FORMAT fixed-width fields."""
    out, flag, kind, width = {}, None, None, None
    with open(path) as fh:
        for line in fh:
            if line.startswith("%VERSION") or line.startswith("%COMMENT"):
                continue
            if line.startswith("%FLAG"):
                flag = line.split()[1]
                out[flag] = []
                kind = None
                continue
            if line.startswith("%FORMAT"):
                m = _FMT.match(line.strip())
                if m is None:  # e.g. (i2,a78) of FORCE_FIELD_TYPE: keep raw strings
                    kind, width = "raw", None
                else:
                    kind, width = m.group(2).lower(), int(m.group(3))
                continue
            if flag is None:
                continue
            body = line.rstrip("\n")
            if kind == "raw":
                out[flag].append(body)
                continue
            for i in range(0, len(body), width):
                tok = body[i:i + width]
                if not tok.strip():
                    continue
                if kind == "a":
                    out[flag].append(tok.strip())
                elif kind == "i":
                    out[flag].append(int(tok))
                else:
                    out[flag].append(float(tok))
    return out


# Charge scaling used by the reference reader (readers.py:1295-1303): prmtop stores q * 18.2223.
PRMTOP_CHARGE_SCALE = 18.2223


def read_prmtop(path: str):
    """AMBER / CHAMBER prmtop -> (Topology, PrmtopForceField).

    Index decoding follows readers.py:1373-1395: stored atom indices are 3*index; a negative third
    dihedral atom means "skip the 1-4 / extra term of a multi-term dihedral"; a negative fourth
    means improper.  Parameters are read *by parameter index* from the prmtop tables (the reference
    goes through parmed's by-type tables; for these files the two are equivalent -- parity for this
    loader is "unpinned").
    """
    sec = _prmtop_sections(path)
    natom, ntypes = sec["POINTERS"][0], sec["POINTERS"][1]
    top = Topology()
    top.name = list(sec["ATOM_NAME"])
    top.atomtype = np.array(sec["AMBER_ATOM_TYPE"], dtype=object)
    top.charge = np.array(sec["CHARGE"], dtype=np.float64) / PRMTOP_CHARGE_SCALE
    top.masses = np.array(sec["MASS"], dtype=np.float64)
    assert len(top.atomtype) == natom

    def rows(flagh, flaga, w):
        return np.concatenate([_int_rows(sec.get(flagh, []), w), _int_rows(sec.get(flaga, []), w)])

    b = rows("BONDS_INC_HYDROGEN", "BONDS_WITHOUT_HYDROGEN", 3)
    a = rows("ANGLES_INC_HYDROGEN", "ANGLES_WITHOUT_HYDROGEN", 4)
    d = rows("DIHEDRALS_INC_HYDROGEN", "DIHEDRALS_WITHOUT_HYDROGEN", 5)
    top.bonds = b[:, :2] // 3
    top.angles = a[:, :3] // 3
    dat = np.stack([d[:, 0] // 3, d[:, 1] // 3, np.abs(d[:, 2]) // 3, np.abs(d[:, 3]) // 3], axis=1)
    is_impr = d[:, 3] < 0
    top.dihedrals = dat[~is_impr]
    top.impropers = dat[is_impr]

    ff = PrmtopForceField(sec, top, bond_rows=b, angle_rows=a, dihedral_rows=d)
    return top, ff


def chamber_urey_bradley(sec):
    """CHAMBER ``CHARMM_UREY_BRADLEY`` -> (pairs (n, 2) 0-based, params (n, 2) = (k, r0)): harmonic 1-3 terms
    ``k (r - r0)^2``, i.e. extra rows of the bond list.  Atom indices in the CHARMM sections are
    plain 1-based, not the 3*index coordinate offsets of the AMBER sections."""
    rows = np.array(sec.get("CHARMM_UREY_BRADLEY", []), dtype=np.int64).reshape(-1, 3)
    k = np.array(sec.get("CHARMM_UREY_BRADLEY_FORCE_CONSTANT", []), dtype=np.float64)
    r0 = np.array(sec.get("CHARMM_UREY_BRADLEY_EQUIL_VALUE", []), dtype=np.float64)
    return rows[:, :2] - 1, np.stack([k[rows[:, 2] - 1], r0[rows[:, 2] - 1]], axis=1)


def chamber_impropers(sec):
    """CHAMBER ``CHARMM_IMPROPERS`` -> (quads (n, 4) 0-based, params (n, 3) = (k, psi0, 0)): harmonic impropers
    ``k (psi - psi0)^2``; periodicity 0 selects the reference's CHARMM branch (forces.py:570-576)."""
    rows = np.array(sec.get("CHARMM_IMPROPERS", []), dtype=np.int64).reshape(-1, 5)
    k = np.array(sec.get("CHARMM_IMPROPER_FORCE_CONSTANT", []), dtype=np.float64)
    p0 = np.array(sec.get("CHARMM_IMPROPER_PHASE", []), dtype=np.float64)
    return rows[:, :4] - 1, np.stack([k[rows[:, 4] - 1], p0[rows[:, 4] - 1], np.zeros(len(rows))], axis=1)


class PrmtopForceField:
    """Per-index parameter look-ups on a parsed prmtop (bond/angle/dihedral by topology row)."""

    def __init__(self, sec, top, bond_rows, angle_rows, dihedral_rows):
        self.sec = sec
        self.top = top
        self.ntypes = sec["POINTERS"][1]
        self.type_index = np.array(sec["ATOM_TYPE_INDEX"], dtype=np.int64) - 1  # (N,) 0-based LJ type
        self._bond = {}
        for i, j, p in bond_rows:
            self._bond[tuple(sorted((i // 3, j // 3)))] = p - 1
        self._angle = {}
        for i, j, k, p in angle_rows:
            key = (i // 3, j // 3, k // 3)
            self._angle[key if key[0] < key[2] else key[::-1]] = p - 1
        # proper dihedrals: keep file order of the terms of each oriented quad; impropers likewise
        self._dihedral, self._improper = {}, {}
        for i, j, k, l, p in dihedral_rows:
            key = (i // 3, j // 3, abs(k) // 3, abs(l) // 3)
            if l < 0:
                self._improper.setdefault(key, []).append(p - 1)
            else:
                key = key if key[0] < key[3] else key[::-1]
                self._dihedral.setdefault(key, []).append(p - 1)

    # LJ tables over the prmtop's own type indices -------------------------------------------
    def lj_tables(self, fourteen=False):
        nt = self.ntypes
        idx = np.array(self.sec["NONBONDED_PARM_INDEX"], dtype=np.int64).reshape(nt, nt) - 1
        suffix = "_14" if fourteen and "LENNARD_JONES_14_ACOEF" in self.sec else ""
        acoef = np.array(self.sec[f"LENNARD_JONES{suffix}_ACOEF"], dtype=np.float64)
        bcoef = np.array(self.sec[f"LENNARD_JONES{suffix}_BCOEF"], dtype=np.float64)
        if np.any(idx < 0):
            raise NotImplementedError("10-12 hydrogen-bond LJ entries are not supported")
        return acoef[idx], bcoef[idx]

    def bond(self, i, j):
        p = self._bond[tuple(sorted((int(i), int(j))))]
        return self.sec["BOND_FORCE_CONSTANT"][p], self.sec["BOND_EQUIL_VALUE"][p]

    def angle(self, i, j, k):
        key = (int(i), int(j), int(k))
        p = self._angle[key if key[0] < key[2] else key[::-1]]
        return self.sec["ANGLE_FORCE_CONSTANT"][p], self.sec["ANGLE_EQUIL_VALUE"][p]

    def dihedral_terms(self, quad):
        key = tuple(int(x) for x in quad)
        key = key if key[0] < key[3] else key[::-1]
        return [(self.sec["DIHEDRAL_FORCE_CONSTANT"][p], self.sec["DIHEDRAL_PHASE"][p],
                 self.sec["DIHEDRAL_PERIODICITY"][p]) for p in self._dihedral[key]]

    def dihedral_scale14(self, quad):
        key = tuple(int(x) for x in quad)
        key = key if key[0] < key[3] else key[::-1]
        p = self._dihedral[key][0]
        scee = self.sec.get("SCEE_SCALE_FACTOR")
        scnb = self.sec.get("SCNB_SCALE_FACTOR")
        # AMBER defaults when the sections are absent: 1.2 / 2.0
        return (scnb[p] if scnb else 2.0), (scee[p] if scee else 1.2)

    def improper_term(self, quad):
        p = self._improper[tuple(int(x) for x in quad)][0]
        return (self.sec["DIHEDRAL_FORCE_CONSTANT"][p], self.sec["DIHEDRAL_PHASE"][p],
                self.sec["DIHEDRAL_PERIODICITY"][p])


# ----------------------------------------------------------------------------------------------
# YAML force field (by atom-type names, with "X" wild-cards)
# ----------------------------------------------------------------------------------------------
class YamlForceField:
    """Restates the look-up rules of the reference's patched ``YamlForcefield`` (ff_yaml.py:12-108).

    Key strings are ``"(A, B, C)"`` (or bare ``"A"`` for single types); a term matches the first
    variant found when variants are ordered by number of ``X`` wild-cards, forward order first, then
    reversed order (bonds / angles / dihedrals) or the centre-preserving permutations (impropers).
    """

    def __init__(self, path: str):
        import yaml

        with open(path) as fh:
            self.prm = yaml.load(fh, Loader=yaml.FullLoader)

    @staticmethod
    def _x_variants(types):
        types = np.array(types, dtype=object)
        masks = sorted(itertools.product([False, True], repeat=len(types)), key=sum)
        out = []
        for m in masks:
            t = types.copy()
            t[np.array(m, dtype=bool)] = "X"
            out.append(t)
        return out

    def lookup(self, term, types):
        types = np.array(types, dtype=object)
        variants = self._x_variants(types)
        if term in ("bonds", "angles", "dihedrals"):
            variants += self._x_variants(types[::-1])
        elif term == "impropers":
            for perm in itertools.permutations((0, 1, 2, 3)):
                if perm[2] == 2:
                    variants += self._x_variants(types[list(perm)])
        variants = sorted(variants, key=lambda v: sum(v == "X"))  # stable, like the reference
        table = self.prm[term]
        for v in variants:
            key = ", ".join(v)
            if len(v) > 1:
                key = "(" + key + ")"
            if key in table:
                return table[key]
        raise RuntimeError(f"{types} doesn't have {term} information in the FF")

    def mass(self, at):
        return self.prm["masses"][at]

    def lj(self, at):
        p = self.lookup("lj", [at])
        return p["sigma"], p["epsilon"]

    def bond(self, a, b):
        p = self.lookup("bonds", [a, b])
        return p["k0"], p["req"]

    def angle(self, a, b, c):
        p = self.lookup("angles", [a, b, c])
        return p["k0"], math.radians(p["theta0"])

    def dihedral_terms(self, a, b, c, d):
        p = self.lookup("dihedrals", [a, b, c, d])
        return [(t["phi_k"], math.radians(t["phase"]), t["per"]) for t in p["terms"].values()]

    def one_four(self, a, b, c, d):
        p = self.lookup("dihedrals", [a, b, c, d])
        lj1, lj4 = self.lookup("lj", [a]), self.lookup("lj", [d])
        return (p.get("scnb", 1), p.get("scee", 1), lj1["sigma14"], lj1["epsilon14"],
                lj4["sigma14"], lj4["epsilon14"])

    def improper(self, a, b, c, d):
        p = self.lookup("impropers", [a, b, c, d])
        return p["phi_k"], math.radians(p["phase"]), p["per"]
