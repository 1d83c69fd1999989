"""torch-CPU restatement of the reference's hot path with the reference's cost profile.

TEST / BASELINE INFRASTRUCTURE ONLY (see oracle/__init__.py).  /root/reference cannot travel to the GPU
box, so the timed CPU baseline ("port") is this file: it evaluates the same terms with the same kind of
torch ops, one Python iteration per replica (the reference's serial loop, TorchMD_Fix/forces.py:104), and
drives them with the restated upstream velocity-Verlet/Langevin loop.  It is written
against the flattened tables of ``_cabi.PackedTopology`` rather than the ``parameters`` object.

Follows: forces.py:111-132 (bonds), :134-145/:504-537 (angles), :147-174/:540-602 (torsions), :176-226
(1-4), :258-316 (non-bonded), :379-413 (LJ + switch), :451-489 (Coulomb / reaction field).
Pinned in tests/test_oracle.py against the golden fixtures (fp64: 1e-10) and the notebook vector.
"""
from __future__ import annotations

import math

import numpy as np
import torch

ELEC_FACTOR = 332.06371307417066
BOLTZMAN = 0.001987191
TIMEFACTOR = 48.88821


class PortTopology:
    """Index / parameter tensors in ``dtype`` built from a PackedTopology's numpy arrays."""

    def __init__(self, packed, dtype=torch.float32):
        a = packed.arrays
        t = lambda x: torch.as_tensor(np.asarray(x)).to(dtype)
        i = lambda x: torch.as_tensor(np.asarray(x)).long()
        self.dtype = dtype
        self.natoms = packed.desc.natoms
        self.mass = t(a["mass"]).reshape(-1, 1)
        self.charge, self.atype = t(a["charge"]), i(a["atom_type"])
        self.A, self.B = t(a["lj_a"]), t(a["lj_b"])
        self.bonds, self.bond_prm = i(a["bond_idx"]), t(a["bond_prm"])
        self.angles, self.angle_prm = i(a["angle_idx"]), t(a["angle_prm"])
        self.idx14, self.prm14 = i(a["idx14"]), t(a["prm14"])
        self.pairs = i(a["pair_idx"])
        self.torsions = {}
        for name in ("dihedral", "improper"):
            self.torsions[name] = (i(a[f"{name}_idx"]), i(a[f"{name}_term_owner"]), t(a[f"{name}_term_prm"]),
                                   torch.as_tensor(a[f"{name}_term_harmonic"]).bool())


def _dist(pos, idx):
    d = pos[idx[:, 0]] - pos[idx[:, 1]]
    r = torch.linalg.norm(d, dim=1)
    return r, d / r[:, None], d


def _lj(r, aa, bb, scale, rs, rc):
    ri = 1 / r
    r6 = ri ** 6
    r12 = r6 * r6
    pot = (aa * r12 - bb * r6) / scale
    frc = (-12 * aa * r12 + 6 * bb * r6) * ri / scale
    if rs is not None and rc is not None:
        m = r > rs
        t = (r[m] - rs) / (rc - rs)
        sv = 1 + t * t * t * (-10 + t * (15 - t * 6))
        sd = t * t * (-30 + t * (60 - t * 30)) / (rc - rs)
        frc[m] = sv * frc[m] + pot[m] * sd / r[m]
        pot[m] = pot[m] * sv
    return pot, frc


def _coulomb(r, qi, qj, scale, rfa, rc, eps):
    if rfa:
        den = 2 * eps + 1
        krf, crf = (1 / rc ** 3) * (eps - 1) / den, (1 / rc) * (3 * eps) / den
        common = ELEC_FACTOR * qi * qj / scale
        return common * (1 / r + krf * r ** 2 - crf), common * (2 * krf * r - 1 / r ** 2)
    pot = ELEC_FACTOR * qi * qj / r / scale
    return pot, -pot / r


def _pair_scatter(f, idx, unit, coeff):
    fv = unit * coeff[:, None]
    f.index_add_(0, idx[:, 0], -fv)
    f.index_add_(0, idx[:, 1], fv)


def _torsion(pos, f, idx, owner, prm, harmonic):
    r12, r23, r34 = pos[idx[:, 0]] - pos[idx[:, 1]], pos[idx[:, 1]] - pos[idx[:, 2]], pos[idx[:, 2]] - pos[idx[:, 3]]
    cA, cB = torch.linalg.cross(r12, r23), torch.linalg.cross(r23, r34)
    cC = torch.linalg.cross(r23, cA)
    nA, nB, nC = (torch.linalg.norm(v, dim=1) for v in (cA, cB, cC))
    ncB = cB / nB[:, None]
    phi = -torch.atan2((cC * ncB).sum(1) / nC, (cA * ncB).sum(1) / nA)
    pot = torch.zeros_like(phi)
    coeff = torch.zeros_like(phi)
    k0, phi0, per = prm[:, 0], prm[:, 1], prm[:, 2]
    amber = ~harmonic
    if amber.any():
        ad = per[amber] * phi[owner[amber]] - phi0[amber]
        pot.scatter_add_(0, owner[amber], k0[amber] * (1 + torch.cos(ad)))
        coeff.scatter_add_(0, owner[amber], -per[amber] * k0[amber] * torch.sin(ad))
    if harmonic.any():
        ad = phi[owner[harmonic]] - phi0[harmonic]
        ad = torch.where(ad < -math.pi, ad + 2 * math.pi, ad)
        ad = torch.where(ad > math.pi, ad - 2 * math.pi, ad)
        pot.scatter_add_(0, owner[harmonic], k0[harmonic] * ad ** 2)
        coeff.scatter_add_(0, owner[harmonic], 2 * k0[harmonic] * ad)
    nd = torch.linalg.norm(r23, dim=1)
    ff0, ff3 = (-coeff * nd) / nA ** 2, (coeff * nd) / nB ** 2
    ff1, ff2 = (r12 * r23).sum(1) / nd ** 2, (r34 * r23).sum(1) / nd ** 2
    f0, f3 = ff0[:, None] * cA, ff3[:, None] * cB
    s = ff1[:, None] * f0 - ff2[:, None] * f3
    f.index_add_(0, idx[:, 0], -f0)
    f.index_add_(0, idx[:, 1], f0 + s)
    f.index_add_(0, idx[:, 2], f3 - s)
    f.index_add_(0, idx[:, 3], -f3)
    return pot.sum()


def compute(top: PortTopology, pos, forces, cutoff=None, rfa=False, solventDielectric=78.5, switch_dist=None,
            smd=None):
    """Forces.compute over (R,N,3) ``pos``: fills ``forces`` in place, returns (R,8) energies."""
    R = pos.shape[0]
    if torch.isnan(pos).any():
        raise RuntimeError("Found NaN coordinates.")
    ene = torch.zeros(R, 8, dtype=pos.dtype)
    forces.zero_()
    for i in range(R):  # serial Python loop over replicas, like forces.py:104
        p, f = pos[i], forces[i]
        if len(top.bonds):
            r, unit, _ = _dist(p, top.bonds)
            idx, prm = top.bonds, top.bond_prm
            if cutoff is not None:
                keep = r <= cutoff
                r, unit, idx, prm = r[keep], unit[keep], idx[keep], prm[keep]
            x = r - prm[:, 1]
            ene[i, 2] += (prm[:, 0] * x ** 2).sum()
            _pair_scatter(f, idx, unit, 2 * prm[:, 0] * x)
        if len(top.angles):
            a = top.angles
            r21, r23 = p[a[:, 0]] - p[a[:, 1]], p[a[:, 2]] - p[a[:, 1]]
            n21i, n23i = 1 / torch.linalg.norm(r21, dim=1), 1 / torch.linalg.norm(r23, dim=1)
            ct = torch.clamp((r23 * r21).sum(1) * n21i * n23i, -1, 1)
            dth = torch.acos(ct) - top.angle_prm[:, 1]
            ene[i, 3] += (top.angle_prm[:, 0] * dth * dth).sum()
            st = torch.sqrt(1.0 - ct * ct)
            coef = torch.where(st != 0, -2.0 * top.angle_prm[:, 0] * dth / st, torch.zeros_like(st))
            f0 = coef[:, None] * (ct[:, None] * r21 * n21i[:, None] - r23 * n23i[:, None]) * n21i[:, None]
            f2 = coef[:, None] * (ct[:, None] * r23 * n23i[:, None] - r21 * n21i[:, None]) * n23i[:, None]
            f.index_add_(0, a[:, 0], f0)
            f.index_add_(0, a[:, 1], -(f0 + f2))
            f.index_add_(0, a[:, 2], f2)
        d_idx, d_owner, d_prm, d_h = top.torsions["dihedral"]
        if len(d_idx):
            ene[i, 4] += _torsion(p, f, d_idx, d_owner, d_prm, d_h)
        if len(top.idx14):
            r, unit, _ = _dist(p, top.idx14)
            e, c = _lj(r, top.prm14[:, 0], top.prm14[:, 1], top.prm14[:, 2], None, None)
            ene[i, 1] += e.sum()
            _pair_scatter(f, top.idx14, unit, c)
            e, c = _coulomb(r, top.charge[top.idx14[:, 0]], top.charge[top.idx14[:, 1]], top.prm14[:, 3], False, None,
                            solventDielectric)
            ene[i, 0] += e.sum()
            _pair_scatter(f, top.idx14, unit, c)
        i_idx, i_owner, i_prm, i_h = top.torsions["improper"]
        if len(i_idx):
            ene[i, 6] += _torsion(p, f, i_idx, i_owner, i_prm, i_h)
        if len(top.pairs):
            r, unit, _ = _dist(p, top.pairs)
            idx = top.pairs
            if cutoff is not None:
                keep = r <= cutoff
                r, unit, idx = r[keep], unit[keep], idx[keep]
            e, c = _coulomb(r, top.charge[idx[:, 0]], top.charge[idx[:, 1]], 1, rfa, cutoff, solventDielectric)
            ene[i, 0] += e.sum()
            _pair_scatter(f, idx, unit, c)
            ty = top.atype[idx]
            e, c = _lj(r, top.A[ty[:, 0], ty[:, 1]], top.B[ty[:, 0], ty[:, 1]], 1, switch_dist, cutoff)
            ene[i, 1] += e.sum()
            _pair_scatter(f, idx, unit, c)
        if smd is not None:
            d = p[smd["atom2"]] - p[smd["atom1"]]
            dist = torch.sqrt((d * d).sum())
            dev = smd["center"] - dist
            ene[i, 7] += 0.5 * smd["fk"] * dev * dev
            g = smd["fk"] * dev / dist * d
            f[smd["atom2"]] += g
            f[smd["atom1"]] -= g
    return ene


def integrate(top: PortTopology, pos, vel, forces, nsteps, dt, gamma=0.0, temperature=0.0, generator=None,
              smd=None, **fkw):
    """Restated upstream Integrator.step(nsteps) on (R,N,3) tensors, in place.  Returns (Ekin, energies)."""
    m = top.mass
    vcoeff = torch.sqrt(2.0 * gamma / m * BOLTZMAN * temperature * dt) if temperature else None
    ene = None
    for s in range(nsteps):
        acc = forces / m
        pos += vel * dt + 0.5 * acc * dt * dt
        vel += 0.5 * dt * acc
        step_smd = None
        if smd is not None:
            step_smd = dict(smd, center=smd["center0"] + smd.get("center_rate", 0.0) * (smd.get("step0", 0) + s + 1))
        ene = compute(top, pos, forces, smd=step_smd, **fkw)
        if temperature:
            vel += -gamma * vel * dt + torch.randn(vel.shape, generator=generator, dtype=vel.dtype) * vcoeff
        vel += 0.5 * dt * (forces / m)
    ekin = (0.5 * (vel * vel).sum(2, keepdim=True) * m).sum(1).reshape(-1)
    return ekin, ene
