"""Batched form of the ``Energy`` adapter every script of the reference carries.

The reference's ML callers evaluate a generated batch one sample at a time and rebuild everything per sample
(``cGAN/cGAN_D_pos-2.3.2/cgan.py:345-353``: ``Energy(psf_file, parameter_file, colvar)`` -> Molecule, ForceField,
Parameters, System and Forces for EVERY sample; 41-105 evaluations/s on a V100).  This class keeps
that adapter's methods (``calc_energy(coords, time)`` -> ``(Epot, phis)``, ``calc_phi``, ``calc_colvar``; class at
``cgan.py:137-244``, ``V_Calculations/Test-6_full_system/potcalc.py:19-97``) and adds ``calc_energy_batch``: the
whole batch in one ``forces_compute`` launch plus one ``dihedral_angles`` launch, differentiable w.r.t. the
coordinates (the generator's loss back-propagates through it, ``cgan.py:354-360``), with the SMD end-to-end restraint
at one centre per sample (``time`` is the frame fraction the generator was conditioned on).
"""
from __future__ import annotations

import numpy as np
import torch

from . import _cabi
from .forces import Forces, SMDBias
from .systems import System


class Energy:
    UNITS = "kcal/mol"

    def __init__(self, psf_file=None, parameter_file=None, colvar=None, device="cuda", precision=torch.float, etype="all",
                 parameters=None, cutoff=9, rfa=True, switch_dist=7.5):
        """``Energy(psf_file, parameter_file, colvar=...)`` as in the scripts (PSF + YAML force field), or
        ``Energy(parameters=par, colvar=...)`` with a prebuilt parameter object.  ``cutoff=9, rfa=True,
        switch_dist=7.5`` are the values the adapter hard-codes (cgan.py:228)."""
        if parameters is None:
            from .parameters import parameters_from_psf_yaml

            parameters = parameters_from_psf_yaml(psf_file, parameter_file, precision=precision, device=device)
        self.etype = etype
        self.dtype = precision
        self.device = torch.device(device)
        self.parameters = parameters
        self.num_atoms = len(parameters.masses)
        self.system = System(self.num_atoms, nreplicas=1, precision=precision, device=self.device)
        self.colvar_name = None
        self.bias = None
        if colvar is not None:
            self.colvar_name = colvar["name"]
            self.bias = SMDBias.from_colvar(colvar)
        # built ONCE: the reference rebuilds this object for every sample
        self.forces = Forces(parameters, cutoff=cutoff, rfa=rfa, switch_dist=switch_dist, external=self.bias)

    def __str__(self):
        return f"Energy type is {self.etype} in units of {self.UNITS}"

    # ---- the adapter's single-sample methods -------------------------------------------------------------------
    def calc_phi(self, forces_obj, spos, sbox):
        """(ndihedrals,) angles of replica 0, as the adapter returns them (cgan.py:166-179)."""
        return forces_obj.dihedral_angles(spos.to(torch.float32).contiguous(), sbox)[0]

    def calc_colvar(self, time):
        """(dist, energy) of the end-to-end restraint for ``self.system.pos`` (cgan.py:209-217)."""
        b = self.bias
        cur_center = (b.cent_1 - b.cent_0) * time + b.cent_0
        p = self.system.pos[0]
        dist = torch.pow(torch.sum(torch.pow(p[b.atom2] - p[b.atom1], 2)), 0.5)
        engy = 0.5 * b.fk * torch.pow(cur_center - dist, 2)
        return dist, engy

    def calc_energy(self, coords, time=None):
        """One conformer, the adapter's return value: ``([dict of 1-element tensors], phis)`` for ``etype='all'``,
        else ``[energy of that term]`` (cgan.py:219-244)."""
        if isinstance(coords, np.ndarray):
            coords = torch.tensor(coords, dtype=self.dtype, device=self.device)
        coords = coords.reshape(self.num_atoms, 3).to(device=self.device, dtype=self.dtype)
        if self.colvar_name is not None and time is None:
            raise ValueError("No time provided.")  # the adapter prints this and exits the process
        ene, phis = self.calc_energy_batch(coords.unsqueeze(0), None if time is None else torch.as_tensor(time).reshape(1))
        self.system.pos = coords.detach().unsqueeze(0)
        epot = {k: v for k, v in ene.items()}
        if self.etype == "all":
            return [epot], phis[0]
        return [epot[self.etype]]

    # ---- the batched call ----------------------------------------------------------------------------------------
    def calc_energy_batch(self, coords, times=None):
        """``coords`` (B, natoms, 3) CUDA float tensor (may require grad), ``times`` (B,) frame fractions in [0, 1]
        (needed with a colvar).  Returns ``(energies, phis)``: ``energies`` maps the term names of the reference
        (``electrostatics, lj, bonds, angles, dihedrals, 1-4, impropers, external`` and the colvar name) to (B,)
        tensors, ``phis`` is (B, ndihedrals).  Two kernel launches for the whole batch; both carry autograd edges."""
        from .autograd import energies_with_grad

        if coords.dim() != 3 or coords.shape[1:] != (self.num_atoms, 3):
            raise ValueError(f"coords must be (B, {self.num_atoms}, 3)")
        B = coords.shape[0]
        pos = coords.to(device=self.device, dtype=torch.float32)
        if self.bias is not None:
            if times is None:
                raise ValueError("No time provided.")
            t = torch.as_tensor(times, dtype=torch.float32, device=self.device).reshape(-1)
            if t.numel() != B:
                raise ValueError("times must have one entry per conformer")
            cen = self.bias.cent_0 + (self.bias.cent_1 - self.bias.cent_0) * t
            self.bias.center = cen if B > 1 else float(cen[0])
        box = torch.zeros((B, 3, 3), dtype=torch.float32, device=self.device)
        frc = torch.empty((B, self.num_atoms, 3), dtype=torch.float32, device=self.device)
        if pos.requires_grad:
            ene = energies_with_grad(self.forces, pos, box, frc)
        else:
            ene = self.forces.compute_batch(pos, box, frc).to(pos.dtype)
        self.last_forces = frc
        keys = list(self.forces.energies) + ["external"]
        cols = ene.unbind(1)  # ONE autograd node for all columns (a select per term costs a zero-fill + copy each in backward)
        out = {k: cols[_cabi.energy_slot(k)] for k in keys}
        if self.colvar_name is not None:  # the adapter books the restraint under the colvar's name
            out[self.colvar_name] = out.pop("external")
        phis = self.forces.dihedral_angles(pos, None)
        return out, phis
