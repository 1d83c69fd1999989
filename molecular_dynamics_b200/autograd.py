"""Differentiable energies for the ML callers.

The reference's TorchMD_Fix patches exist so that per-term energies stay differentiable w.r.t. the
positions (systems.py:54, forces.py:341-344; consumed at cGAN/cGAN_D_pos-2.3.2/cgan.py:351-360).  Here
the energies come from a kernel, so the graph edge is provided by an autograd.Function whose backward is

    d(sum_k g_k E_k)/d pos = - sum_k g_k F_k

When every enabled term of a replica receives the same upstream weight the total force of the forward
launch is reused; otherwise columns that share a weight form a group, the most expensive group is taken from the
forward's total force, and each further group costs one force-only launch (``energies=False``: no FP64 energy pass):
``sum_g w_g F_g = w_0 F_tot + sum_{g > 0} (w_g - w_0) F_g``.
"""
from __future__ import annotations

import torch

from . import _cabi

_B = _cabi.TERM_BITS
# energy column -> term bits whose forces belong to that column (1-4 energies are booked under lj /
# electrostatics, forces.py:202-226)
_COLUMN_TERMS = {
    0: lambda m: (m & _B["electrostatics"]) | (m & _B["1-4"]),
    1: lambda m: (m & (_B["lj"] | _B["repulsion"] | _B["repulsioncg"])) | (m & _B["1-4"]),
    2: lambda m: m & _B["bonds"],
    3: lambda m: m & _B["angles"],
    4: lambda m: m & _B["dihedrals"],
    6: lambda m: m & _B["impropers"],
}


def _term_mask(mask, col):
    """Term bits of the launch that reproduces the forces behind energy column ``col`` (the kernels evaluate the 1-4
    pairs of any launch whose mask carries 1-4 together with lj / electrostatics; dihedrals need not be on)."""
    return _COLUMN_TERMS[col](mask)


class _EnergyFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pos, fobj, box, forces):
        from .forces import SMDBias

        ene = fobj.compute_batch(pos, box, forces)
        ctx.fobj = fobj
        # the SMD centre of THIS forward: Energy.calc_energy_batch overwrites bias.center on every call, and a loss
        # may hold several forwards before its backward (real and fake batch, gradient accumulation)
        ctx.smd_center = None
        if isinstance(fobj.external, SMDBias):
            c = fobj.external.center
            ctx.smd_center = c.detach().clone() if isinstance(c, torch.Tensor) else float(c)
        ftot = forces.detach().to(torch.float32).clone()
        if fobj._last_ext_force is not None:
            # a generic external object's force is already inside `forces`, and its own torch ops carry its gradient
            # (energies_with_grad): keep it out of the force this Function differentiates with
            ftot -= fobj._last_ext_force.to(ftot.dtype)
        ctx.save_for_backward(pos.detach().to(torch.float32).contiguous(), box.detach(), ftot)
        ctx.in_dtype = pos.dtype
        return ene.to(pos.dtype)

    @staticmethod
    def backward(ctx, g):
        pos, box, ftot = ctx.saved_tensors
        fobj = ctx.fobj
        from .forces import SMDBias

        mask = fobj._terms_mask
        cols = [c for c, fn in _COLUMN_TERMS.items() if fn(mask)]
        smd = isinstance(fobj.external, SMDBias)
        if smd:
            cols.append(7)
        extra = list(fobj._extra_pair_terms)  # further lj-family terms: columns 8, 9 of the energy tensor
        allcols = cols + [_cabi.NUM_ENERGIES + k for k in range(len(extra))]
        if not allcols:
            return torch.zeros_like(pos).to(ctx.in_dtype), None, None, None
        # which columns carry the same weight on every replica: ONE small comparison and one device -> host read (each
        # torch.equal / bool() here was a stream synchronisation of its own; a loss like cgan.py:351-353 paid ten per step)
        gsel = g[:, allcols]
        eq = (gsel.unsqueeze(2) == gsel.unsqueeze(1)).all(0).cpu()
        if bool(eq[0].all()):  # uniform weight per replica: reuse the forward launch
            grad = -gsel[:, :1, None].to(ftot.dtype) * ftot
            return grad.to(ctx.in_dtype), None, None, None
        b = box.to(device=pos.device, dtype=torch.float32).contiguous()
        tmp = torch.empty_like(pos)
        # columns that carry the same weight on every replica share a launch: the ML callers' loss
        # sum(terms) + 200 * E2End (cgan.py:351-353) needs two, not one per column
        groups = []
        for i, c in enumerate(cols):
            for grp in groups:
                if bool(eq[i, cols.index(grp[0])]):
                    grp.append(c)
                    break
            else:
                groups.append([c])
        # The forward launch already holds the sum of ALL per-column forces (ftot), so the most expensive group needs no
        # launch of its own: with weights w_g,  sum_g w_g F_g = w_0 F_tot + sum_{g > 0} (w_g - w_0) F_g.  The ML callers'
        # loss (all terms at one weight, the restraint at another) costs ONE extra launch -- of the restraint alone.
        def cost(grp):
            tm = 0
            for c in grp:
                if c != 7:
                    tm |= _term_mask(mask, c)
            return bin(tm).count("1") + (8 if tm & (_B["lj"] | _B["electrostatics"] | _B["repulsion"] | _B["repulsioncg"]) else 0)

        launches = [("cols", grp) for grp in groups] + [("extra", k) for k in range(len(extra))]
        base = max(range(len(groups)), key=lambda i: cost(groups[i]))
        w0 = g[:, groups[base][0]:groups[base][0] + 1, None].to(ftot.dtype)
        grad = -w0 * ftot
        for kind, item in launches:
            if kind == "cols":
                if item is groups[base]:
                    continue
                tmask = 0
                for c in item:
                    if c != 7:
                        tmask |= _term_mask(mask, c)
                fobj._launch(pos, b, tmp, terms_mask=tmask, with_smd=7 in item, energies=False, smd_center=ctx.smd_center)
                wg = g[:, item[0]:item[0] + 1, None].to(tmp.dtype)
            else:
                fobj._launch(pos, b, tmp, terms_mask=_B[extra[item]], with_smd=False, energies=False)
                wg = g[:, _cabi.NUM_ENERGIES + item:_cabi.NUM_ENERGIES + item + 1, None].to(tmp.dtype)
            grad -= (wg - w0) * tmp
        return grad.to(ctx.in_dtype), None, None, None


def energies_with_grad(fobj, pos, box, forces):
    """(R, 8) energies in pos.dtype, differentiable w.r.t. ``pos``; fills ``forces`` like compute()."""
    from .forces import SMDBias

    ene = _EnergyFunction.apply(pos, fobj, box, forces)
    ext = fobj.external
    if ext is not None and not isinstance(ext, SMDBias):
        # generic external object: its own torch ops carry the graph (forces.py:318-323); the forward above
        # already added its detached energy/force, so only the graph edge is added here
        ext_ene, _ = ext.calculate(pos, box)
        col = torch.zeros_like(ene)
        col[:, 7] = ext_ene.reshape(-1).to(ene.dtype) - ext_ene.detach().reshape(-1).to(ene.dtype)
        ene = ene + col
    return ene


class _DihedralAngles(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pos, fobj, box):
        p = pos.detach().to(torch.float32).contiguous()
        ctx.fobj, ctx.box, ctx.in_dtype = fobj, box, pos.dtype
        ctx.save_for_backward(p)
        return fobj._phi_launch(p, box).to(pos.dtype)

    @staticmethod
    def backward(ctx, g):
        (p,) = ctx.saved_tensors
        return ctx.fobj._phi_launch(p, ctx.box, grad_phi=g.to(torch.float32).contiguous()).to(ctx.in_dtype), None, None


def dihedral_angles_with_grad(fobj, pos, box):
    return _DihedralAngles.apply(pos, fobj, box)
