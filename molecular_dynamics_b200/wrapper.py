"""``Wrapper(natoms, bonds, device).wrap(pos, box, wrapidx=None)`` -- upstream ``torchmd.wrapper`` as the tutorial uses it
(V_Calculations/Test-1/tutorial.ipynb cell 18: ``wrapper.wrap(system.pos, system.box)`` after every ``step``; SURVEY.md
8f N3).  Moves whole bonded groups back into ``[0, box)`` by their centroid (so molecules are not split) and
non-bonded atoms individually; a box whose diagonal is all zero (vacuum, the ALA10 case) is a no-op.  Restated from
upstream TorchMD 1.0.x ("upstream, unpinned": the file is not in the reference).  Torch plumbing on whatever device
``pos`` lives on; the kernels apply the minimum image themselves and never need wrapped coordinates.
"""
from __future__ import annotations

import numpy as np
import torch


def _connected_groups(natoms, bonds):
    parent = list(range(natoms))

    def find(a):
        while parent[a] != a:
            parent[a] = parent[parent[a]]
            a = parent[a]
        return a

    for a, b in np.asarray(bonds, dtype=np.int64).reshape(-1, 2):
        ra, rb = find(int(a)), find(int(b))
        if ra != rb:
            parent[max(ra, rb)] = min(ra, rb)
    groups = {}
    for a in range(natoms):
        groups.setdefault(find(a), []).append(a)
    return [g for g in groups.values() if len(g) > 1], [g[0] for g in groups.values() if len(g) == 1]


class Wrapper:
    def __init__(self, natoms, bonds, device="cpu"):
        bonds = bonds.detach().cpu().numpy() if isinstance(bonds, torch.Tensor) else np.asarray(bonds)
        groups, single = _connected_groups(natoms, bonds if bonds is not None and len(bonds) else np.zeros((0, 2), int))
        self.groups = [torch.tensor(g, device=device) for g in groups]
        self.nongrouped = torch.tensor(single, dtype=torch.long, device=device)

    def wrap(self, pos, box, wrapidx=None):
        diag = torch.diagonal(box, dim1=1, dim2=2).to(pos.dtype)  # (R, 3): only the diagonal, like the force code
        if torch.all(diag == 0):
            return
        safe = torch.where(diag == 0, torch.ones_like(diag), diag)
        live = (diag != 0).to(pos.dtype)
        if wrapidx is not None:  # put the centroid of `wrapidx` at the box centre
            idx = torch.as_tensor(wrapidx, device=pos.device)
            com = pos[:, idx].mean(dim=1)
            pos -= ((com - diag / 2) * live).unsqueeze(1)
        for g in self.groups:
            g = g.to(pos.device)
            com = pos[:, g].mean(dim=1)
            offset = torch.floor(com / safe) * diag
            pos[:, g] -= offset.unsqueeze(1)
        if len(self.nongrouped):
            ng = self.nongrouped.to(pos.device)
            offset = torch.floor(pos[:, ng] / safe.unsqueeze(1)) * diag.unsqueeze(1)
            pos[:, ng] -= offset
