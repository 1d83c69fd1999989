"""Surroundings of the step in the upstream TorchMD tutorial the reference ships (V_Calculations/Test-1/tutorial.ipynb,
cells 14-18): ``LogWriter`` (``torchmd.utils``), trajectory writer, checkpoint / restart.

None of this is on the hot path: it is host plumbing around ``Integrator.step``.  Upstream ``torchmd.utils`` is not
vendored by the reference ("upstream, unpinned"); the behaviour restated here is what the tutorial cells use.
"""
from __future__ import annotations

import csv
import os

import numpy as np
import torch


class LogWriter:
    """CSV monitor file, ``LogWriter(path, keys, header=None, name='monitor.csv').write_row(dict)``
    (tutorial.ipynb cell 16: keys ``('iter','ns','epot','ekin','etot','T')``).  One row per call; array / tensor
    values are written as their first element when they hold one item, else as a ``;``-joined list (one per replica)."""

    def __init__(self, path, keys, header=None, name="monitor.csv"):
        self.path, self.keys = path, tuple(keys)
        os.makedirs(path, exist_ok=True)
        self.fname = os.path.join(path, name)
        with open(self.fname, "w", newline="") as fh:
            if header is not None:
                fh.write(str(header) + "\n")
            csv.writer(fh).writerow(self.keys)

    @staticmethod
    def _cell(v):
        if hasattr(v, "tensor"):  # Integrator.step's per-replica energy sequence
            v = v.tensor
        if isinstance(v, torch.Tensor):
            v = v.detach().cpu().numpy()
        if isinstance(v, (list, tuple)):
            v = np.array([float(x) for x in v])
        if isinstance(v, np.ndarray):
            v = v.reshape(-1)
            return repr(float(v[0])) if v.size == 1 else ";".join(repr(float(x)) for x in v)
        return v

    def write_row(self, rowdict):
        with open(self.fname, "a", newline="") as fh:
            csv.writer(fh).writerow([self._cell(rowdict[k]) for k in self.keys])


class TrajectoryWriter:
    """Frames of ``system.pos`` collected in the reference's trajectory layout: ``np.save(path, (frames, natoms, 3))``
    float32 per replica file (Data_Processing_Scripts/dcd_to_npy-2.0.py:24-40), or the tutorial's stacked layout
    ``(replicas, natoms, 3, frames)`` (cell 18: ``np.stack(traj, axis=2)``) with ``layout='tutorial'``."""

    def __init__(self, path, layout="frames"):
        if layout not in ("frames", "tutorial"):
            raise ValueError("layout must be 'frames' or 'tutorial'")
        self.path, self.layout, self.frames = path, layout, []

    def append(self, pos):
        p = pos.detach().to("cpu", torch.float32).numpy().copy() if isinstance(pos, torch.Tensor) else np.asarray(pos, np.float32).copy()
        self.frames.append(p)

    def extend_from_snapshots(self, snapshots):
        """Integrator.snapshots (nsnap, R, N, 3) -> one frame per snapshot."""
        for s in snapshots:
            self.append(s)

    def save(self):
        if not self.frames:
            raise RuntimeError("no frames collected")
        if self.layout == "tutorial":
            np.save(self.path, np.stack(self.frames, axis=3))
            return [self.path]
        arr = np.stack(self.frames, axis=0)  # (frames, R, N, 3)
        if arr.shape[1] == 1:
            np.save(self.path, arr[:, 0])
            return [self.path]
        base, ext = os.path.splitext(self.path)
        out = []
        for r in range(arr.shape[1]):
            out.append(f"{base}_rep-{r}{ext or '.npy'}")
            np.save(out[-1], arr[:, r])
        return out


def save_checkpoint(path, system, integrator=None, extra=None):
    """pos / vel / box / forces (+ the integrator's counter-based RNG state: seed, global step, first global replica).
    With the fused integrator's deterministic reductions a run resumed from a checkpoint continues bit-identically.
    The file holds tensors, numbers and plain containers only (``extra`` included), so that it loads with
    ``weights_only=True``: a checkpoint from somewhere else cannot execute code when it is read."""
    state = {"pos": system.pos.detach().cpu(), "vel": system.vel.detach().cpu(), "box": system.box.detach().cpu(),
             "forces": system.forces.detach().cpu(), "extra": extra}
    if integrator is not None:
        state["integrator"] = {"seed": integrator.seed, "step_count": integrator.step_count, "replica0": integrator.replica0,
                               "dt": integrator.dt, "gamma": integrator.gamma, "T": integrator.T,
                               "smd_work": None if integrator.smd_work is None else integrator.smd_work.detach().cpu()}
    torch.save(state, path)


def load_checkpoint(path, system, integrator=None):
    state = torch.load(path, map_location="cpu", weights_only=True)  # tensors / numbers / containers: no unpickling of code
    for name in ("pos", "vel", "box", "forces"):
        t = getattr(system, name)
        src = state[name]
        if tuple(t.shape) != tuple(src.shape):
            raise ValueError(f"checkpoint {name} has shape {tuple(src.shape)}, system has {tuple(t.shape)}")
        setattr(system, name, src.to(device=t.device, dtype=t.dtype).contiguous())
    if integrator is not None and "integrator" in state:
        ig = state["integrator"]
        integrator.seed, integrator.step_count, integrator.replica0 = ig["seed"], ig["step_count"], ig["replica0"]
        integrator.smd_work = ig.get("smd_work")
    return state.get("extra")


def write_colvars_traj(path, steps, dist, center, work, name="End2EndDist", bias="harmonic1"):
    """One replica's SMD record in the layout of the data generator's ``smd_out.colvars.traj``
    (All_ML_Training_Data/210905_SMD_decaalanine/SMD/output/N/smd_out.colvars.traj: ``# step  End2EndDist
    x0_End2EndDist  W_harmonic1``, one row per report, ``%.14e`` columns).  ``steps`` (n,) ints, the rest (n,) floats:
    ``Integrator.observables[:, r, OBS_E2E_DIST]``, the restraint centre and ``[:, r, OBS_SMD_WORK]``."""
    steps = np.asarray(steps).reshape(-1)
    cols = [np.asarray(c, dtype=np.float64).reshape(-1) for c in (dist, center, work)]
    if any(c.shape != steps.shape for c in cols):
        raise ValueError("steps, dist, center and work must have one entry per report")
    w = 23
    with open(path, "w") as fh:
        fh.write("# " + "step".ljust(11) + name.rjust(w - 2) + ("x0_" + name).rjust(w) + ("W_" + bias).rjust(w) + "\n")
        for i, s in enumerate(steps):
            fh.write(f"{int(s):>11d} " + " ".join(f"{c[i]: .14e}".rjust(w - 1) for c in cols) + "\n")


def read_colvars_traj(path):
    """(steps, dist, center, work) arrays from a colvars.traj file (the reference's or ``write_colvars_traj``'s)."""
    rows = [ln.split() for ln in open(path) if ln.strip() and not ln.startswith("#")]
    a = np.array(rows, dtype=np.float64)
    return a[:, 0].astype(np.int64), a[:, 1], a[:, 2], a[:, 3]
