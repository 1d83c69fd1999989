"""TEST / BASELINE INFRASTRUCTURE -- never imported by molecular_dynamics_b200/.

The reference's OWN Forces.compute (TorchMD_Fix/forces.py:82-344, copied unmodified into oracle/_ref/ by
tools/make_oracle_ref.py) driven by the restated upstream velocity-Verlet loop (upstream
torchmd.integrator is not part of the reference repo).  Used by bench.py's `--impl reference` arm and cpu_baseline leg
as the timed CPU implementation of the path; `available()` is False where oracle/_ref was not shipped.
"""
import importlib.util
import os

import torch

_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")
_MOD = {}


def available():
    return os.path.exists(os.path.join(_DIR, "forces.py"))


def module(name="forces"):
    if name not in _MOD:
        spec = importlib.util.spec_from_file_location(f"b200md_oracle_ref_{name}", os.path.join(_DIR, name + ".py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        _MOD[name] = mod
    return _MOD[name]


def integrate(forces_obj, masses, pos, vel, frc, box, nsteps, dt):
    """upstream Integrator.step(nsteps), NVE, around the reference's Forces.compute; (R, N, 3) tensors in place."""
    m = masses.reshape(1, -1, 1).to(pos.dtype)
    pot = None
    for _ in range(nsteps):
        acc = frc / m
        pos += vel * dt + 0.5 * acc * dt * dt  # _first_VV
        vel += 0.5 * dt * acc
        pot = forces_obj.compute(pos, box, frc)
        vel += 0.5 * dt * (frc / m)  # _second_VV
    return pot
