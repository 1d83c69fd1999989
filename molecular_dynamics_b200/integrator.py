"""Drop-in for upstream ``torchmd.integrator`` (not vendored by the reference; used at
V_Calculations/Test-1/tutorial.ipynb:212,313,374): ``Integrator(systems, forces, timestep, device,
gamma, T).step(niter) -> (Ekin, pot, T)`` and ``maxwell_boltzmann(masses, T, replicas)``.

``step(niter)`` is ONE kernel launch: velocity Verlet (+ Langevin kick) fused with the force evaluation,
the replica's positions / velocities / forces staying on chip for whole slices of the ``niter`` steps
(b200md_integrate, include/b200md.h).  A ``Forces`` object the fused kernel cannot express (a generic ``external``
object with its own ``calculate``, several pair terms of the lj family at once) takes the same loop as separate
launches per step, ``Forces.compute_batch`` in the middle, on the same stream.  Semantics restated from upstream
(SURVEY.md 3.2):

    for _ in range(niter):
        pos += vel*dt + 0.5*(F/m)*dt^2 ; vel += 0.5*dt*F/m          # _first_VV
        pot = forces.compute(pos, box, forces)
        if T: vel += -gamma*vel*dt + sqrt(2*gamma/m*kB*T*dt) * N(0,1)  # langevin
        vel += 0.5*dt*F/m                                            # _second_VV
"""
from __future__ import annotations

import numpy as np
import torch

from . import _cabi
from .forces import SMDBias, _integrate

TIMEFACTOR = 48.88821  # fs per internal time unit
BOLTZMAN = 0.001987191  # kcal/mol/K
PICOSEC2TIMEU = 1000.0 / TIMEFACTOR


def kinetic_energy(masses, vel):
    return torch.sum(0.5 * torch.sum(vel * vel, dim=2, keepdim=True) * masses, dim=1)


def kinetic_to_temp(Ekin, natoms):
    return 2.0 / (3.0 * natoms * BOLTZMAN) * Ekin


def maxwell_boltzmann(masses, T, replicas=1, generator=None):
    """vel = sqrt(T*kB/m) * N(0,1), shape (replicas, natoms, 3), dtype/device of ``masses``."""
    natoms = len(masses)
    noise = torch.randn((replicas, natoms, 3), generator=generator,
                        device=masses.device if generator is None else generator.device).to(masses)
    return torch.sqrt(T * BOLTZMAN / masses).unsqueeze(0) * noise


class _ReplicaEnergies:
    """What upstream returns as ``pot`` is a Python list with one tensor per replica; building that list costs
    ~1.5 us per replica per call (12 ms for 8 192 replicas, more than the transfers of a whole launch).  This is the same
    sequence, backed by one (R,) tensor and indexed lazily: ``pot[i]``, ``len(pot)``, iteration, ``pot.tensor``."""

    def __init__(self, tensor):
        self.tensor = tensor

    def __len__(self):
        return self.tensor.shape[0]

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self.tensor[k] for k in range(*i.indices(len(self)))]
        return self.tensor[i]

    def __iter__(self):
        return iter(self.tensor)

    def __repr__(self):
        return f"_ReplicaEnergies({self.tensor!r})"

    # the tutorial loop logs ``Epot + Ekin`` with Ekin a numpy array (tutorial.ipynb cell 18)
    def __array__(self, dtype=None, copy=None):
        a = self.tensor.detach().cpu().numpy()
        return a.astype(dtype) if dtype is not None else a

    def __add__(self, other):
        return np.asarray(self) + np.asarray(other)

    __radd__ = __add__


class Integrator:
    def __init__(self, systems, forces, timestep, device=None, gamma=None, T=None, seed=0, rigid_bonds=False):
        """Upstream's signature plus ``seed`` (counter-based RNG) and ``rigid_bonds``: hold every bond with a hydrogen
        end at its equilibrium length (NAMD ``rigidBonds all``, the setting of the reference's data generator, which
        is what allows its 2 fs step): SHAKE / RATTLE inside the fused kernel.  A start structure off the constraint
        surface is projected onto it by the first step (``apply_constraints`` does the same without stepping)."""
        self.dt = timestep / TIMEFACTOR
        self.systems = systems
        self.forces = forces
        self.device = device
        self.gamma = (gamma if gamma is not None else 0.0) / PICOSEC2TIMEU
        self.T = T
        if T:
            M = self.forces.par.masses
            self.vcoeff = torch.sqrt(2.0 * self.gamma / M * BOLTZMAN * T * self.dt).to(device)
        # counter-based RNG state: (seed, global step, global replica) -> reproducible under any sharding
        self.seed = int(seed)
        self.step_count = 0
        self.replica0 = 0
        self.energy_every = 0  # 0: record observables after the last step of each step() call only
        self.snapshot_every = 0
        # (nreports, R, 11) float64 after step(): 8 energy slots, Ekin, |r_a2 - r_a1|, SMD work since step 0
        self.observables = None
        self.smd_work = None  # (R,) float64: work done by the moving SMD restraint over all step() calls so far
        self.snapshots = None  # (nsnap, R, natoms, 3) float32 when snapshot_every > 0
        self.check_nan = True
        self.rigid_bonds = bool(rigid_bonds)
        self.num_constraints = 0  # known after the first step(): rigid bonds held, 3 N - this degrees of freedom

    def apply_constraints(self):
        """Put ``systems.pos`` / ``systems.vel`` on the constraint surface of ``rigid_bonds`` (positions along their
        present X-H bond vectors, bond-parallel velocity components removed) without taking a step: for start structures
        that were not produced by a constrained run, before a minimisation, an energy evaluation or a checkpoint.  (The
        first rigid step would otherwise do the same projection.)  One kernel launch for all replicas."""
        import ctypes

        s, f = self.systems, self.forces
        for name in ("pos", "vel"):
            t = getattr(s, name)
            if not t.is_cuda or t.dtype != torch.float32:
                raise RuntimeError(f"System.{name} must be a float32 CUDA tensor: the B200 path has no CPU fallback")
            if not t.is_contiguous():
                setattr(s, name, t.detach().contiguous())
        dev = s.pos.device
        R = s.pos.shape[0]
        status = torch.zeros(R, dtype=torch.int32, device=dev)
        lib = _cabi.load_library()
        with torch.cuda.device(dev):
            rc = lib.b200md_constrain(ctypes.c_void_p(f._handle(dev)), R, ctypes.c_void_p(s.pos.data_ptr()),
                                      ctypes.c_void_p(s.vel.data_ptr()), ctypes.c_void_p(status.data_ptr()),
                                      ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
        if rc == _cabi.ERR_UNSUPPORTED:
            raise NotImplementedError(_cabi.last_error())
        if rc != _cabi.OK:
            raise RuntimeError(f"b200md error {rc}: {_cabi.last_error()}")
        if R and int(status.max().item()) & _cabi.FLAG_CONSTRAINT:
            raise RuntimeError("SHAKE did not converge for a hydrogen group: the start structure is too far from its bond lengths")

    def step(self, niter=1):
        s, f = self.systems, self.forces
        for name in ("pos", "vel", "forces"):
            t = getattr(s, name)
            if not t.is_cuda:
                raise RuntimeError(f"System.{name} must be a CUDA tensor: the B200 path has no CPU fallback")
            if t.dtype != torch.float32:
                raise TypeError("the CUDA path integrates in float32")
            if not t.is_contiguous():  # set_positions rebinding (Q17) may have left a permuted view
                setattr(s, name, t.detach().contiguous())
        pos, vel, frc = s.pos, s.vel, s.forces
        R, N, _ = pos.shape
        dev = pos.device
        box = s.box.to(device=dev, dtype=torch.float32).contiguous()
        ext = f.external
        if (ext is not None and not isinstance(ext, SMDBias)) or f._extra_pair_terms:
            return self._step_unfused(int(niter))
        energy_every = self.energy_every if 0 < self.energy_every <= niter else 0  # shorter call: report at its end
        nrep = -(-niter // energy_every) if energy_every > 0 else 1  # + a record after the last step of the call
        nsnap = niter // self.snapshot_every if self.snapshot_every > 0 else 0
        obs = torch.zeros((nrep, R, _cabi.NUM_OBSERVABLES), dtype=torch.float64, device=dev)
        snaps = torch.empty((nsnap, R, N, 3), dtype=torch.float32, device=dev)
        status = torch.empty(R, dtype=torch.int32, device=dev)
        if isinstance(ext, SMDBias):  # centre of the restraint after global step g: center + rate * g
            c = ext.center
            if isinstance(c, torch.Tensor):
                if c.numel() > 1:
                    raise ValueError("the fused integrator moves ONE restraint centre (center + rate_per_step * step) for "
                                     "all replicas: SMDBias.center is a per-replica tensor (left by "
                                     "Energy.calc_energy_batch?); set a scalar centre first")
                c = float(c.reshape(-1)[0])
            smd = (True, ext.atom1, ext.atom2, ext.fk, float(c), ext.rate_per_step)
        else:
            smd = (False, 0, 0, 0.0, 0.0, 0.0)
        smd_c0 = smd[4]
        call = torch.ops.b200md.integrate if torch.compiler.is_compiling() else _integrate
        call(pos, vel, frc, box, obs, snaps, status, f._handle(dev), *f._opts(), smd[0], smd[1], smd[2], smd[3], smd_c0,
             smd[5], float(self.dt), float(self.gamma), float(self.T) if self.T else 0.0, self.seed, self.step_count,
             self.replica0, int(niter), int(energy_every), int(self.snapshot_every), self.rigid_bonds)
        self.step_count += int(niter)
        if self.rigid_bonds:
            import ctypes

            self.num_constraints = max(0, _cabi.load_library().b200md_topology_num_constraints(ctypes.c_void_p(f._handle(dev))))
        if self.check_nan:
            flags = int(status.max().item()) if R else 0
            if flags & _cabi.FLAG_NAN_POS:
                raise RuntimeError("Found NaN coordinates.")
            if flags & _cabi.FLAG_CONSTRAINT:
                raise RuntimeError("SHAKE / RATTLE did not converge for a hydrogen group: the rigid bonds were far from "
                                   "their lengths (start from a constrained structure or reduce the timestep)")
            if flags & _cabi.FLAG_FORCE_RANGE:
                raise RuntimeError("the bonded force on an atom exceeded 4096 kcal/mol/A: the fused integrator accumulates "
                                   "bonded forces in 2^-17 fixed point and the system has blown up (reduce the timestep "
                                   "or relax the start structure)")
        if smd[0]:  # the kernel counts work from the first step of the launch: carry it across calls
            if self.smd_work is not None and self.smd_work.shape[0] == R:
                obs[:, :, _cabi.OBS_SMD_WORK] += self.smd_work.to(dev)
            self.smd_work = obs[-1, :, _cabi.OBS_SMD_WORK].clone()  # the last record is the end of the call
        self.observables, self.snapshots = obs, snaps
        last = obs[-1]
        Ekin = last[:, _cabi.OBS_EKIN].cpu().numpy()
        T = kinetic_to_temp(Ekin, N)
        if self.num_constraints:  # each rigid bond removes a degree of freedom
            T = T * (3.0 * N / (3.0 * N - self.num_constraints))
        keys = list(f.energies) + ["external"]
        cols = [_cabi.energy_slot(k) for k in keys]
        pot = last[:, cols].sum(dim=1).to(pos.dtype)
        return Ekin, _ReplicaEnergies(pot), T

    def _step_unfused(self, niter):
        """Upstream's loop as launches per step, for what the fused kernel cannot hold: a generic ``external`` object
        (forces.py:318-323: anything with ``calculate(pos, box) -> (energy (R,), force (R, N, 3))``) and several pair
        terms of the lj family in one ``Forces`` object (forces.py:268-316).  Everything stays on the GPU and on the
        current stream: two elementwise updates and one ``Forces.compute_batch`` (kernel launches + the hook's torch
        ops) per step; the Langevin noise comes from a torch generator keyed by (seed, step)."""
        s, f = self.systems, self.forces
        if self.rigid_bonds:
            raise NotImplementedError("rigid_bonds lives in the fused kernel: not available with a generic external "
                                      "object or several lj-family terms")
        pos, vel, frc = s.pos, s.vel, s.forces
        R, N, _ = pos.shape
        dev = pos.device
        m = f.par.masses.to(device=dev, dtype=torch.float32).reshape(1, N, 1)
        dt = float(self.dt)
        ene = None
        for _ in range(niter):
            vel += (0.5 * dt) * frc / m  # _first_VV: pos += vel dt + 0.5 a dt^2 ; vel += 0.5 dt a
            pos += dt * vel
            ene = f.compute_batch(pos, s.box, frc)
            if self.T:
                gen = torch.Generator(device=dev).manual_seed((self.seed * 1000003 + self.step_count) & 0x7FFFFFFFFFFFFFFF)
                noise = torch.randn(pos.shape, generator=gen, device=dev, dtype=torch.float32)
                vel += -self.gamma * dt * vel + self.vcoeff.to(dev).reshape(1, N, 1) * noise
            vel += (0.5 * dt) * frc / m  # _second_VV
            self.step_count += 1
        obs = torch.zeros((1, R, _cabi.NUM_OBSERVABLES), dtype=torch.float64, device=dev)
        obs[0, :, :_cabi.NUM_ENERGIES] = ene[:, :_cabi.NUM_ENERGIES]
        for k, _term in enumerate(f._extra_pair_terms):  # further lj-family terms: folded into the "lj" column of the record
            obs[0, :, 1] += ene[:, _cabi.NUM_ENERGIES + k]
        ekin = kinetic_energy(m.to(torch.float64), vel.to(torch.float64)).reshape(-1)
        obs[0, :, _cabi.OBS_EKIN] = ekin
        self.observables = obs
        self.snapshots = torch.empty((0, R, N, 3), dtype=torch.float32, device=dev)
        Ekin = ekin.cpu().numpy()
        cols = [f.column(k) for k in list(f.energies) + ["external"]]
        return Ekin, _ReplicaEnergies(ene[:, cols].sum(dim=1).to(pos.dtype)), kinetic_to_temp(Ekin, N)
