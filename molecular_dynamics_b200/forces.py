"""Drop-in ``Forces`` for the reference's TorchMD_Fix/forces.py, backed by the sm_100a kernels.

Same constructor, attributes and ``compute`` contract as the reference class (forces.py:28-73,
:82-344; behaviours Q1-Q15 of SURVEY.md 7.4).  The arithmetic runs in ``libb200md.so`` through the
PyTorch custom ops ``b200md::forces_compute`` / ``b200md::integrate``; there is no CPU fallback: CPU
tensors, a missing library or an unsupported option raise.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _cabi

__all__ = ["Forces", "SMDBias"]


# ----------------------------------------------------------------------------------------------
# custom ops: thin shims over the C-ABI (pointers + sizes only)
# ----------------------------------------------------------------------------------------------
def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream_ptr(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _check(rc):
    if rc != _cabi.OK:
        msg = _cabi.last_error()
        if rc == _cabi.ERR_NAN_COORDINATES:
            raise RuntimeError("Found NaN coordinates.")
        if rc == _cabi.ERR_UNSUPPORTED:
            raise NotImplementedError(msg)
        if rc in (_cabi.ERR_INVALID_ARGUMENT, _cabi.ERR_TOO_MANY_ATOMS):
            raise ValueError(msg)
        raise RuntimeError(f"b200md error {rc}: {msg}")


def _forces_compute(pos: torch.Tensor, box: torch.Tensor, forces: torch.Tensor, energies: torch.Tensor,
                    status: torch.Tensor, handle: int, terms: int, cutoff: float, rfa: bool, eps: float,
                    switch_dist: float, smd_on: bool, smd_a1: int, smd_a2: int, smd_fk: float, smd_c0: float,
                    smd_centers: torch.Tensor) -> None:
    """ctypes call of b200md_forces_compute on the current stream.  The eager classes call this directly; the
    ``b200md::forces_compute`` custom op below is the same call for traced / compiled graphs (its Python dispatch costs
    ~0.2 ms per call, twice the kernels of a 128-conformer batch)."""
    lib = _cabi.load_library()
    fo = _cabi.ForceOpts(terms, cutoff, int(rfa), eps, switch_dist)
    cen = smd_centers.data_ptr() if smd_centers.numel() else None
    so = _cabi.SmdOpts(int(smd_on), smd_a1, smd_a2, smd_fk, smd_c0, 0.0, cen)
    ene = _ptr(energies) if energies.numel() else None  # empty: forces only (the library skips the FP64 energy pass)
    with torch.cuda.device(pos.device):
        _check(lib.b200md_forces_compute(C.c_void_p(handle), C.byref(fo), C.byref(so), pos.shape[0], _ptr(pos),
                                         _ptr(box), _ptr(forces), ene, _ptr(status), _stream_ptr(pos.device)))


@torch.library.custom_op("b200md::forces_compute", mutates_args=("forces", "energies", "status"))
def _forces_compute_op(pos: torch.Tensor, box: torch.Tensor, forces: torch.Tensor, energies: torch.Tensor,
                       status: torch.Tensor, handle: int, terms: int, cutoff: float, rfa: bool, eps: float,
                       switch_dist: float, smd_on: bool, smd_a1: int, smd_a2: int, smd_fk: float, smd_c0: float,
                       smd_centers: torch.Tensor) -> None:
    _forces_compute(pos, box, forces, energies, status, handle, terms, cutoff, rfa, eps, switch_dist, smd_on, smd_a1,
                    smd_a2, smd_fk, smd_c0, smd_centers)


def _integrate(pos: torch.Tensor, vel: torch.Tensor, forces: torch.Tensor, box: torch.Tensor,
               observables: torch.Tensor, snapshots: torch.Tensor, status: torch.Tensor, handle: int, terms: int,
               cutoff: float, rfa: bool, eps: float, switch_dist: float, smd_on: bool, smd_a1: int, smd_a2: int,
               smd_fk: float, smd_c0: float, smd_rate: float, dt: float, gamma: float, temperature: float,
               seed: int, step0: int, replica0: int, nsteps: int, energy_every: int, snapshot_every: int,
               rigid_bonds: bool = False) -> None:
    """ctypes call of b200md_integrate on the current stream (eager path; ``b200md::integrate`` below for traced graphs)."""
    lib = _cabi.load_library()
    fo = _cabi.ForceOpts(terms, cutoff, int(rfa), eps, switch_dist)
    so = _cabi.SmdOpts(int(smd_on), smd_a1, smd_a2, smd_fk, smd_c0, smd_rate, None)
    io = _cabi.IntegrateOpts(dt, gamma, temperature, seed, step0, replica0, nsteps, energy_every, snapshot_every,
                             int(rigid_bonds))
    snaps = _ptr(snapshots) if snapshots.numel() else None
    with torch.cuda.device(pos.device):
        _check(lib.b200md_integrate(C.c_void_p(handle), C.byref(fo), C.byref(so), C.byref(io), pos.shape[0], _ptr(pos),
                                    _ptr(vel), _ptr(forces), _ptr(box), _ptr(observables), snaps, _ptr(status),
                                    _stream_ptr(pos.device)))


@torch.library.custom_op("b200md::integrate", mutates_args=("pos", "vel", "forces", "observables", "snapshots", "status"))
def _integrate_op(pos: torch.Tensor, vel: torch.Tensor, forces: torch.Tensor, box: torch.Tensor,
                  observables: torch.Tensor, snapshots: torch.Tensor, status: torch.Tensor, handle: int, terms: int,
                  cutoff: float, rfa: bool, eps: float, switch_dist: float, smd_on: bool, smd_a1: int, smd_a2: int,
                  smd_fk: float, smd_c0: float, smd_rate: float, dt: float, gamma: float, temperature: float,
                  seed: int, step0: int, replica0: int, nsteps: int, energy_every: int, snapshot_every: int,
                  rigid_bonds: bool = False) -> None:
    _integrate(pos, vel, forces, box, observables, snapshots, status, handle, terms, cutoff, rfa, eps, switch_dist, smd_on,
               smd_a1, smd_a2, smd_fk, smd_c0, smd_rate, dt, gamma, temperature, seed, step0, replica0, nsteps, energy_every,
               snapshot_every, rigid_bonds)


class SMDBias:
    """Steered-MD end-to-end harmonic restraint, the bias the reference's ``Energy`` adapter adds on
    the caller side (V_Calculations/Test-6_full_system/potcalc.py:75-89; cGAN/.../cgan.py:209-217):

        c(t) = cent_0 + (cent_1 - cent_0) * t / T ;  E = 0.5 * fk * (c - |r[group2[0]] - r[group1[0]]|)^2

    Pass it as ``Forces(..., external=SMDBias(...))``: the kernels evaluate it fused (energy slot
    "external", force = -dE/dr).  ``center`` may be a float or a per-replica tensor; in the integrator the
    centre moves by ``rate_per_step`` each step.
    """

    def __init__(self, group1, group2, fk=1.0, cent_0=12.0, cent_1=34.0, T=1.0, rate_per_step=0.0):
        self.atom1 = int(group1[0] if hasattr(group1, "__len__") else group1)
        self.atom2 = int(group2[0] if hasattr(group2, "__len__") else group2)
        self.fk, self.cent_0, self.cent_1, self.T = float(fk), float(cent_0), float(cent_1), float(T)
        self.rate_per_step = float(rate_per_step)
        self.center = self.cent_0

    @classmethod
    def from_colvar(cls, colvar):
        """From the reference's ``colvar`` dict (potcalc.py:99-107)."""
        return cls(colvar["group1"], colvar["group2"], colvar["fk"], colvar["cent_0"], colvar["cent_1"], colvar["T"])

    def set_time(self, t):
        """Frame-index form of Q19: centre = cent_0 + (cent_1 - cent_0) * t / T (t: float or (R,) tensor)."""
        self.center = self.cent_0 + (self.cent_1 - self.cent_0) * t / self.T
        return self


class Forces:
    """Same interface as the reference class; see module docstring."""

    bonded = ["bonds", "angles", "dihedrals", "impropers", "1-4"]
    nonbonded = ["electrostatics", "lj", "repulsion", "repulsioncg"]
    terms = bonded + nonbonded

    def __init__(self, parameters, terms=None, external=None, cutoff=None, rfa=False, solventDielectric=78.5,
                 switch_dist=None, exclusions=("bonds", "angles", "1-4")):
        self.par = parameters
        if terms is None:
            terms = ("electrostatics", "lj", "bonds", "angles", "dihedrals", "1-4", "impropers")
        self.energies = [ene.lower() for ene in terms]
        for et in self.energies:
            if et not in Forces.terms:
                raise ValueError(f"Force term {et} is not implemented.")
        if "1-4" in self.energies and "dihedrals" not in self.energies:
            raise RuntimeError("You cannot enable 1-4 interactions without enabling dihedrals")
        # forces.py:268-316 evaluates every enabled pair term of the lj family over the same pair list, one after the
        # other.  The kernels have ONE pair-table slot per launch: the first family term rides with everything else,
        # each further one gets its own launch (that term alone) whose forces are added and whose energy is reported
        # in its own column behind the 8 standard ones.
        family = [et for et in self.energies if et in ("lj", "repulsion", "repulsioncg")]
        self._extra_pair_terms = family[1:]
        self.natoms = len(parameters.masses)
        self.require_distances = any(f in self.nonbonded for f in self.energies)
        self._pair_idx = (_cabi.make_pair_index(self.natoms, parameters.get_exclusions(exclusions))
                          if self.require_distances else None)
        self.ava_idx = (torch.tensor(self._pair_idx).to(parameters.device) if self.require_distances else None)
        self.external = external
        self.cutoff = cutoff
        self.rfa = rfa
        self.solventDielectric = solventDielectric
        self.switch_dist = switch_dist
        self.check_nan = True  # forces.py:91-92 costs one device->host flag read per call
        if rfa and cutoff is None:
            raise ValueError("rfa=True needs a cutoff (forces.py:467 divides by cutoff**3)")
        self._terms_mask = 0
        for et in self.energies:
            if et not in self._extra_pair_terms:
                self._terms_mask |= _cabi.TERM_BITS[et]
        self._packed = _cabi.PackedTopology(parameters, self._pair_idx)
        self._handles = {}  # device index -> native handle
        self.last_energies = None  # (R, 8) float64 tensor of the last compute (FP64 energy pass)
        self._last_ext_force = None  # detached force of a generic external object in the last compute_batch
        self._host_state = {}  # device index -> staging buffers and streams of compute_batch_host

    # -- native handle management ----------------------------------------------------------------
    def _handle(self, device):
        idx = device.index if device.index is not None else torch.cuda.current_device()
        if idx not in self._handles:
            lib = _cabi.load_library()
            h = lib.b200md_topology_create(C.byref(self._packed.desc), idx)
            if not h:
                msg = _cabi.last_error()
                if "natoms" in msg:
                    raise ValueError(msg)
                raise RuntimeError(f"b200md_topology_create failed: {msg}")
            self._handles[idx] = h
        return self._handles[idx]

    def __del__(self):
        try:
            lib = _cabi.load_library()
            for h in self._handles.values():
                lib.b200md_topology_destroy(C.c_void_p(h))
        except Exception:
            pass

    def exclusion_mask(self, device="cuda"):
        """(natoms, W) uint32 bitmask the kernels use (bit j % 32 of word j // 32 of row i set <=> pair (i, j) not
        evaluated); W = 4 up to 128 atoms, ceil(natoms / 32) for the large-solute kernels."""
        h = self._handle(torch.device(device))
        words = _cabi.load_library().b200md_topology_mask_words(C.c_void_p(h))
        out = np.zeros((self.natoms, words), dtype=np.uint32)
        _check(_cabi.load_library().b200md_topology_exclusion_mask(C.c_void_p(h), out.ctypes.data_as(C.c_void_p)))
        return out

    def _opts(self):
        return (self._terms_mask, float(self.cutoff) if self.cutoff is not None else 0.0, bool(self.rfa),
                float(self.solventDielectric), float(self.switch_dist) if self.switch_dist is not None else 0.0)

    def column(self, term):
        """Column of the energy tensor ``compute_batch`` returns that holds ``term``: the 8 standard slots
        (``_cabi.ENERGY_SLOTS``; the first enabled term of the lj family uses the "lj" column), then one column per
        further lj-family term."""
        if term in self._extra_pair_terms:
            return _cabi.NUM_ENERGIES + self._extra_pair_terms.index(term)
        return _cabi.energy_slot(term)

    def _smd_args(self, R, device, center=None):
        ext = self.external
        empty = torch.empty(0, dtype=torch.float32, device=device)
        if not isinstance(ext, SMDBias):
            return (False, 0, 0, 0.0, 0.0), empty
        c = ext.center if center is None else center
        if isinstance(c, torch.Tensor) and c.numel() > 1:
            cen = c.to(device=device, dtype=torch.float32).contiguous()
            if cen.numel() != R:
                raise ValueError("per-replica SMD centres must have one entry per replica")
            return (True, ext.atom1, ext.atom2, ext.fk, 0.0), cen
        return (True, ext.atom1, ext.atom2, ext.fk, float(c)), empty

    @staticmethod
    def _require_cuda(t, what):
        if not t.is_cuda:
            raise RuntimeError(f"{what} must be a CUDA tensor: the B200 path has no CPU fallback")

    # -- low-level launch ------------------------------------------------------------------------
    def _launch(self, pos, box, forces, terms_mask=None, with_smd=True, energies=True, smd_center=None):
        """One b200md::forces_compute launch; ``forces`` must be a contiguous float32 CUDA tensor.  ``energies=False``
        skips the FP64 energy pass (returns an empty tensor); ``smd_center`` overrides ``external.center`` (the
        autograd backward passes the centre its forward used)."""
        R = pos.shape[0]
        dev = pos.device
        ene = torch.empty((R if energies else 0, _cabi.NUM_ENERGIES), dtype=torch.float64, device=dev)
        status = torch.empty(R, dtype=torch.int32, device=dev)
        opts = list(self._opts())
        if terms_mask is not None:
            opts[0] = terms_mask
        (smd_on, a1, a2, fk, c0), cen = self._smd_args(R, dev, smd_center)
        if not with_smd:
            smd_on, cen = False, torch.empty(0, dtype=torch.float32, device=dev)
        call = torch.ops.b200md.forces_compute if torch.compiler.is_compiling() else _forces_compute
        call(pos, box, forces, ene, status, self._handle(dev), *opts, smd_on, a1, a2, fk, c0, cen)
        return ene, status

    # -- batched fast path -----------------------------------------------------------------------
    def compute_batch(self, pos, box, forces, energies=True):
        """Evaluate all replicas; fills ``forces`` in place and returns the (R, 8) float64 energy tensor (columns in
        ``_cabi.ENERGY_SLOTS`` order; see ``column`` for objects with several lj-family terms).  Energies are evaluated
        in FP64 from the fp32 positions.  ``energies=False``: forces only, returns None.  No Python work per replica.
        ``pos`` / ``forces`` may be float64 tensors (``precision=torch.double`` systems, systems.py:32-36): the kernels
        compute forces in fp32, so they are converted on the way in and out."""
        self._require_cuda(pos, "pos")
        self._require_cuda(forces, "forces")
        if pos.dtype not in (torch.float32, torch.float64):
            raise TypeError("positions must be float32 or float64 tensors")
        R, N, _ = pos.shape
        if N != self.natoms:
            raise ValueError(f"pos has {N} atoms, parameters have {self.natoms}")
        dev = pos.device
        p = pos.detach().to(torch.float32).contiguous()  # System.set_positions hands over a permuted view (Q17)
        direct = forces.is_contiguous() and forces.dtype == torch.float32 and forces.device == dev
        f = forces if direct else torch.empty((R, N, 3), dtype=torch.float32, device=dev)
        b = box.detach().to(device=dev, dtype=torch.float32).contiguous()
        ene, status = self._launch(p, b, f, energies=energies)
        for term in self._extra_pair_terms:  # forces.py:268-316: the further pair terms of the lj family, one launch each
            f2 = torch.empty_like(f)
            e2, _ = self._launch(p, b, f2, terms_mask=_cabi.TERM_BITS[term], with_smd=False, energies=energies)
            f += f2
            if energies:
                ene = torch.cat([ene, e2[:, _cabi.energy_slot(term):_cabi.energy_slot(term) + 1]], dim=1)
        if self.check_nan:
            flags = int(status.max().item()) if R else 0
            if flags & _cabi.FLAG_NAN_POS:
                raise RuntimeError("Found NaN coordinates.")
        if not direct:
            forces.copy_(f)
        self._last_ext_force = None
        if self.external is not None and not isinstance(self.external, SMDBias):  # forces.py:318-323
            ext_ene, ext_force = self.external.calculate(pos.detach(), box)
            if energies:
                ene[:, 7] += ext_ene.detach().to(ene.dtype).reshape(-1)
            forces += ext_force.detach().to(forces.dtype)
            self._last_ext_force = ext_force.detach()
        if not energies:
            return None
        self.last_energies = ene
        return ene

    # -- host buffers in, host buffers out: copies overlapped with the kernels -----------------------
    def compute_batch_host(self, pos, forces_out=None, energies_out=None, box=None, energies=True, chunks=8, device="cuda"):
        """``Forces.compute`` for a batch that lives in HOST memory (pinned for full speed): ``pos`` (R, N, 3) float32 CPU
        tensor → ``(forces_out, energies_out)`` CPU tensors (allocated pinned when not given; ``energies_out`` is None with
        ``energies=False``).  The batch is cut into ``chunks`` pieces that flow through three streams -- host→device copy,
        kernels, device→host copy -- so the two PCIe directions and the kernels overlap instead of running back to back
        (65 536 ALA10 conformers: 164 MB over PCIe against 2.3 ms of kernels).  Device staging buffers are kept on the
        object between calls.  Synchronous: returns when the results are in the host buffers; raises on NaN coordinates
        like ``compute``.  SMD per-replica centres and generic ``external`` objects: use ``compute_batch``."""
        if pos.is_cuda or pos.dtype != torch.float32 or pos.dim() != 3 or pos.shape[1] != self.natoms:
            raise ValueError(f"pos must be a float32 CPU tensor of shape (R, {self.natoms}, 3)")
        if self.external is not None and not (isinstance(self.external, SMDBias) and not isinstance(self.external.center, torch.Tensor)):
            raise NotImplementedError("compute_batch_host supports no external term beyond an SMDBias with one scalar centre")
        if self._extra_pair_terms:
            raise NotImplementedError("compute_batch_host evaluates one pair term of the lj family; use compute_batch")
        dev = torch.device(device)
        if dev.index is None:
            dev = torch.device("cuda", torch.cuda.current_device())
        R, N, _ = pos.shape
        pos = pos.contiguous()
        if forces_out is None:
            forces_out = torch.empty((R, N, 3), dtype=torch.float32).pin_memory()
        if energies and energies_out is None:
            energies_out = torch.empty((R, _cabi.NUM_ENERGIES), dtype=torch.float64).pin_memory()
        hbox = None if box is None else box.to(dtype=torch.float32).contiguous()
        st = self._host_state.get(dev.index)
        if st is None or st["R"] < R:
            st = {"R": R, "pos": torch.empty((R, N, 3), dtype=torch.float32, device=dev),
                  "frc": torch.empty((R, N, 3), dtype=torch.float32, device=dev),
                  "box": torch.zeros((R, 3, 3), dtype=torch.float32, device=dev),
                  "s_in": torch.cuda.Stream(dev), "s_k": torch.cuda.Stream(dev), "s_out": torch.cuda.Stream(dev)}
            self._host_state[dev.index] = st
        s_in, s_k, s_out = st["s_in"], st["s_k"], st["s_out"]
        start = torch.cuda.current_stream(dev).record_event()  # work queued by the caller before this call comes first
        for s_ in (s_in, s_k, s_out):
            s_.wait_event(start)
        nchunk = max(1, min(int(chunks), R))
        bounds = [R * k // nchunk for k in range(nchunk + 1)]
        keep, flags = [], []
        for k in range(nchunk):
            lo, hi = bounds[k], bounds[k + 1]
            if hi == lo:
                continue
            with torch.cuda.stream(s_in):
                st["pos"][lo:hi].copy_(pos[lo:hi], non_blocking=True)
                if hbox is not None:
                    st["box"][lo:hi].copy_(hbox[lo:hi], non_blocking=True)
                elif k == 0 and st.get("box_dirty"):
                    st["box"].zero_()
                e_in = s_in.record_event()
            with torch.cuda.stream(s_k):
                s_k.wait_event(e_in)
                ene, status = self._launch(st["pos"][lo:hi], st["box"][lo:hi], st["frc"][lo:hi], energies=energies)
                e_k = s_k.record_event()
            with torch.cuda.stream(s_out):
                s_out.wait_event(e_k)
                forces_out[lo:hi].copy_(st["frc"][lo:hi], non_blocking=True)
                if energies:
                    energies_out[lo:hi].copy_(ene, non_blocking=True)
                flags.append(status.max())
            keep.append((ene, status))  # alive until the synchronisation below (they were allocated on s_k, read on s_out)
        st["box_dirty"] = hbox is not None
        s_out.synchronize()
        torch.cuda.current_stream(dev).wait_stream(s_out)
        if self.check_nan and flags and int(torch.stack(flags).max().item()) & _cabi.FLAG_NAN_POS:
            raise RuntimeError("Found NaN coordinates.")
        return forces_out, (energies_out if energies else None)

    # -- Energy.calc_phi of the ML callers ---------------------------------------------------------
    def dihedral_angles(self, pos, box=None):
        """Proper-dihedral angles of every replica, (R, ndihedrals) in ``pos.dtype``, rows in ``par.dihedrals`` order
        (``Energy.calc_phi``, cGAN/cGAN_D_pos-2.3.2/cgan.py:166-207: phi = -atan2(sin, cos) as in forces.py:541-552).
        One launch for the whole batch; differentiable w.r.t. ``pos`` when it requires grad."""
        from .autograd import dihedral_angles_with_grad

        self._require_cuda(pos, "pos")
        if pos.dtype not in (torch.float32, torch.float64):
            raise TypeError("positions must be float32 or float64 tensors")
        if pos.shape[1] != self.natoms:
            raise ValueError(f"pos has {pos.shape[1]} atoms, parameters have {self.natoms}")
        return dihedral_angles_with_grad(self, pos, box)

    def _phi_launch(self, pos, box, grad_phi=None):
        lib = _cabi.load_library()
        dev = pos.device
        h = C.c_void_p(self._handle(dev))
        R = pos.shape[0]
        nd = lib.b200md_topology_ndihedrals(h)
        b = None if box is None else box.detach().to(device=dev, dtype=torch.float32).contiguous()
        bp = _ptr(b) if b is not None else None
        with torch.cuda.device(dev):
            if grad_phi is None:
                out = torch.empty((R, nd), dtype=torch.float32, device=dev)
                _check(lib.b200md_dihedral_angles(h, R, _ptr(pos), bp, _ptr(out), _stream_ptr(dev)))
            else:
                out = torch.empty((R, self.natoms, 3), dtype=torch.float32, device=dev)
                _check(lib.b200md_dihedral_angles_backward(h, R, _ptr(pos), bp, _ptr(grad_phi), _ptr(out), _stream_ptr(dev)))
        return out

    # -- reference-shaped entry point --------------------------------------------------------------
    def compute(self, pos, box, forces, returnDetails=False, explicit_forces=True):
        """Forces.compute of the reference (forces.py:82-344): zeroes and fills ``forces`` in place and
        returns per-replica energies (list of dicts of 1-element tensors when ``returnDetails``)."""
        if pos.requires_grad or not explicit_forces:
            from .autograd import energies_with_grad

            ene = energies_with_grad(self, pos, box, forces)  # (R, 8), differentiable w.r.t. pos
        else:
            ene = self.compute_batch(pos, box, forces).to(pos.dtype)
        keys = list(self.energies) + ["external"]
        cols = [self.column(k) for k in keys]
        if not explicit_forces:  # Q13: return the summed tensor
            return ene[:, cols].sum().reshape(1)
        pot = []
        for i in range(ene.shape[0]):
            row = ene[i]
            pot.append({k: row[c:c + 1] for k, c in zip(keys, cols)})
        if returnDetails:
            return pot
        return [sum(pp.values()) for pp in pot]
