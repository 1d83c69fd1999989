#!/usr/bin/env python
"""bench.py -- ALA10 replica-steps/s of the fused force + velocity-Verlet kernel (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--replicas R] [--inner S]

Workload (BASELINE.json configs[2]): ALA10 (104 atoms, CHAMBER prmtop parameters from
tests/golden/ala10.npz), NVE velocity Verlet, dt = 1 fs, plain all-pairs LJ + Coulomb, R = 8 192 replicas
PER GPU (weak scaling; replicas never interact), synthetic perturbed conformers: base frame r mod 101 +
0.05 A Gaussian noise (seed 1234), Maxwell-Boltzmann velocities at 300 K (seed 5678).  One bench "step"
= one Integrator.step(niter=S) call = one kernel launch advancing every replica S MD steps (default
S = 1000, energies recorded every 100).  value = replica-steps/s over all GPUs with state resident in HBM;
e2e = the same through the public API with pinned HOST buffers (H2D of pos/vel/forces and D2H of
pos/vel/observables inside the timed region).

--impl reference times the reference's CPU implementation of the same path on all host cores, each step being a
bounded sample of the same workload: the reference's OWN TorchMD_Fix/forces.py (copied unmodified into oracle/_ref/ by
tools/make_oracle_ref.py; git-ignored, shipped to the GPU box) under the restated upstream velocity-Verlet loop
(cpu_baseline.kind "reference"), or, where oracle/_ref is missing, the torch-op port oracle/torchmd_port.py ("port").

--workload forces is BASELINE.json configs[1] as a first-class line: one Forces.compute launch over 65 536 conformers
(energies in FP64 + forces; --no-energies for the force-only launch the autograd backward uses).
"""
import argparse
import json
import os
import statistics
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

FLOP_PER_REPLICA_STEP = 240824  # SURVEY.md 8d, ALA10 plain NVE (force-only evaluation: 236144)
FLOP_PER_REPLICA_EVAL = 236144
NOMINAL_FP32_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12  # 74.5
TIMEFACTOR = 48.88821
def hbm_view(R, S, energy_every, ms_per_launch):
    """Algorithmic HBM bytes of one integrate launch (state read + written once, observable records) against the
    measured copy bandwidth of MEASURED_PEAKS.json (driver-written), when that file is present."""
    nrep = S // energy_every if energy_every > 0 else 1
    nbytes = R * (2 * 3 * 104 * 12 + nrep * 11 * 8 + 4)
    peak = None
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "MEASURED_PEAKS.json")) as fh:
            peak = float(json.load(fh)["hbm_gbs"])
    except Exception:
        pass
    ach = nbytes / (ms_per_launch * 1e-3) / 1e9
    return {"algorithmic_bytes_per_launch": nbytes, "achieved": ach, "peak": peak, "unit": "GB/s",
            "frac": ach / peak if peak else None, "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peak else None}


NCU_DRAM_BYTES_PER_LAUNCH = 31.1e6  # profiles/r01_v22_integrate_kernel.md (v23 identical): 31.12 MB read + 0.05 MB written
BOLTZMAN = 0.001987191


def load_ala10():
    with np.load(os.path.join(ROOT, "tests", "golden", "ala10.npz")) as z:
        return {k: z[k] for k in z.files}


def synth_inputs(g, R, replica0=0):
    """Deterministic per global replica id: independent of how replicas are sharded over GPUs."""
    base = g["base_frames"]
    mass = g["par_masses"].reshape(-1).astype(np.float64)
    pos = np.empty((R, 104, 3), dtype=np.float32)
    vel = np.empty((R, 104, 3), dtype=np.float32)
    for k in range(R):
        r = replica0 + k
        rng = np.random.default_rng([1234, r])
        pos[k] = (base[r % len(base)] + 0.05 * rng.standard_normal((104, 3))).astype(np.float32)
        rngv = np.random.default_rng([5678, r])
        vel[k] = (np.sqrt(300.0 * BOLTZMAN / mass)[:, None] * rngv.standard_normal((104, 3))).astype(np.float32)
    return pos, vel


class ClockSampler:
    """Samples SM clock + throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                 "hw_power_brake": 0x80, "sync_boost": 0x10}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for n, bit in names.items():
                    if r & bit:
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.05)

    def __enter__(self):
        if self.nv is not None:
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thread is not None:
            self._thread.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


# ---------------------------------------------------------------------------------------------------
# CPU legs: the oracle is executed here only as the *reported baseline* (never on the product path)
# ---------------------------------------------------------------------------------------------------
_WORKER_CACHE = {}


def cpu_kind():
    """"reference": oracle/_ref holds the reference's own forces.py; "port": only the torch restatement is here."""
    from oracle import ref_arm

    return "reference" if ref_arm.available() else "port"


def _port_worker(args):
    """One host core: `nrep` replicas x `nsteps` MD steps (or, nsteps == 0, `nrep` Forces.compute evaluations) with
    the reference's own forces.py when oracle/_ref is present, else with the torch port (1 torch thread)."""
    import torch

    torch.set_num_threads(1)
    from molecular_dynamics_b200 import _cabi
    from molecular_dynamics_b200.parameters import Parameters
    from oracle import ref_arm
    from oracle import torchmd_port as port

    replica0, nrep, nsteps, dt, fkw = args
    use_ref = ref_arm.available()
    key = ("ref" if use_ref else "port", tuple(sorted(fkw.items())))
    if key not in _WORKER_CACHE:
        g = load_ala10()
        par = Parameters.from_arrays({k[4:]: v for k, v in g.items() if k.startswith("par_")})
        if use_ref:
            _WORKER_CACHE[key] = (g, par, ref_arm.module("forces").Forces(par, **fkw))
        else:
            _WORKER_CACHE[key] = (g, par, port.PortTopology(_cabi.PackedTopology(par, g["ava_idx"]), torch.float32))
    g, par, top = _WORKER_CACHE[key]
    pos, vel = synth_inputs(g, nrep, replica0)
    pos, vel = torch.tensor(pos), torch.tensor(vel)
    frc = torch.zeros_like(pos)
    box = torch.zeros((nrep, 3, 3))
    if use_ref:
        top.compute(pos, box, frc)
        t0 = time.perf_counter()
        if nsteps:
            ref_arm.integrate(top, par.masses, pos, vel, frc, box, nsteps, dt)
        else:
            top.compute(pos, box, frc, returnDetails=True)
    else:
        port.compute(top, pos, frc, **fkw)
        t0 = time.perf_counter()
        if nsteps:
            port.integrate(top, pos, vel, frc, nsteps, dt, **fkw)
        else:
            port.compute(top, pos, frc, **fkw)
    return time.perf_counter() - t0


def cpu_port_rate(cores, nrep_per_core, nsteps, pool=None, fkw=None):
    """replica-steps/s (nsteps == 0: replica-evaluations/s) of the CPU implementation over `cores` worker processes
    (max over workers' wall time)."""
    import multiprocessing as mp

    dt = 1.0 / TIMEFACTOR
    jobs = [(c * nrep_per_core, nrep_per_core, nsteps, dt, dict(fkw or {})) for c in range(cores)]
    own = pool is None
    if own:
        pool = mp.get_context("spawn").Pool(cores)
    try:
        t0 = time.perf_counter()
        times = pool.map(_port_worker, jobs)
        wall = time.perf_counter() - t0
    finally:
        if own:
            pool.close()
            pool.join()
    return cores * nrep_per_core * max(nsteps, 1) / max(times), max(times), wall


def cpu_c_oracle_rate(nrep, nsteps):
    """replica-steps/s of the FP64 C oracle (OpenMP over replicas, all host cores)."""
    import oracle
    from molecular_dynamics_b200 import _cabi
    from molecular_dynamics_b200.parameters import Parameters

    g = load_ala10()
    par = Parameters.from_arrays({k[4:]: v for k, v in g.items() if k.startswith("par_")})
    packed = _cabi.PackedTopology(par, g["ava_idx"])
    pos, vel = synth_inputs(g, nrep)
    f0, _ = oracle.forces(packed, pos)
    t0 = time.perf_counter()
    oracle.integrate(packed, pos, vel, f0, nsteps=nsteps, dt=1.0 / TIMEFACTOR)
    return nrep * nsteps / (time.perf_counter() - t0)


# identical in both arms' `config` (the driver compares the two dicts); the CPU arm's bounded sample is described in its
# cpu_baseline.sample and top-level "sample"
L2_NOTE = "GPU arm: flushed between timed steps (256 MiB fill, outside the event pair); CPU arm: not applicable"


def bench_config(args, R, S):
    """`config` of the JSON line: the same dict in both arms."""
    if args.workload == "forces":
        return {"workload": WORKLOAD_NAME_FORCES, "replicas_per_gpu": R, "mode": args.forces_mode,
                "energies": not args.no_energies, "natoms": 104, "l2": L2_NOTE}
    return {"workload": WORKLOAD_NAME_C4 if args.workload == "langevin_smd" else WORKLOAD_NAME, "replicas_per_gpu": R,
            "md_steps_per_launch": S, "energy_every": args.energy_every, "dt_fs": 1.0, "natoms": 104, "l2": L2_NOTE}


FORCES_MODES = {"plain": {}, "rf": {"cutoff": 9.0, "rfa": True, "switch_dist": 7.5}}


def run_reference(args):
    """--impl reference: the CPU path on all host cores, same metric/config, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp

    cores = os.cpu_count() or 1
    kind = cpu_kind()
    what = ("the reference's own TorchMD_Fix/forces.py (oracle/_ref, unmodified)" if kind == "reference"
            else "torch port of TorchMD_Fix/forces.py (oracle/torchmd_port.py)")
    if args.workload == "forces":
        nrep, nsteps, fkw = 256, 0, FORCES_MODES[args.forces_mode]  # 256 evaluations per core per bench step (~1 s)
        unit, metric = "replica-evaluations/s", "ALA10 replica-evaluations/s"
        sample = f"{cores} procs x {nrep} Forces.compute evaluations per bench step ({what}, fp32, 1 thread each)"
    else:
        nrep, nsteps, fkw = 4, 400, {}  # per core and per bench step: 1 600 replica-steps x ~1.3 ms = ~2 s of CPU work
        unit, metric = "replica-steps/s", "ALA10 replica-steps/s"
        sample = (f"{cores} procs x {nrep} replicas x {nsteps} MD steps per bench step ({what} + upstream VV loop, fp32, "
                  f"1 thread each)")
    pool = mp.get_context("spawn").Pool(cores)
    try:
        for _ in range(max(1, min(args.warmup, 1))):
            cpu_port_rate(cores, nrep, nsteps, pool, fkw)
        units, wall = 0, 0.0
        for _ in range(args.steps):
            _, slowest, _ = cpu_port_rate(cores, nrep, nsteps, pool, fkw)
            wall += slowest  # slowest worker's compute time: the step ends when the last core finishes
            units += cores * nrep * max(nsteps, 1)
    finally:
        pool.close()
        pool.join()
    value = units / wall
    cfg = bench_config(args, args.replicas, args.inner)
    print(json.dumps({
        "impl": "reference", "metric": metric, "value": value, "unit": unit,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * wall / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": cfg, "sample": sample,
        "cpu_baseline": {"value": value, "unit": unit, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


WORKLOAD_NAME_C4 = ("configs[3]: ALA10 Langevin 300 K (gamma 1/ps) with the steered-MD end-to-end restraint (atoms 3 / 98, "
                    "fk 1, centre 12 -> 34 A), fused persistent kernel, dt=1fs, plain all-pairs LJ+Coulomb")
WORKLOAD_NAME = ("configs[2]: ALA10 NVE velocity-Verlet, 8192 replicas per GPU, fused persistent kernel, "
                 "dt=1fs, plain all-pairs LJ+Coulomb")
WORKLOAD_NAME_FORCES = ("configs[1]: ALA10 batched energy+force evaluation (Forces.compute) over 65536 synthetic perturbed "
                        "conformers per GPU, one launch")


# ---------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    import __graft_entry__ as entry

    if not os.path.exists(entry.LIB):
        entry.build()
    from molecular_dynamics_b200 import Forces, Integrator, Parameters, System, _cabi

    lib = _cabi.load_library()
    g = load_ala10()
    par = Parameters.from_arrays({k[4:]: v for k, v in g.items() if k.startswith("par_")}, device=dev)
    R, S = args.replicas, args.inner
    replica0 = rank * R
    pos_h, vel_h = synth_inputs(g, R, replica0)
    config4 = args.workload == "langevin_smd"
    if config4:  # BASELINE.json configs[3]: Langevin 300 K, gamma 1/ps, SMD restraint on atoms 3 / 98 moving 12 -> 34 A
        from molecular_dynamics_b200 import SMDBias
        total_steps = (args.warmup + args.steps + 8) * S
        forces = Forces(par, external=SMDBias([3], [98], fk=1.0, cent_0=12.0, rate_per_step=22.0 / total_steps))
    else:
        forces = Forces(par)
    forces.check_nan = False
    system = System(104, R, torch.float, dev)
    system.pos[:] = torch.tensor(pos_h, device=dev)
    system.vel[:] = torch.tensor(vel_h, device=dev)
    forces.compute_batch(system.pos, system.box, system.forces)
    frc_h = system.forces.cpu().numpy()
    if config4:
        integ = Integrator(system, forces, timestep=1.0, device=dev, gamma=1.0, T=300.0, seed=42)
    else:
        integ = Integrator(system, forces, timestep=1.0, device=dev)
    integ.energy_every = args.energy_every
    integ.replica0 = replica0
    integ.check_nan = False

    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)  # > 126 MB L2
    from molecular_dynamics_b200.sharding import gather_observables

    def one_step():
        integ.step(niter=S)
        if world > 1:  # reporting interval: all-gather the per-replica observables (the only collective)
            gather_observables(integ.observables[-1], world * R)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(args.warmup):
        one_step()
    barrier()

    # ---- device-timed region: K steps, each bracketed by CUDA events, L2 flushed in between ----------
    launches0 = lib.b200md_launch_count()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    with ClockSampler(local) as clocks:
        barrier()
        for a, b in ev:
            flush.fill_(1)
            a.record()
            one_step()
            b.record()
        barrier()
    launches = lib.b200md_launch_count() - launches0
    per_launch = [a.elapsed_time(b) for a, b in ev]
    ms = sum(per_launch)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    units = world * R * S * args.steps
    value = units / (ms * 1e-3)
    nan_count = int(torch.isnan(system.pos).any(dim=2).any(dim=1).sum().item())

    # ---- end-to-end through the public API with pinned host buffers ----------------------------------
    hp = torch.tensor(pos_h).pin_memory()
    hv = torch.tensor(vel_h).pin_memory()
    hf = torch.tensor(frc_h).pin_memory()
    out_p, out_v = torch.empty_like(hp).pin_memory(), torch.empty_like(hv).pin_memory()
    out_o = torch.empty((max(1, S // args.energy_every) if args.energy_every else 1, R, _cabi.NUM_OBSERVABLES),
                        dtype=torch.float64).pin_memory()
    h2d = hp.numel() * 4 * 3
    d2h = out_p.numel() * 4 * 2 + out_o.numel() * 8

    forces.check_nan = integ.check_nan = True  # the public call as a user makes it: the NaN / range flag read is inside

    def e2e_step():
        system.pos.copy_(hp, non_blocking=True)
        system.vel.copy_(hv, non_blocking=True)
        system.forces.copy_(hf, non_blocking=True)
        ekin, pot, temp = integ.step(niter=S)  # returns host Ekin/T (device->host read inside)
        out_p.copy_(system.pos, non_blocking=True)
        out_v.copy_(system.vel, non_blocking=True)
        out_o.copy_(integ.observables, non_blocking=True)
        torch.cuda.synchronize()
        return float(ekin[0])

    e2e_steps = max(2, min(args.steps, 5))
    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * R * S * e2e_steps / float(t.item())
    forces.check_nan = integ.check_nan = False

    # ---- strong scaling: a FIXED total (configs[2]'s 8 192 replicas) split over the ranks, same launches ------------
    strong = None
    if args.strong_total > 0 and not config4:
        from molecular_dynamics_b200.sharding import shard_range
        lo, hi = shard_range(args.strong_total, rank, world)
        Rs = hi - lo
        ps, vs = synth_inputs(g, Rs, lo)
        sys_s = System(104, Rs, torch.float, dev)
        sys_s.pos[:] = torch.tensor(ps, device=dev)
        sys_s.vel[:] = torch.tensor(vs, device=dev)
        forces.compute_batch(sys_s.pos, sys_s.box, sys_s.forces)
        integ_s = Integrator(sys_s, forces, timestep=1.0, device=dev)
        integ_s.energy_every, integ_s.replica0, integ_s.check_nan = args.energy_every, lo, False

        def strong_step():
            integ_s.step(niter=S)
            if world > 1:
                gather_observables(integ_s.observables[-1], args.strong_total)

        for _ in range(2):
            strong_step()
        nst = max(3, min(args.steps, 5))
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(nst)]
        barrier()
        for a, b in evs:
            flush.fill_(1)
            a.record()
            strong_step()
            b.record()
        barrier()
        ts = torch.tensor([sum(a.elapsed_time(b) for a, b in evs)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ts, op=dist.ReduceOp.MAX)
        strong = {"replicas_total": args.strong_total, "replicas_per_gpu": Rs, "md_steps_per_launch": S, "steps": nst,
                  "value": args.strong_total * S * nst / (float(ts.item()) * 1e-3), "unit": "replica-steps/s",
                  "ms_per_step": float(ts.item()) / nst, "scaling": "strong",
                  "note": "the same fused launches over a fixed total of replicas split across the ranks; device-timed, "
                          "max over ranks, L2 flushed between launches"}
        del sys_s, integ_s

    # ---- configs[4]: replica-count sweep, TOTAL replicas split over the ranks (1 000 NVE steps per launch) ----------
    sweep = []
    if args.sweep and not config4:
        from molecular_dynamics_b200.sharding import shard_range
        for Rt in [int(v) for v in args.sweep.split(",") if v]:
            lo, hi = shard_range(Rt, rank, world)
            Rs = hi - lo
            idx = (np.arange(lo, hi) % R)  # inputs of global replica g = those of replica g mod R of the main run
            sys_w = System(104, Rs, torch.float, dev)
            sys_w.pos[:] = torch.tensor(pos_h[idx], device=dev)
            sys_w.vel[:] = torch.tensor(vel_h[idx], device=dev)
            forces.compute_batch(sys_w.pos, sys_w.box, sys_w.forces)
            integ_w = Integrator(sys_w, forces, timestep=1.0, device=dev)
            integ_w.energy_every, integ_w.replica0, integ_w.check_nan = args.energy_every, lo, False

            def sweep_step():
                integ_w.step(niter=S)
                if world > 1:
                    gather_observables(integ_w.observables[-1], Rt)

            sweep_step()
            nsw = 3
            evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(nsw)]
            barrier()
            for a, b in evs:
                flush.fill_(1)
                a.record()
                sweep_step()
                b.record()
            barrier()
            tw = torch.tensor([sum(a.elapsed_time(b) for a, b in evs)], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(tw, op=dist.ReduceOp.MAX)
            sweep.append({"replicas_total": Rt, "replicas_per_gpu": Rs, "value": Rt * S * nsw / (float(tw.item()) * 1e-3),
                          "unit": "replica-steps/s", "ms_per_launch": float(tw.item()) / nsw})
            del sys_w, integ_w

    # ---- secondary workloads (BASELINE.json configs[1] and configs[3]); reported under "extra", N = 1 only -----
    extra = {}
    if world == 1 and not args.no_extra:
        def timed(fn, reps):
            """seconds per call: median over three event-timed groups of `reps` calls (the shared boxes show a host
            interference spike of tens of ms about once a second, which a single short group can catch)"""
            fn()
            torch.cuda.synchronize()
            groups = []
            for _ in range(3):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                for _ in range(reps):
                    fn()
                b.record()
                torch.cuda.synchronize()
                groups.append(a.elapsed_time(b) * 1e-3 / reps)
            return statistics.median(groups)

        # configs[1]: batched energy + force evaluation over 65 536 conformers (Forces.compute, one launch)
        R2 = 65536
        reps2 = -(-R2 // R)
        big = torch.tensor(pos_h, device=dev).repeat(reps2, 1, 1)[:R2].contiguous()
        big += 0.01 * torch.randn(big.shape, device=dev, generator=torch.Generator(device=dev).manual_seed(7))
        sys2 = System(104, R2, torch.float, dev)
        sys2.pos[:] = big
        for label, kw in (("plain", {}), ("cutoff9_rfa_switch7.5", {"cutoff": 9.0, "rfa": True, "switch_dist": 7.5})):
            f2 = Forces(par, **kw)
            f2.check_nan = False
            t2 = timed(lambda: f2.compute_batch(sys2.pos, sys2.box, sys2.forces), 5)
            flop = FLOP_PER_REPLICA_EVAL + (8 * 4820 if kw else 0)
            t2f = timed(lambda: f2.compute_batch(sys2.pos, sys2.box, sys2.forces, energies=False), 5)
            extra[f"forces_compute_65536_{label}"] = {"value": R2 / t2, "unit": "replica-evaluations/s", "ms_per_launch": 1e3 * t2,
                                                      "tflops": R2 / t2 * flop / 1e12,
                                                      "note": "forces fp32 + all 8 energy terms in FP64 (energy pass)",
                                                      "force_only": {"value": R2 / t2f, "ms_per_launch": 1e3 * t2f,
                                                                     "tflops": R2 / t2f * flop / 1e12}}
        del sys2, big
        # configs[3]: Langevin 300 K, gamma = 1/ps, steered-MD end-to-end restraint moving 12 -> 34 A
        from molecular_dynamics_b200 import SMDBias
        nst = min(S, 500)
        f3 = Forces(par, external=SMDBias([3], [98], fk=1.0, cent_0=12.0, cent_1=34.0, rate_per_step=22.0 / 500000))
        f3.check_nan = False
        sys3 = System(104, R, torch.float, dev)
        sys3.pos[:] = torch.tensor(pos_h, device=dev)
        sys3.vel[:] = torch.tensor(vel_h, device=dev)
        f3.compute_batch(sys3.pos, sys3.box, sys3.forces)
        integ3 = Integrator(sys3, f3, timestep=1.0, device=dev, gamma=1.0, T=300.0, seed=42)
        integ3.energy_every, integ3.check_nan = args.energy_every, False
        t3 = timed(lambda: integ3.step(niter=nst), 2)
        extra["langevin_smd"] = {"value": R * nst / t3, "unit": "replica-steps/s", "replicas": R, "md_steps_per_launch": nst}
        # the fused NVE step with the settings every script of the reference uses (cutoff 9 A, reaction field, switch 7.5 A)
        sys7 = System(104, R, torch.float, dev)
        sys7.pos[:] = torch.tensor(pos_h, device=dev)
        sys7.vel[:] = torch.tensor(vel_h, device=dev)
        f7 = Forces(par, cutoff=9.0, rfa=True, switch_dist=7.5)
        f7.check_nan = False
        f7.compute_batch(sys7.pos, sys7.box, sys7.forces)
        integ7 = Integrator(sys7, f7, timestep=1.0, device=dev)
        integ7.energy_every, integ7.check_nan = args.energy_every, False
        t7 = timed(lambda: integ7.step(niter=nst), 2)
        extra["nve_cutoff9_rfa_switch7.5"] = {"value": R * nst / t7, "unit": "replica-steps/s", "replicas": R,
                                              "md_steps_per_launch": nst, "tflops": R * nst / t7 * (FLOP_PER_REPLICA_STEP + 8 * 4820) / 1e12}
        del sys7, integ7
        # snapshot-write path (north star: "achieved HBM GB/s for the snapshot-write path"): positions of every replica
        # written every 10 / every step (observables only after the last step, so that the cost seen is the write path's)
        for every in (10, 1):
            sys6 = System(104, R, torch.float, dev)
            sys6.pos[:] = torch.tensor(pos_h, device=dev)
            sys6.vel[:] = torch.tensor(vel_h, device=dev)
            forces.compute_batch(sys6.pos, sys6.box, sys6.forces)
            integ6 = Integrator(sys6, forces, timestep=1.0, device=dev)
            integ6.energy_every, integ6.snapshot_every, integ6.check_nan = 0, every, False
            n6 = 200
            t6 = timed(lambda: integ6.step(niter=n6), 2)
            wbytes = (n6 // every) * R * 104 * 12
            extra[f"snapshots_every_{every}"] = {"value": R * n6 / t6, "unit": "replica-steps/s", "hbm_write_GBps": wbytes / t6 / 1e9,
                                                "bytes_per_snapshot_per_replica": 104 * 12}
            del sys6, integ6
        # the ML callers' loss (cgan.py:345-360): per-term energies + torsion angles of a generated batch, forward and
        # backward through the batched Energy adapter (the reference rebuilds an Energy object per sample: 41-105
        # evaluations/s on a V100)
        from molecular_dynamics_b200 import Energy
        colvar = {"name": "E2End Harm", "fk": 1.0, "cent_0": 12.0, "cent_1": 34.0, "T": 1.0, "group1": [3], "group2": [98]}
        en = Energy(parameters=par, colvar=colvar, device=dev)
        en.forces.check_nan = False
        B5 = 4096
        x5 = torch.tensor(pos_h[:B5] if len(pos_h) >= B5 else np.resize(pos_h, (B5, 104, 3)), device=dev).requires_grad_(True)
        t5 = torch.linspace(0.0, 1.0, B5, device=dev)

        def adapter_step():
            x5.grad = None
            ene, phis = en.calc_energy_batch(x5, t5)
            loss = sum(v for k, v in ene.items() if k != "E2End Harm").sum() / B5 + 200.0 * ene["E2End Harm"].sum() / B5
            loss.backward()

        t5s = timed(adapter_step, 5)
        extra["energy_adapter_forward_backward"] = {"value": B5 / t5s, "unit": "conformers/s", "batch": B5,
                                                    "ms_per_batch": 1e3 * t5s, "launches_per_batch": 4}
        # the data generator's own protocol (smd.namd: rigidBonds ALL, 2 fs, Langevin 300 K, SMD): SHAKE / RATTLE in the kernel
        sys4 = System(104, R, torch.float, dev)
        sys4.pos[:] = torch.tensor(pos_h, device=dev)
        sys4.vel[:] = torch.tensor(vel_h, device=dev)
        f3.compute_batch(sys4.pos, sys4.box, sys4.forces)
        integ4 = Integrator(sys4, f3, timestep=2.0, device=dev, gamma=1.0, T=300.0, seed=43, rigid_bonds=True)
        integ4.energy_every, integ4.check_nan = args.energy_every, False
        integ4.step(niter=20)  # the synthetic start is not constrained: let the first steps project it
        t4 = timed(lambda: integ4.step(niter=nst), 2)
        extra["langevin_smd_rigid_bonds_2fs"] = {"value": R * nst / t4, "unit": "replica-steps/s", "replicas": R,
                                                 "md_steps_per_launch": nst, "dt_fs": 2.0,
                                                 "ns_per_day_aggregate": R * nst / t4 * 2.0 * 86400 * 1e-6}

        # SURVEY 8f N4, solutes above 128 atoms (csrc/b200md_large.cuh): a bundle of 10 deca-alanine chains as ONE
        # 1 040-atom topology (Parameters.tiled), chains 40 A apart on a line; Forces.compute and Langevin steps
        copies, RL = 10, 256
        parL = par.tiled(copies)
        NL = 104 * copies
        shift = torch.arange(copies, device=dev, dtype=torch.float32).repeat_interleave(104)[None, :, None] * \
            torch.tensor([40.0, 0.0, 0.0], device=dev)
        posL = torch.tensor(pos_h[:RL], device=dev).repeat(1, copies, 1) + shift
        sysL = System(NL, RL, torch.float, dev)
        sysL.pos[:] = posL
        sysL.vel[:] = torch.tensor(vel_h[:RL], device=dev).repeat(1, copies, 1)
        fL = Forces(parL)
        fL.check_nan = False
        npairL = int(fL.ava_idx.shape[0])
        tL = timed(lambda: fL.compute_batch(sysL.pos, sysL.box, sysL.forces), 3)
        flopL = 36 * npairL + copies * (FLOP_PER_REPLICA_EVAL - 36 * 4820)
        extra["large_solute_1040_atoms_forces_compute"] = {
            "value": RL / tL, "unit": "replica-evaluations/s", "replicas": RL, "natoms": NL, "unique_pairs": npairL,
            "ms_per_call": 1e3 * tL, "tflops": RL / tL * flopL / 1e12,
            "note": "half-shell pair rectangles + bonded windows (fp32 forces) and all 8 energy terms in FP64 (two more kernels on a second stream)"}
        tLf = timed(lambda: fL.compute_batch(sysL.pos, sysL.box, sysL.forces, energies=False), 3)
        extra["large_solute_1040_atoms_forces_compute"]["force_only"] = {"value": RL / tLf, "ms_per_call": 1e3 * tLf,
                                                                       "tflops": RL / tLf * flopL / 1e12}
        integL = Integrator(sysL, fL, timestep=1.0, device=dev, gamma=1.0, T=300.0, seed=44)
        integL.energy_every, integL.check_nan = 10, False
        tL2 = timed(lambda: integL.step(niter=20), 2)
        extra["large_solute_1040_atoms_langevin"] = {"value": RL * 20 / tL2, "unit": "replica-steps/s", "replicas": RL,
                                                    "natoms": NL, "md_steps_per_call": 20, "energy_every": 10,
                                                    "tflops": RL * 20 / tL2 * (flopL + 60 * NL) / 1e12}
        integL.energy_every = 100
        tL3 = timed(lambda: integL.step(niter=100), 2)
        extra["large_solute_1040_atoms_langevin"]["energy_every_100"] = {"value": RL * 100 / tL3, "md_steps_per_call": 100,
                                                                        "us_per_step": 1e6 * tL3 / 100}
        # the same bundle with the 9 A cutoff / reaction field / 7.5 A switch (the mode the tile-box skips exist for)
        fLc = Forces(parL, cutoff=9.0, rfa=True, switch_dist=7.5)
        fLc.check_nan = False
        integLc = Integrator(sysL, fLc, timestep=1.0, device=dev, gamma=1.0, T=300.0, seed=45)
        integLc.energy_every, integLc.check_nan = 100, False
        integLc.step(niter=20)
        tL4 = timed(lambda: integLc.step(niter=100), 2)
        extra["large_solute_1040_atoms_langevin_cutoff9_rfa_switch7.5"] = {
            "value": RL * 100 / tL4, "unit": "replica-steps/s", "replicas": RL, "natoms": NL, "md_steps_per_call": 100,
            "energy_every": 100, "us_per_step": 1e6 * tL4 / 100}
        del sysL, posL

    if rank == 0:
        peak = lib.b200md_fp32_peak_tflops(local, 4096)
        flop_step = FLOP_PER_REPLICA_STEP + (15 * 104 if config4 else 0)  # +15 N_atom for the Langevin kick
        achieved = value / world * flop_step / 1e12  # per GPU
        line = {
            "metric": "ALA10 replica-steps/s", "value": value, "unit": "replica-steps/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": bench_config(args, R, S),
            "detail": {"ns_per_day_aggregate": value * 1.0 * 86400 * 1e-6,
                       "ms_per_step_median_rank0": statistics.median(per_launch), "ms_per_step_max_rank0": max(per_launch),
                       "replicas_with_nan_after_run": nan_count},
            "clocks": clocks.summary(),
            "e2e": {"value": e2e_value, "unit": "replica-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": e2e_steps},
            "gpu_launches": int(launches),
            "roofline": {"bound": "fp32", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                         "frac": achieved / peak if peak > 0 else None, "traffic": NCU_DRAM_BYTES_PER_LAUNCH,
                         "traffic_source": "ncu --set full of one 50-step launch, profiles/ (dram read + write; the state is read once per launch whatever its length)",
                         "peak_source": "builder-measured FMA peak, live in this run: b200md_fp32_peak_tflops (dependent-FMA kernel, "
                                        "all SMs); MEASURED_PEAKS.json carries no FP32 CUDA-core figure, see frac_of_nominal",
                         "nominal_peak": NOMINAL_FP32_TFLOPS, "frac_of_nominal": achieved / NOMINAL_FP32_TFLOPS,
                         "flop_per_replica_step": flop_step, "kernel": "integrate_kernel<0>",
                         "kernel_ms_per_launch": ms / args.steps,
                         "bound_note": "FP32 CUDA-core kernel: neither the HBM nor the tensor-core roofline binds it; the "
                                       "HBM view of the same launch is given under 'hbm'",
                         "hbm": hbm_view(R, S, args.energy_every, ms / args.steps)},
        }
        if strong is not None:
            extra["strong_scaling"] = strong
        if sweep:
            extra["replica_sweep"] = sweep
        if extra:
            line["extra"] = extra
        if world == 1 and not args.no_cpu:
            cores = os.cpu_count() or 1
            rate, slowest, _ = cpu_port_rate(cores, 4, 3000)
            kind = cpu_kind()
            what = ("the reference's own TorchMD_Fix/forces.py (oracle/_ref, unmodified)" if kind == "reference"
                    else "torch port of TorchMD_Fix/forces.py (oracle/torchmd_port.py)")
            line["cpu_baseline"] = {
                "value": rate, "unit": "replica-steps/s", "cores": cores, "kind": kind,
                "sample": f"{cores} procs x 4 replicas x 3000 MD steps of the same workload, {what} + upstream VV loop, "
                          f"fp32, 1 thread each; slowest worker {slowest:.1f} s"}
            try:
                line["cpu_baseline_c_oracle"] = {
                    "value": cpu_c_oracle_rate(64 * cores, 20), "unit": "replica-steps/s", "cores": cores,
                    "kind": "port", "sample": f"{64 * cores} replicas x 20 MD steps, oracle/md_oracle.c FP64, OpenMP"}
            except Exception as exc:  # the C oracle is optional here
                line["cpu_baseline_c_oracle"] = {"error": str(exc)}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def run_ours_forces(args):
    """--workload forces: BASELINE.json configs[1].  One bench step = one Forces.compute launch over R conformers
    (forces_kernel<MODE>: fp32 forces + the FP64 energy pass); R = 65 536 per GPU, replicas sharded over the ranks with
    no collective on the data path (the per-replica energies are all-gathered as the reporting step)."""
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import __graft_entry__ as entry

    if not os.path.exists(entry.LIB):
        entry.build()
    from molecular_dynamics_b200 import Forces, Parameters, System, _cabi
    from molecular_dynamics_b200.sharding import gather_observables

    lib = _cabi.load_library()
    g = load_ala10()
    par = Parameters.from_arrays({k[4:]: v for k, v in g.items() if k.startswith("par_")}, device=dev)
    R = args.replicas
    want_e = not args.no_energies
    pos_h, _ = synth_inputs(g, R, rank * R)
    kw = FORCES_MODES[args.forces_mode]
    F = Forces(par, **kw)
    F.check_nan = False
    system = System(104, R, torch.float, dev)
    system.pos[:] = torch.tensor(pos_h, device=dev)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)

    def one_step():
        e = F.compute_batch(system.pos, system.box, system.forces, energies=want_e)
        if world > 1 and e is not None:
            gather_observables(e, world * R)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(args.warmup):
        one_step()
    barrier()
    launches0 = lib.b200md_launch_count()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    with ClockSampler(local) as clocks:
        barrier()
        for a, b in ev:
            flush.fill_(1)
            a.record()
            one_step()
            b.record()
        barrier()
    launches = lib.b200md_launch_count() - launches0
    per_launch = [a.elapsed_time(b) for a, b in ev]
    t = torch.tensor([sum(per_launch)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    value = world * R * args.steps / (ms * 1e-3)

    # end to end: pinned host positions in, forces + energies back to the host, NaN flag read inside
    hp = torch.tensor(pos_h).pin_memory()
    out_f = torch.empty_like(hp).pin_memory()
    out_e = torch.empty((R, 8), dtype=torch.float64).pin_memory()
    F.check_nan = True

    def e2e_step():
        # the public call for a batch in host memory: chunked copies on their own streams overlap both PCIe directions
        # with the kernels (Forces.compute_batch_host); synchronous, NaN flag read inside
        F.compute_batch_host(hp, out_f, out_e if want_e else None, energies=want_e, device=dev)

    e2e_steps = max(3, min(args.steps, 10))
    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * R * e2e_steps / float(t.item())
    if rank == 0:
        peak = lib.b200md_fp32_peak_tflops(local, 4096)
        flop = FLOP_PER_REPLICA_EVAL + (8 * 4820 if kw else 0)
        achieved = value / world * flop / 1e12
        hbm_bytes = R * (2 * 104 * 12 + (64 if want_e else 0) + 4)
        line = {
            "metric": "ALA10 replica-evaluations/s", "value": value, "unit": "replica-evaluations/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": bench_config(args, R, 0),
            "detail": {"ms_per_step_median_rank0": statistics.median(per_launch)},
            "clocks": clocks.summary(),
            "e2e": {"value": e2e_value, "unit": "replica-evaluations/s", "h2d_bytes_per_step": hp.numel() * 4,
                    "d2h_bytes_per_step": out_f.numel() * 4 + (out_e.numel() * 8 if want_e else 0), "steps": e2e_steps},
            "gpu_launches": int(launches),
            "roofline": {"bound": "fp32", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                         "frac": achieved / peak if peak > 0 else None, "traffic": None,
                         "peak_source": "builder-measured FMA peak, live in this run (b200md_fp32_peak_tflops)",
                         "nominal_peak": NOMINAL_FP32_TFLOPS, "frac_of_nominal": achieved / NOMINAL_FP32_TFLOPS,
                         "flop_per_replica_evaluation": flop, "kernel": f"forces_kernel<{1 if kw else 0}>",
                         "kernel_ms_per_launch": ms / args.steps,
                         "hbm": {"algorithmic_bytes_per_launch": hbm_bytes, "achieved_GBps": hbm_bytes / (ms / args.steps * 1e-3) / 1e9},
                         "bound_note": "FP32 CUDA-core kernel (+ an FP64 energy pass when energies are on); the flop count is "
                                       "SURVEY 8d's force-evaluation figure and gives the FP64 energy arithmetic no credit"},
        }
        if world == 1 and not args.no_cpu:
            cores = os.cpu_count() or 1
            rate, slowest, _ = cpu_port_rate(cores, 1024, 0, fkw=kw)
            kind = cpu_kind()
            line["cpu_baseline"] = {"value": rate, "unit": "replica-evaluations/s", "cores": cores, "kind": kind,
                                    "sample": f"{cores} procs x 1024 Forces.compute evaluations of the same conformers "
                                              f"({'oracle/_ref forces.py, unmodified' if kind == 'reference' else 'torch port'}, fp32, "
                                              f"1 thread each); slowest worker {slowest:.1f} s"}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--replicas", type=int, default=8192, help="replicas per GPU")
    ap.add_argument("--inner", type=int, default=1000, help="MD steps per launch (Integrator.step niter)")
    ap.add_argument("--energy-every", type=int, default=100)
    ap.add_argument("--workload", default="nve", choices=["nve", "langevin_smd", "forces"],
                    help="nve = configs[2] (the metric's configuration, default); langevin_smd = configs[3] "
                         "(use --replicas 4096 at 8 GPUs for its 32 768 replicas); forces = configs[1] (one "
                         "Forces.compute launch over --replicas conformers, default 65 536)")
    ap.add_argument("--forces-mode", default="plain", choices=sorted(FORCES_MODES),
                    help="--workload forces: plain all-pairs, or rf = cutoff 9 A + reaction field + switch 7.5 A")
    ap.add_argument("--no-energies", action="store_true", help="--workload forces: force-only launches (no FP64 energy pass)")
    ap.add_argument("--sweep", default="", help="comma-separated TOTAL replica counts of the configs[4] sweep (split over the ranks), "
                                                "e.g. 1024,4096,16384,65536,262144; reported under extra.replica_sweep")
    ap.add_argument("--strong-total", type=int, default=8192,
                    help="replicas of the strong-scaling leg, split over the ranks (0: skip the leg)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-extra", action="store_true", help="skip the secondary workloads (configs[1], configs[3])")
    args = ap.parse_args()
    if args.workload == "forces" and args.replicas == 8192:
        args.replicas = 65536
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "forces":
        run_ours_forces(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
