// b200md.cu -- kernels + C-ABI (include/b200md.h) of the B200-native TorchMD force-field/integrator path.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -shared -Xcompiler -fPIC
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <mutex>
#include <string>
#include <vector>

#include "b200md_forces.cuh"
#include "b200md_large.cuh"
#include "b200md_minimize.cuh"

using namespace b200md;

// =================================================================================================
// kernels
// =================================================================================================
namespace b200md {

// Stage the replica-independent tables into shared memory once per CTA (the CTA then loops over
// replicas).  Term switches (forces.py:40-50) are applied here: "lj" off -> A = B = 0.
template <bool FIXED>
__device__ __forceinline__ void stage_tables(const Smem& S, const DevTopo& T, const ForceParams& P, int tid) {
  for (int t = 0; t < T.ntypes; ++t)
    for (int j = tid; j < kTabStride; j += kThreads) {
      const float2 v = j < kMaxAtoms ? __ldg(T.ab + t * kAbStride + j) : make_float2(0.f, 0.f);
      S.ta[t * kTabStride + j] = v.x * P.tab_a;
      S.tb[t * kTabStride + j] = v.y * P.tab_b;
    }
  for (int k = tid; k < kMaxAtoms; k += kThreads) {
    S.ab_row_off[k] = __ldg(T.ab_row_off + k);
    S.diagmask[k] = __ldg(T.diagmask + k);
    S.pos[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    S.sx[k] = S.sx[kMaxAtoms + k] = S.sx[2 * kMaxAtoms + k] = S.sx[3 * kMaxAtoms + k] = 0.0f;
  }
  for (int k = tid; k < kMaxChunks * kMaxAtoms / 4; k += kThreads)
    ((uint32_t*)S.mask8)[k] = __ldg((const uint32_t*)T.mask8 + k);
  for (int k = tid; k <= kMaxAtoms; k += kThreads) S.gstart[k] = __ldg(T.gather_start + k);
  for (int k = tid; k < kWarps * kMaxTasks; k += kThreads) S.tasks[k] = __ldg(T.tasks + k);
  if (tid < 8) { S.smd[tid] = 0.0f; S.box[tid] = 0.0f; }
  // slots of disabled / dropped terms are never written: keep them at zero for the gather
  if (FIXED) {
    for (int k = tid; k < 3 * kMaxAtoms; k += kThreads) S.fbi[k] = 0;
  } else {
    for (int k = tid; k < 3 * S.fbs; k += kThreads) S.fb[k] = 0.0f;
  }
  // chunks without a tile task (no non-bonded term enabled) must read as zero in the gather
  for (int k = tid; k < 3 * kMaxAtoms; k += kThreads) S.fj[k] = 0.0f;
  for (int k = tid; k < kWarps * 3 * kMaxAtoms; k += kThreads) S.fi[k] = 0.0f;  // i-side partials: armed here, re-armed by the gather
}

// Positions in double for the energy pass (S.pd overlays vel + sx: callers put a barrier on either side).
__device__ __forceinline__ void stage_pd(const Smem& S, int tid, float x, float y, float z, float q) {
  S.pd[tid] = (double)x; S.pd[kMaxAtoms + tid] = (double)y; S.pd[2 * kMaxAtoms + tid] = (double)z;
  S.pd[3 * kMaxAtoms + tid] = (double)q * 18.222615429025844;  // q sqrt(K), K = 332.06371307417066: pair products are K qi qj
}

// Forces.compute (forces.py:82-344), the FORCES of a batch of replicas: one CTA per replica (grid-stride).  The energies
// are a kernel of their own (energy_kernel below), launched behind this one when the caller wants them.
template <int MODE>
__global__ void __launch_bounds__(kThreads, 4)
forces_kernel(const DevTopo T, const ForceParams P, const long long R, const float* __restrict__ pos,
              const float* __restrict__ box, float* __restrict__ forces, uint32_t* __restrict__ status) {
  extern __shared__ __align__(16) char smem_raw[];
  Smem S;
  smem_layout(&S, smem_raw, T.ntypes, T.nslots, false);
  // pinned in registers: under register pressure the compiler otherwise re-reads SR_TID.X (a ~25-cycle S2R)
  // at the top of every task
  int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  asm volatile("" : "+r"(tid), "+r"(warp), "+r"(lane));
  const int N = T.natoms;
  stage_tables<false>(S, T, P, tid);
  const float q = tid < N ? __ldg(T.charge + tid) * P.el_on : 0.0f;
  // the positions of the CTA's next replica are fetched while the current one is evaluated
  float nx = 0.f, ny = 0.f, nz = 0.f;
  if (tid < N && (long long)blockIdx.x < R) {
    const float* p = pos + ((long long)blockIdx.x * N + tid) * 3;
    nx = __ldg(p); ny = __ldg(p + 1); nz = __ldg(p + 2);
  }
  for (long long r = blockIdx.x; r < R; r += gridDim.x) {
    __syncthreads();  // previous replica fully consumed (and tables staged)
    uint32_t flag = 0;
    const float x = nx, y = ny, z = nz;
    if (tid < N) {
      if (x != x || y != y || z != z) flag |= B200MD_FLAG_NAN_POS;
      S.pos[tid] = make_float4(x, y, z, q);
      S.sx[tid] = x; S.sx[kMaxAtoms + tid] = y; S.sx[2 * kMaxAtoms + tid] = z; S.sx[3 * kMaxAtoms + tid] = q;
      if (r + gridDim.x < R) {
        const float* p = pos + ((r + gridDim.x) * N + tid) * 3;
        nx = __ldg(p); ny = __ldg(p + 1); nz = __ldg(p + 2);
      }
    }
    // box diagonal (forces.py:107); all zero = vacuum, no wrap (forces.py:359)
    const float b0 = box ? __ldg(box + r * 9) : 0.0f, b1 = box ? __ldg(box + r * 9 + 4) : 0.0f, b2 = box ? __ldg(box + r * 9 + 8) : 0.0f;
    const bool pbc = b0 != 0.0f || b1 != 0.0f || b2 != 0.0f;
    if (tid < 3) {
      const float b = tid == 0 ? b0 : (tid == 1 ? b1 : b2);
      S.box[tid] = -b;
      S.box[4 + tid] = b != 0.0f ? 1.0f / b : 0.0f;
    }
    if (status != nullptr) {
      if (tid == 0) status[r] = 0;
    }
    __syncthreads();
    if (flag && status != nullptr) atomicOr(status + r, flag);
    if (pbc) run_tasks<MODE, false, true>(S, T, P, warp, lane);
    else run_tasks<MODE, false, false>(S, T, P, warp, lane);
    if (P.smd_on && tid == kThreads - 1) smd_task(S, P, P.smd_centers ? __ldg(P.smd_centers + r) : P.smd_c0);
    __syncthreads();
    if (tid < N) {
      float fx, fy, fz;
      gather_forces<false>(S, P, tid, fx, fy, fz, flag);
      float* f = forces + (r * N + tid) * 3;
      f[0] = fx; f[1] = fy; f[2] = fz;
    }
  }
}

// The ENERGIES of Forces.compute (forces.py:82-344) for a batch of replicas, every term in FP64 from the fp32 positions
// (energy_pass_impl, b200md_forces.cuh): one CTA per replica (grid-stride), positions converted to double in shared
// memory once, item lists read through L1.  A kernel of its own rather than a tail of forces_kernel: it wants what that
// kernel cannot give -- 4 KB of shared memory per CTA, so that the ~40 KB of item lists stay L1-resident (beside
// forces_kernel's 4 x 52 KB of shared memory the unified L1 keeps 20 KB and every list read went to L2: 2.9 stall cycles
// per issue on the long scoreboard, profiles/r02_forces_kernel.md), and twice the resident warps to cover the FP64
// dependency chains.  FP64-pipe bound (B200: 64 DFMA / clock / SM).
template <int MODE>
__global__ void __launch_bounds__(kThreads, 6)
energy_kernel(const DevTopo T, const ForceParams P, const long long R, const float* __restrict__ pos,
              const float* __restrict__ box, const long long box_mod, double* __restrict__ energies, const int estride,
              const int nslots) {
  __shared__ __align__(16) double spd[4 * kMaxAtoms];
  __shared__ double sered[kWarps * kNumReduced];
  __shared__ float sbox[8];
  Smem S{};
  S.pd = spd; S.ered = sered; S.box = sbox;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, N = T.natoms;
  const float q = tid < N ? __ldg(T.charge + tid) * P.el_on : 0.0f;
  for (long long r = blockIdx.x; r < R; r += gridDim.x) {
    float x = 0.f, y = 0.f, z = 0.f;
    if (tid < N) {
      const float* p = pos + (r * N + tid) * 3;
      x = __ldg(p); y = __ldg(p + 1); z = __ldg(p + 2);
    }
    const long long rb = r % box_mod;  // records of the fused integrator: record (slot, replica) uses the replica's box
    const float b0 = box ? __ldg(box + rb * 9) : 0.0f, b1 = box ? __ldg(box + rb * 9 + 4) : 0.0f, b2 = box ? __ldg(box + rb * 9 + 8) : 0.0f;
    const bool pbc = b0 != 0.0f || b1 != 0.0f || b2 != 0.0f;
    __syncthreads();  // the previous replica's reads of spd / sered are done
    stage_pd(S, tid, x, y, z, q);
    if (tid < 3) sbox[tid] = -(tid == 0 ? b0 : (tid == 1 ? b1 : b2));
    __syncthreads();
    Energies E;
    E.clear();
    energy_pass_impl<MODE>(S, T, P, pbc, tid, E);
    if (P.smd_on && tid == 0 && nslots == B200MD_NUM_ENERGIES) {  // potcalc.py:75-89, as smd_task evaluates it for the force
      const float center = P.smd_centers ? __ldg(P.smd_centers + r) : P.smd_c0;
      const double dx = spd[P.smd_a2] - spd[P.smd_a1], dy = spd[kMaxAtoms + P.smd_a2] - spd[kMaxAtoms + P.smd_a1],
                   dz = spd[2 * kMaxAtoms + P.smd_a2] - spd[2 * kMaxAtoms + P.smd_a1];
      const double dev = (double)center - sqrt(dx * dx + dy * dy + dz * dz);
      E.e[B200MD_E_EXTERNAL] += 0.5 * (double)P.smd_fk * dev * dev;
    }
    reduce_energies<B200MD_NUM_ENERGIES>(S, E, warp, lane);
    if (tid == 0) {
#pragma unroll
      for (int k = 0; k < B200MD_NUM_ENERGIES; ++k)
        if (k < nslots) energies[r * estride + k] = E.e[k];
    }
  }
}

__device__ __forceinline__ unsigned int ld_acquire_u32(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_u32(unsigned int* p, unsigned int v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Integrator.step(niter) of upstream TorchMD fused with Forces.compute: positions, velocities and forces of a
// replica stay in registers / shared memory for a whole SLICE of steps.
//
// Work distribution (time slicing): the launch is `nslices` x R units, a unit = `slice_len` consecutive steps of one
// replica.  The persistent CTAs (one wave: grid <= resident CTAs) draw units slice-major from a global ticket
// counter; between its slices a replica's state parks in the caller's own pos / vel / forces arrays (3.7 KB, L2).  A
// CTA that draws slice s of replica r waits for `slice_done[r] >= s` -- the unit with that ticket was drawn earlier,
// so whoever holds it is running: no deadlock, and a wait only happens when R is within a wave of the resident
// count.  The arithmetic of a step does not depend on the slicing, so results are bit-identical for every slice
// length; what it buys is a tail of one slice instead of one whole launch per replica (592 resident CTAs: 1 024
// replicas x 1 000 steps took two full waves).
// 5 CTAs per SM (<= 102 registers): the task loops fit without spills (ptxas moves ~200 bytes of per-launch state to local
// memory, outside them), and the fifth CTA is worth +2 % (9.08e7 vs 8.91e7 replica-steps/s); at 6 (80 registers) the pair
// loops spill and the rate drops to 7.1e7.  The cutoff mode and the rigid-bond instantiation need their 128 registers
// (at 5 CTAs: 6.77e7 vs 7.23e7 and 6.38e7 vs 6.64e7) and stay at 4.
template <int MODE, bool RIGID>
__global__ void __launch_bounds__(kThreads, (MODE == 0 && !RIGID) ? 5 : 4)
integrate_kernel(const DevTopo T, const ForceParams P, const StepParams I, const long long R, float* __restrict__ pos,
                 float* __restrict__ vel, float* __restrict__ forces, const float* __restrict__ box,
                 double* __restrict__ observables, float* __restrict__ snapshots, uint32_t* __restrict__ status) {
  extern __shared__ __align__(16) char smem_raw[];
  Smem S;
  smem_layout(&S, smem_raw, T.ntypes, T.nslots, true);
  // pinned in registers: under register pressure the compiler otherwise re-reads SR_TID.X (a ~25-cycle S2R)
  // at the top of every task
  int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  asm volatile("" : "+r"(tid), "+r"(warp), "+r"(lane));
  const int N = T.natoms;
  stage_tables<true>(S, T, P, tid);
  const bool live = tid < N;
  const float q = live ? __ldg(T.charge + tid) * P.el_on : 0.0f;
  const float mass = live ? __ldg(T.mass + tid) : 1.0f;
  const float invm = 1.0f / mass;
  const float dt = I.dt, hdt_m = 0.5f * dt * invm;
  const bool langevin = I.temperature > 0.0f;
  const float vcoef = langevin ? sqrtf(2.0f * I.gamma * invm * kBoltzman * I.temperature * dt) : 0.0f;
  const float vdamp = I.gamma * dt;
  const float inv_dt = 1.0f / dt;
  const int cmax = RIGID ? (int)((T.cnmax >> (8 * warp)) & 0xffu) : 0;
  const unsigned long long nunits = (unsigned long long)R * (unsigned long long)I.nslices;
  for (;;) {
    __syncthreads();  // the previous unit is complete in every thread (S.tick may be rewritten)
    if (tid == 0) *S.tick = atomicAdd(I.ticket, 1ull);
    __syncthreads();
    const unsigned long long unit = *S.tick;
    if (unit >= nunits) break;
    const int slice = (int)(unit / (unsigned long long)R);
    const long long r = (long long)(unit - (unsigned long long)slice * (unsigned long long)R);
    const int s_begin = slice * I.slice_len, s_end = min(I.nsteps, s_begin + I.slice_len);
    if (slice > 0) {  // the replica's previous slice must have parked its state
      if (tid == 0) {
        while (ld_acquire_u32(I.slice_done + r) < (unsigned int)slice) __nanosleep(100);
      }
      __syncthreads();
      __threadfence();
    }
    float x = 0.f, y = 0.f, z = 0.f, vx = 0.f, vy = 0.f, vz = 0.f, fx = 0.f, fy = 0.f, fz = 0.f;
    uint32_t flag = 0;
    const long long base = (r * N + tid) * 3;
    if (live) {  // L2 loads: another SM may have written this state a moment ago
      x = __ldcg(pos + base); y = __ldcg(pos + base + 1); z = __ldcg(pos + base + 2);
      vx = __ldcg(vel + base); vy = __ldcg(vel + base + 1); vz = __ldcg(vel + base + 2);
      fx = __ldcg(forces + base); fy = __ldcg(forces + base + 1); fz = __ldcg(forces + base + 2);
      if (x != x || y != y || z != z) flag |= B200MD_FLAG_NAN_POS;
    }
    const float b0 = box ? __ldg(box + r * 9) : 0.0f, b1 = box ? __ldg(box + r * 9 + 4) : 0.0f, b2 = box ? __ldg(box + r * 9 + 8) : 0.0f;
    const bool pbc = b0 != 0.0f || b1 != 0.0f || b2 != 0.0f;  // forces.py:359: all zero = vacuum, no wrap
    if (tid < 3) {
      const float b = tid == 0 ? b0 : (tid == 1 ? b1 : b2);
      S.box[tid] = -b;
      S.box[4 + tid] = b != 0.0f ? 1.0f / b : 0.0f;
    }
    if (status != nullptr && tid == 0 && slice == 0) status[r] = 0;
    // work of the moving restraint since the first step of the launch; same thread as smd_task: no barrier needed
    if (tid == kThreads - 1) S.smd_energy[1] = (slice > 0 && P.smd_on) ? __ldcg(I.smd_work + r) : 0.0;
    if (RIGID && live) S.pos[tid] = make_float4(x, y, z, q);  // SHAKE reads x(t) from here (visible after its barrier)
    for (int s = s_begin; s < s_end; ++s) {
      const unsigned long long gstep = I.step0 + (unsigned long long)s;
      // _first_VV: pos += vel dt + 0.5 a dt^2 ; vel += 0.5 dt a
      vx = fmaf(hdt_m, fx, vx); vy = fmaf(hdt_m, fy, vy); vz = fmaf(hdt_m, fz, vz);
      x = fmaf(dt, vx, x); y = fmaf(dt, vy, y); z = fmaf(dt, vz, z);
      if (RIGID) {  // SHAKE: S.pos still holds x(t); publish the unconstrained x(t + dt), collect displacements
        S.sx[tid] = x; S.sx[kMaxAtoms + tid] = y; S.sx[2 * kMaxAtoms + tid] = z;
        S.vel[tid] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        __syncthreads();
        if (tid < T.ngroups) {
          const bool ok = cmax > 1 ? shake_group<3>(S, T, tid) : shake_group<1>(S, T, tid);
          if (!ok) flag |= B200MD_FLAG_CONSTRAINT;
        }
        __syncthreads();
        const float4 d = S.vel[tid];
        x += d.x; y += d.y; z += d.z;
        vx = fmaf(d.x, inv_dt, vx); vy = fmaf(d.y, inv_dt, vy); vz = fmaf(d.z, inv_dt, vz);
      }
      if (live) {
        S.pos[tid] = make_float4(x, y, z, q);
        S.sx[tid] = x; S.sx[kMaxAtoms + tid] = y; S.sx[2 * kMaxAtoms + tid] = z; S.sx[3 * kMaxAtoms + tid] = q;
      }
      S.vel[tid] = make_float4(vx, vy, vz, 0.0f);  // parked: frees 6 registers across the task phase
      __syncthreads();
      // a record every energy_every steps and after the last step of the call (what step() returns is the end state)
      const bool report = (I.energy_every > 0 && (s + 1) % I.energy_every == 0) || s == I.nsteps - 1;
      if (pbc) run_tasks<MODE, true, true, !RIGID>(S, T, P, warp, lane);
      else run_tasks<MODE, true, false, !RIGID>(S, T, P, warp, lane);
      if (P.smd_on && tid == kThreads - 1) smd_task(S, P, (float)((double)P.smd_c0 + (double)P.smd_rate * (double)(gstep + 1)));
      __syncthreads();
      {
        const float4 pp = S.pos[tid], vv = S.vel[tid];
        x = pp.x; y = pp.y; z = pp.z; vx = vv.x; vy = vv.y; vz = vv.z;
      }
      if (live) gather_forces<true>(S, P, tid, fx, fy, fz, flag);
      if (langevin) {  // vel += -gamma vel dt + vcoeff N(0,1)
        float n[3];
        normal3(I.seed, (unsigned long long)(I.replica0 + r), (uint32_t)tid, gstep, n);
        vx += fmaf(vcoef, n[0], -vdamp * vx); vy += fmaf(vcoef, n[1], -vdamp * vy); vz += fmaf(vcoef, n[2], -vdamp * vz);
      }
      // _second_VV
      vx = fmaf(hdt_m, fx, vx); vy = fmaf(hdt_m, fy, vy); vz = fmaf(hdt_m, fz, vz);
      if (RIGID) {  // RATTLE: take the components along the rigid bonds out of v(t + dt)
        S.vel[tid] = make_float4(vx, vy, vz, 0.0f);
        __syncthreads();
        if (tid < T.ngroups) {
          if (cmax > 1) rattle_group<3>(S, T, tid); else rattle_group<1>(S, T, tid);
        }
        __syncthreads();
        const float4 vv = S.vel[tid];
        vx = vv.x; vy = vv.y; vz = vv.z;
      }
      if (report) {
        const long long slot = I.energy_every > 0 ? s / I.energy_every : 0;
        if (I.report_pos != nullptr) {
          // Deferred energies: the record's positions go to HBM (coalesced, 1 248 B) and energy_kernel evaluates the seven
          // force-field terms of every record behind this launch, at its own occupancy and with L1-resident item lists.
          // Here only what needs the velocities or the restraint state: Ekin (fixed-order FP64 block sum), the
          // restraint's energy, distance and work.
          if (live) {
            float* o = I.report_pos + ((slot * R + r) * N + tid) * 3;
            o[0] = x; o[1] = y; o[2] = z;
          }
          double ek = live ? 0.5 * (double)mass * ((double)vx * vx + (double)vy * vy + (double)vz * vz) : 0.0;
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) ek += __shfl_xor_sync(0xffffffffu, ek, o);
          if (lane == 0) S.ered[warp] = ek;
          __syncthreads();
          if (tid == 0 && observables != nullptr) {
            double* o = observables + (slot * R + r) * B200MD_NUM_OBSERVABLES;
            o[B200MD_E_EXTERNAL] = P.smd_on ? S.smd_energy[0] : 0.0;
            o[B200MD_OBS_EKIN] = (S.ered[0] + S.ered[1]) + (S.ered[2] + S.ered[3]);
            o[B200MD_OBS_E2E_DIST] = P.smd_on ? (double)S.smd[4] : 0.0;
            o[B200MD_OBS_SMD_WORK] = P.smd_on ? S.smd_energy[1] : 0.0;
          }
        } else {  // FP64 energies of x(t + dt) in place: the parked velocities and the SoA copy are dead until the next step
          __syncthreads();  // every thread has read its S.vel entry
          stage_pd(S, tid, x, y, z, live ? q : 0.0f);
          __syncthreads();
          Energies E;
          E.clear();
          energy_pass<MODE>(S, T, P, pbc, tid, E);
          if (live) E.e[B200MD_OBS_EKIN] = 0.5 * (double)mass * ((double)vx * vx + (double)vy * vy + (double)vz * vz);
          if (P.smd_on && tid == 0) { E.e[B200MD_E_EXTERNAL] += S.smd_energy[0]; E.e[B200MD_OBS_E2E_DIST] = (double)S.smd[4]; }
          reduce_energies<kNumReduced>(S, E, warp, lane);
          if (tid == 0 && observables != nullptr) {
            double* o = observables + (slot * R + r) * B200MD_NUM_OBSERVABLES;
#pragma unroll
            for (int k = 0; k < kNumReduced; ++k) o[k] = E.e[k];
            o[B200MD_OBS_SMD_WORK] = P.smd_on ? S.smd_energy[1] : 0.0;
          }
        }
      }
      if (I.snapshot_every > 0 && snapshots != nullptr && (s + 1) % I.snapshot_every == 0 && live) {
        const long long slot = (s + 1) / I.snapshot_every - 1;
        float* o = snapshots + ((slot * R + r) * N + tid) * 3;
        o[0] = x; o[1] = y; o[2] = z;
      }
    }
    if (live) {
      pos[base] = x; pos[base + 1] = y; pos[base + 2] = z;
      vel[base] = vx; vel[base + 1] = vy; vel[base + 2] = vz;
      forces[base] = fx; forces[base + 1] = fy; forces[base + 2] = fz;
      if (x != x || y != y || z != z) flag |= B200MD_FLAG_NAN_POS;
    }
    if (flag && status != nullptr) atomicOr(status + r, flag);
    if (I.nslices > 1) {  // park: publish the state, then the slice count (release)
      if (P.smd_on && tid == kThreads - 1) I.smd_work[r] = S.smd_energy[1];
      __threadfence();
      __syncthreads();
      if (tid == 0) st_release_u32(I.slice_done + r, (unsigned int)(slice + 1));
    }
  }
}

// ---- Energy.calc_phi (cGAN/cGAN_D_pos-2.3.2/cgan.py:166-207, the angle of forces.py:541-552) -------------------
// One CTA per replica: positions staged in shared memory by coalesced loads, lane d evaluates dihedral d, output
// rows are written coalesced.  HBM-bound: 12 N bytes in, 4 ndihedrals out per replica.
struct PhiGeom {
  float ax, ay, az, bx, by, bz;      // crossA = r12 x r23, crossB = r23 x r34
  float r12x, r12y, r12z, r23x, r23y, r23z, r34x, r34y, r34z;
};

__device__ __forceinline__ float wrap_scalar(float d, float b) { return b != 0.0f ? d - b * rintf(d / b) : d; }

__device__ __forceinline__ PhiGeom phi_geometry(const float4* pos, uchar4 q, const float* bx) {
  const float4 p0 = pos[q.x], p1 = pos[q.y], p2 = pos[q.z], p3 = pos[q.w];
  PhiGeom g;
  g.r12x = wrap_scalar(p0.x - p1.x, bx[0]); g.r12y = wrap_scalar(p0.y - p1.y, bx[1]); g.r12z = wrap_scalar(p0.z - p1.z, bx[2]);
  g.r23x = wrap_scalar(p1.x - p2.x, bx[0]); g.r23y = wrap_scalar(p1.y - p2.y, bx[1]); g.r23z = wrap_scalar(p1.z - p2.z, bx[2]);
  g.r34x = wrap_scalar(p2.x - p3.x, bx[0]); g.r34y = wrap_scalar(p2.y - p3.y, bx[1]); g.r34z = wrap_scalar(p2.z - p3.z, bx[2]);
  g.ax = g.r12y * g.r23z - g.r12z * g.r23y; g.ay = g.r12z * g.r23x - g.r12x * g.r23z; g.az = g.r12x * g.r23y - g.r12y * g.r23x;
  g.bx = g.r23y * g.r34z - g.r23z * g.r34y; g.by = g.r23z * g.r34x - g.r23x * g.r34z; g.bz = g.r23x * g.r34y - g.r23y * g.r34x;
  return g;
}

// crossC = r23 x crossA;  cos = A.B / (|A||B|),  sin = C.B / (|C||B|),  phi = -atan2(sin, cos)
__device__ __forceinline__ float phi_angle(const PhiGeom& g) {
  const float cx = g.r23y * g.az - g.r23z * g.ay, cy = g.r23z * g.ax - g.r23x * g.az, cz = g.r23x * g.ay - g.r23y * g.ax;
  const float na = sqrtf(g.ax * g.ax + g.ay * g.ay + g.az * g.az), nb = sqrtf(g.bx * g.bx + g.by * g.by + g.bz * g.bz);
  const float nc = sqrtf(cx * cx + cy * cy + cz * cz);
  const float cosp = (g.ax * g.bx + g.ay * g.by + g.az * g.bz) / nb / na;
  const float sinp = (cx * g.bx + cy * g.by + cz * g.bz) / nb / nc;
  return -atan2f(sinp, cosp);
}
// d phi / d x of the atom at position `role` of the quadruple (see dihedral_angles_backward_kernel)
__device__ __forceinline__ void phi_gradient(const PhiGeom& g, const int role, float& dx, float& dy, float& dz) {
  const float n23sq = g.r23x * g.r23x + g.r23y * g.r23y + g.r23z * g.r23z, n23 = sqrtf(n23sq);
  const float f0 = -n23 / (g.ax * g.ax + g.ay * g.ay + g.az * g.az), f3 = n23 / (g.bx * g.bx + g.by * g.by + g.bz * g.bz);
  const float g0x = f0 * g.ax, g0y = f0 * g.ay, g0z = f0 * g.az, g3x = f3 * g.bx, g3y = f3 * g.by, g3z = f3 * g.bz;
  const float ff1 = (g.r12x * g.r23x + g.r12y * g.r23y + g.r12z * g.r23z) / n23sq;
  const float ff2 = (g.r34x * g.r23x + g.r34y * g.r23y + g.r34z * g.r23z) / n23sq;
  const float sx = ff1 * g0x - ff2 * g3x, sy = ff1 * g0y - ff2 * g3y, sz = ff1 * g0z - ff2 * g3z;
  if (role == 0) { dx = g0x; dy = g0y; dz = g0z; }
  else if (role == 1) { dx = -g0x - sx; dy = -g0y - sy; dz = -g0z - sz; }
  else if (role == 2) { dx = -g3x + sx; dy = -g3y + sy; dz = -g3z + sz; }
  else { dx = g3x; dy = g3y; dz = g3z; }
}

__device__ __forceinline__ void stage_replica(float4* spos, float* sbox, const float* pos, const float* box, long long r,
                                              int N, int tid) {
  if (tid < N) {
    const long long base = (r * N + tid) * 3;
    spos[tid] = make_float4(pos[base], pos[base + 1], pos[base + 2], 0.0f);
  }
  if (tid < 3) sbox[tid] = box ? __ldg(box + r * 9 + 4 * tid) : 0.0f;
  // forces.py:359 / cgan.py:185: a box of zeros is vacuum; the reference wraps with every component otherwise
}

__global__ void __launch_bounds__(kThreads) dihedral_angles_kernel(const DevTopo T, const long long R,
                                                                   const float* __restrict__ pos,
                                                                   const float* __restrict__ box, float* __restrict__ phi) {
  __shared__ float4 spos[kMaxAtoms];
  __shared__ float sbox[4];
  const int tid = threadIdx.x;
  for (long long r = blockIdx.x; r < R; r += gridDim.x) {
    __syncthreads();
    stage_replica(spos, sbox, pos, box, r, T.natoms, tid);
    __syncthreads();
    for (int d = tid; d < T.nphi; d += kThreads) {
      phi[r * T.nphi + d] = phi_angle(phi_geometry(spos, T.phi_idx[d], sbox));
    }
  }
}

// grad_pos[a] = sum over the dihedrals d that contain a of grad_phi[d] * d phi_d / d x_a.  Atom-major gather (each
// lane recomputes the geometry of the dihedrals its atom sits in): no atomics, fixed order, bit-reproducible.
// d phi / d x follows from the reference's torsion forces with coeff = 1 (forces.py:578-600):
//   g0 = -|r23| / |A|^2 A,  g3 = |r23| / |B|^2 B,  s = (r12.r23 / |r23|^2) g0 - (r34.r23 / |r23|^2) g3,
//   d phi / d x = (g0, -g0 - s, -g3 + s, g3).
__global__ void __launch_bounds__(kThreads) dihedral_angles_backward_kernel(const DevTopo T, const long long R,
                                                                            const float* __restrict__ pos,
                                                                            const float* __restrict__ box,
                                                                            const float* __restrict__ grad_phi,
                                                                            float* __restrict__ grad_pos) {
  __shared__ float4 spos[kMaxAtoms];
  __shared__ float sbox[4];
  __shared__ float sg[512];  // grad_phi row of the replica (host guarantees nphi <= 512 for this path)
  const int tid = threadIdx.x;
  for (long long r = blockIdx.x; r < R; r += gridDim.x) {
    __syncthreads();
    stage_replica(spos, sbox, pos, box, r, T.natoms, tid);
    for (int d = tid; d < T.nphi; d += kThreads) sg[d] = grad_phi[r * T.nphi + d];
    __syncthreads();
    if (tid < T.natoms) {
      float gx = 0.0f, gy = 0.0f, gz = 0.0f;
      const int i1 = T.phi_start[tid + 1];
      for (int i = T.phi_start[tid]; i < i1; ++i) {
        const int item = T.phi_items[i], d = item & 0x3fff, role = item >> 14;
        float dx, dy, dz;
        phi_gradient(phi_geometry(spos, T.phi_idx[d], sbox), role, dx, dy, dz);
        const float w = sg[d];
        gx = fmaf(w, dx, gx); gy = fmaf(w, dy, gy); gz = fmaf(w, dz, gz);
      }
      const long long base = (r * T.natoms + tid) * 3;
      grad_pos[base] = gx; grad_pos[base + 1] = gy; grad_pos[base + 2] = gz;
    }
  }
}

// The same two kernels for solutes above 128 atoms: the replica does not fit shared memory, every thread reads the four
// atoms of its dihedral from HBM (one thread per (replica, dihedral) forward, per (replica, atom) backward; same arithmetic).
__device__ __forceinline__ PhiGeom large_phi_geometry(const float* __restrict__ p, const int4 q, const float* bx) {
  float4 lp[4];
  const int a[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
  for (int k = 0; k < 4; ++k) lp[k] = make_float4(p[3 * a[k]], p[3 * a[k] + 1], p[3 * a[k] + 2], 0.0f);
  return phi_geometry(lp, make_uchar4(0, 1, 2, 3), bx);
}
__global__ void large_dihedral_angles_kernel(const LargeTopo T, const long long R, const float* __restrict__ pos,
                                             const float* __restrict__ box, float* __restrict__ phi) {
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < R * T.ndih; g += (long long)gridDim.x * blockDim.x) {
    const long long r = g / T.ndih;
    const int d = (int)(g - r * T.ndih);
    const float bx[3] = {box ? __ldg(box + r * 9) : 0.0f, box ? __ldg(box + r * 9 + 4) : 0.0f, box ? __ldg(box + r * 9 + 8) : 0.0f};
    phi[g] = phi_angle(large_phi_geometry(pos + r * T.natoms * 3, __ldg(T.dih_idx + d), bx));
  }
}
__global__ void large_dihedral_angles_backward_kernel(const LargeTopo T, const long long R, const float* __restrict__ pos,
                                                      const float* __restrict__ box, const float* __restrict__ grad_phi,
                                                      float* __restrict__ grad_pos) {
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < R * T.natoms; g += (long long)gridDim.x * blockDim.x) {
    const long long r = g / T.natoms;
    const int a = (int)(g - r * T.natoms);
    const float bx[3] = {box ? __ldg(box + r * 9) : 0.0f, box ? __ldg(box + r * 9 + 4) : 0.0f, box ? __ldg(box + r * 9 + 8) : 0.0f};
    float gx = 0.0f, gy = 0.0f, gz = 0.0f;
    const int i1 = __ldg(T.phi_start + a + 1);
    for (int i = __ldg(T.phi_start + a); i < i1; ++i) {
      const int item = __ldg(T.phi_items + i), d = item >> 2, role = item & 3;
      float dx, dy, dz;
      phi_gradient(large_phi_geometry(pos + r * T.natoms * 3, __ldg(T.dih_idx + d), bx), role, dx, dy, dz);
      const float w = grad_phi[r * T.ndih + d];
      gx = fmaf(w, dx, gx); gy = fmaf(w, dy, gy); gz = fmaf(w, dz, gz);
    }
    grad_pos[3 * g] = gx; grad_pos[3 * g + 1] = gy; grad_pos[3 * g + 2] = gz;
  }
}

// Projection of a start structure onto the rigid-bond constraints (b200md_constrain): positions by SHAKE along the
// present bond vectors (x(t) = the unconstrained x itself), velocities by RATTLE.  One CTA per replica; only the
// three arrays the group solvers read are set up in the shared-memory view.
__global__ void __launch_bounds__(kThreads) constrain_kernel(const DevTopo T, const long long R, float* __restrict__ pos,
                                                             float* __restrict__ vel, uint32_t* __restrict__ status) {
  __shared__ float4 spos[kMaxAtoms];
  __shared__ float4 svel[kMaxAtoms];
  __shared__ float ssx[4 * kMaxAtoms];
  Smem S{};
  S.pos = spos; S.vel = svel; S.sx = ssx;
  const int tid = threadIdx.x, warp = tid >> 5, N = T.natoms;
  const int cmax = (int)((T.cnmax >> (8 * warp)) & 0xffu);
  const bool live = tid < N;
  for (long long r = blockIdx.x; r < R; r += gridDim.x) {
    __syncthreads();
    const long long base = (r * N + tid) * 3;
    float x = 0.f, y = 0.f, z = 0.f, vx = 0.f, vy = 0.f, vz = 0.f;
    if (live) {
      x = pos[base]; y = pos[base + 1]; z = pos[base + 2];
      if (vel) { vx = vel[base]; vy = vel[base + 1]; vz = vel[base + 2]; }
    }
    uint32_t flag = 0;
    bool ok = true;
    for (int sweep = 0; sweep < 4; ++sweep) {  // far-off starts: the linearisation point follows the projection
      spos[tid] = make_float4(x, y, z, 0.0f);
      ssx[tid] = x; ssx[kMaxAtoms + tid] = y; ssx[2 * kMaxAtoms + tid] = z;
      svel[tid] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
      __syncthreads();
      ok = true;
      if (tid < T.ngroups) ok = cmax > 1 ? shake_group<3>(S, T, tid) : shake_group<1>(S, T, tid);
      __syncthreads();
      const float4 d = svel[tid];
      x += d.x; y += d.y; z += d.z;
      __syncthreads();
    }
    if (!ok) flag |= B200MD_FLAG_CONSTRAINT;
    spos[tid] = make_float4(x, y, z, 0.0f);
    svel[tid] = make_float4(vx, vy, vz, 0.0f);
    __syncthreads();
    if (tid < T.ngroups) { if (cmax > 1) rattle_group<3>(S, T, tid); else rattle_group<1>(S, T, tid); }
    __syncthreads();
    if (live) {
      const float4 v = svel[tid];
      pos[base] = x; pos[base + 1] = y; pos[base + 2] = z;
      if (vel) { vel[base] = v.x; vel[base + 1] = v.y; vel[base + 2] = v.z; }
    }
    if (status != nullptr) {
      if (tid == 0) status[r] = 0;
      __syncthreads();
      if (flag) atomicOr(status + r, flag);
    }
  }
}

// FP32 FMA throughput micro-benchmark (the roofline denominator): 8 independent chains per thread.
__global__ void __launch_bounds__(256) fma_peak_kernel(float* out, int iters, float a, float b) {
  float v[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) v[k] = (float)(threadIdx.x + k) * 1e-3f;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = fmaf(v[k], a, b);
    }
  }
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += v[k];
  if (s == 12345.678f) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

}  // namespace b200md

// =================================================================================================
// host side: topology packing (Forces.__init__, forces.py:28-73) and the C-ABI
// =================================================================================================
static thread_local std::string g_last_error;
static std::atomic<long long> g_launches{0};

static int fail(int code, const std::string& msg) {
  g_last_error = msg;
  return code;
}
#define CUDA_TRY(expr)                                                                         \
  do {                                                                                         \
    cudaError_t _e = (expr);                                                                   \
    if (_e != cudaSuccess)                                                                     \
      return fail(B200MD_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e));        \
  } while (0)

// Entry points run on the topology's device and put the caller's current device back afterwards (a host that drives
// several GPUs from one thread must not find its current device changed by a library call).
struct DeviceGuard {
  int prev = -1;
  cudaError_t err;
  explicit DeviceGuard(int device) {
    err = cudaGetDevice(&prev);
    if (err == cudaSuccess && prev != device) err = cudaSetDevice(device); else if (err == cudaSuccess) prev = -1;
  }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};
#define B200MD_ON_DEVICE(dev)                                                                  \
  DeviceGuard _guard(dev);                                                                     \
  if (_guard.err != cudaSuccess) return fail(B200MD_ERR_CUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(_guard.err))

struct b200md_topology {
  int device = 0;
  int natoms = 0;
  int num_sms = 0;
  DevTopo dev{};
  void* blob = nullptr;
  std::vector<uint32_t> mask_rows;  // [natoms][4], bit j set <=> pair (i, j) not evaluated
  int occ_forces[2] = {0, 0}, occ_integrate[2] = {0, 0}, occ_energy[2] = {1, 1};
  int occ_steps[2] = {0, 0};  // resident CTAs per SM of large_steps_kernel<MODE> (0: not available)
  // solutes above 128 atoms (b200md_large.cuh): the per-replica state no longer fits one CTA
  bool large = false;
  LargeTopo ldev{};
  int mask_words = 4;  // uint32 words per row of mask_rows
  // grow-only device workspace of the large-solute calls (partial force rows, energy partials, tile boxes ...), reused from
  // call to call: stream-ordered on `ws_stream`; a call on another stream first waits for that one
  mutable void* ws = nullptr;
  mutable size_t ws_bytes = 0;
  mutable cudaStream_t ws_stream = nullptr;
  // second stream of the large-solute calls: the latency-bound bonded kernel (and the FP64 energy kernels) run beside the throughput-bound pair
  // kernel (fork / join by events, so the caller's stream still orders the call as a whole)
  mutable cudaStream_t aux = nullptr;
  mutable cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_bonded = nullptr;
};

namespace {

struct BlobBuilder {
  std::vector<char> bytes;
  template <typename Tv>
  size_t add(const std::vector<Tv>& v, size_t min_elems = 1) {
    size_t off = (bytes.size() + 15) & ~size_t(15);
    size_t n = std::max(v.size(), min_elems);
    bytes.resize(off + n * sizeof(Tv), 0);
    if (!v.empty()) memcpy(bytes.data() + off, v.data(), v.size() * sizeof(Tv));
    return off;
  }
};

uint32_t pack4(const int32_t* a, int n) {
  uint32_t q = 0;
  for (int k = 0; k < n; ++k) q |= (uint32_t)(a[k] & 0xff) << (8 * k);
  return q;
}

// flatten par.dihedral_params slices into per-dihedral [first, count] ranges (slice order kept).
// term = (k0, per | marker, c0, s0): integer periodicity 1..8 -> (cos phi0, sin phi0) so the kernel never
// forms phi; harmonic slices (forces.py:565 failed) and non-integer periodicities keep phi0.
void build_terms(int nitems, int nterms, const int32_t* owner, const float* prm, const int32_t* harmonic,
                 std::vector<uint2>& ranges, std::vector<float4>& out) {
  ranges.assign(std::max(nitems, 1), make_uint2(0, 0));
  out.clear();
  for (int d = 0; d < nitems; ++d) {
    uint32_t first = (uint32_t)out.size();
    for (int t = 0; t < nterms; ++t) {
      if (owner[t] != d) continue;
      const float k0 = prm[3 * t], phi0 = prm[3 * t + 1], per = prm[3 * t + 2];
      if (harmonic[t]) {
        out.push_back(make_float4(k0, kTermHarmonic, phi0, 0.0f));
      } else if (per >= 1.0f && per <= 8.0f && per == floorf(per)) {
        out.push_back(make_float4(k0, per, (float)cos((double)phi0), (float)sin((double)phi0)));
      } else {
        out.push_back(make_float4(k0, kTermGeneral, phi0, per));
      }
    }
    ranges[d] = make_uint2(first, (uint32_t)out.size() - first);
  }
}

// Orders the items of one bonded kind into consecutive lane sets (`sets`: sizes, <= 32 each) whose members touch
// distinct shared-memory banks (atom mod 32) at every atom position: the packed bonded tasks gather positions with
// LDS.32 and accumulate forces with shared-memory integer atomics, and both serialise on bank collisions (2-6
// wavefronts per instruction in topology order, profiles/r01_v14).  Greedy fill, then pairwise swaps between sets
// while they lower the collision count (deterministic: fixed LCG seed).
int bank_cost(const std::vector<int>& seq, size_t begin, int size, const int32_t* idx, int arity) {
  int cost = 0;
  for (int k = 0; k < arity; ++k) {
    int cnt[32] = {0};
    for (int l = 0; l < size; ++l) cnt[idx[arity * seq[begin + l] + k] & 31]++;
    for (int b = 0; b < 32; ++b) cost += cnt[b] > 1 ? (cnt[b] - 1) * (cnt[b] - 1) : 0;
  }
  return cost;
}
std::vector<int> spread_banks(const std::vector<int>& order, const int32_t* idx, int arity, const std::vector<int>& sets) {
  std::vector<int> out;
  std::vector<char> used(order.size(), 0);
  for (int size : sets) {
    uint32_t busy[4] = {0, 0, 0, 0};
    for (int fill = 0; fill < size; ++fill) {
      long best = -1;
      int bestc = 1 << 30;
      for (size_t m = 0; m < order.size() && bestc > 0; ++m) {
        if (used[m]) continue;
        int c = 0;
        for (int k = 0; k < arity; ++k) c += (busy[k] >> (idx[arity * order[m] + k] & 31)) & 1u;
        if (c < bestc) { bestc = c; best = (long)m; }
      }
      if (best < 0) break;
      used[best] = 1;
      out.push_back(order[best]);
      for (int k = 0; k < arity; ++k) busy[k] |= 1u << (idx[arity * order[best] + k] & 31);
    }
  }
  std::vector<size_t> begin;
  size_t pos = 0;
  for (int size : sets) { begin.push_back(pos); pos += size; }
  if (out.size() != pos || sets.size() < 2) return out;
  auto set_of = [&](size_t m) { size_t s2 = 0; while (s2 + 1 < begin.size() && begin[s2 + 1] <= m) ++s2; return s2; };
  uint32_t rng = 12345u;
  auto next = [&]() { rng = rng * 1664525u + 1013904223u; return rng >> 8; };
  for (int it = 0; it < 40000; ++it) {
    const size_t m1 = next() % out.size(), m2 = next() % out.size();
    const size_t s1 = set_of(m1), s2 = set_of(m2);
    if (s1 == s2) continue;
    const int before = bank_cost(out, begin[s1], sets[s1], idx, arity) + bank_cost(out, begin[s2], sets[s2], idx, arity);
    if (before == 0) continue;
    std::swap(out[m1], out[m2]);
    const int after = bank_cost(out, begin[s1], sets[s1], idx, arity) + bank_cost(out, begin[s2], sets[s2], idx, arity);
    if (after >= before) std::swap(out[m1], out[m2]);
  }
  if (getenv("B200MD_DEBUG")) {
    int total = 0;
    for (size_t s2 = 0; s2 < sets.size(); ++s2) total += bank_cost(out, begin[s2], sets[s2], idx, arity);
    fprintf(stderr, "b200md: spread_banks arity %d items %zu sets %zu residual collision cost %d\n", arity, out.size(), sets.size(), total);
  }
  return out;
}

template <typename K>
int query_occupancy(K kernel, int smem) {
  if (smem > 48 * 1024) cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  int occ = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, kThreads, smem);
  return occ;
}

}  // namespace

static b200md_topology* create_large(const b200md_topology_desc* d, int device);

extern "C" b200md_topology* b200md_topology_create(const b200md_topology_desc* d, int device) {
  if (!d || d->natoms <= 0 || d->ntypes <= 0) { fail(B200MD_ERR_INVALID_ARGUMENT, "empty topology"); return nullptr; }
  if (d->natoms > B200MD_MAX_ATOMS_LARGE) {
    fail(B200MD_ERR_TOO_MANY_ATOMS, "natoms > 32768: the all-pairs kernels keep a dense exclusion bit matrix");
    return nullptr;
  }
  const int N = d->natoms, NT = d->ntypes;
  auto bad_index = [&](const int32_t* idx, int n) {
    for (int k = 0; k < n; ++k) if (idx[k] < 0 || idx[k] >= N) return true;
    return false;
  };
  if (bad_index(d->bond_idx, 2 * d->nbonds) || bad_index(d->angle_idx, 3 * d->nangles) ||
      bad_index(d->dihedral_idx, 4 * d->ndihedrals) || bad_index(d->improper_idx, 4 * d->nimpropers) ||
      bad_index(d->idx14, 2 * d->n14) || bad_index(d->pair_idx, 2 * d->npairs)) {
    fail(B200MD_ERR_INVALID_ARGUMENT, "atom index out of range in topology");
    return nullptr;
  }
  for (int a = 0; a < N; ++a)
    if (d->atom_type[a] < 0 || d->atom_type[a] >= NT) { fail(B200MD_ERR_INVALID_ARGUMENT, "atom type out of range"); return nullptr; }

  if (N > kMaxAtoms) return create_large(d, device);

  auto* topo = new b200md_topology();
  topo->device = device;
  topo->natoms = N;
  DevTopo& T = topo->dev;
  T.natoms = N; T.ntypes = NT; T.nchunks = (N + kChunk - 1) / kChunk;

  // ---- exclusion mask from the pair list (= Forces.ava_idx) ----------------------------------
  std::vector<uint8_t> allowed((size_t)kMaxAtoms * kMaxAtoms, 0);
  for (int k = 0; k < d->npairs; ++k) {
    int i = d->pair_idx[2 * k], j = d->pair_idx[2 * k + 1];
    if (i > j) std::swap(i, j);
    if (i != j) allowed[(size_t)i * kMaxAtoms + j] = 1;
  }
  topo->mask_rows.assign((size_t)N * 4, 0);
  for (int i = 0; i < N; ++i)
    for (int j = 0; j < kMaxAtoms; ++j) {
      const int lo = std::min(i, j), hi = std::max(i, j);
      if (!(i != j && hi < N && allowed[(size_t)lo * kMaxAtoms + hi])) topo->mask_rows[i * 4 + j / 32] |= 1u << (j % 32);
    }
  std::vector<uint8_t> mask8((size_t)kMaxChunks * kMaxAtoms, 0xff);
  for (int c = 0; c < kMaxChunks; ++c)
    for (int i = 0; i < kMaxAtoms; ++i) {
      uint8_t m = 0;
      for (int s = 0; s < kChunk; ++s) {
        const int j = c * kChunk + s;
        const bool eval = i < j && j < N && allowed[(size_t)i * kMaxAtoms + j];
        if (!eval) m |= (uint8_t)(1u << s);
      }
      mask8[(size_t)c * kMaxAtoms + i] = m;
    }

  // diagonal-block masks (tile_diag_task): bit j of atom i = pair (i, 32*block(i) + j) is not evaluated
  std::vector<uint32_t> diagmask(kMaxAtoms, 0xffffffffu);
  for (int i = 0; i < N; ++i) diagmask[i] = topo->mask_rows[(size_t)i * 4 + i / 32];

  // ---- per-atom tables ---------------------------------------------------------------------------
  std::vector<float2> ab((size_t)NT * kAbStride, make_float2(0.f, 0.f));
  for (int t = 0; t < NT; ++t)
    for (int j = 0; j < N; ++j)
      ab[(size_t)t * kAbStride + j] = make_float2(12.0f * d->lj_a[t * NT + d->atom_type[j]], 6.0f * d->lj_b[t * NT + d->atom_type[j]]);
  std::vector<float> charge(kMaxAtoms, 0.f), mass(kMaxAtoms, 1.f);
  std::vector<int> rowoff(kMaxAtoms, 0);
  for (int a = 0; a < N; ++a) {
    charge[a] = d->charge ? d->charge[a] : 0.f;
    mass[a] = d->mass[a];
    rowoff[a] = d->atom_type[a] * kTabStride * (int)sizeof(float);
  }

  // ---- bonded items ------------------------------------------------------------------------------
  // Bonds, angles, proper torsions with integer periodicity and 1-4 pairs are evaluated TWO ITEMS PER LANE in
  // packed FP32 (b200md_forces.cuh): lane l of chunk c holds items A = seq[64c + l] and B = seq[64c + 32 + l].
  // The 32 A items (and the 32 B items) of a chunk are chosen so that they touch distinct shared-memory banks
  // (spread_banks).  Torsions the packed path cannot express (CHARMM harmonic slices, non-integer or > 6
  // periodicities) and all impropers go to the "general" list, evaluated one item per lane by torsion_task.
  const int ND = d->ndihedrals;
  std::vector<std::vector<int>> dterms(ND);
  for (int t = 0; t < d->ndihedral_terms; ++t) {
    const int o = d->dihedral_term_owner[t];
    if (o < 0 || o >= ND) { delete topo; fail(B200MD_ERR_INVALID_ARGUMENT, "dihedral term owner out of range"); return nullptr; }
    dterms[o].push_back(t);
  }
  for (int t = 0; t < d->nimproper_terms; ++t)
    if (d->improper_term_owner[t] < 0 || d->improper_term_owner[t] >= d->nimpropers) {
      delete topo; fail(B200MD_ERR_INVALID_ARGUMENT, "improper term owner out of range"); return nullptr;
    }
  std::vector<int> dclass(ND, 0), dmaxper(ND, 1);  // 0: packed series, 1: general, 2: every k is 0 (only its 1-4 pair survives)
  for (int q = 0; q < ND; ++q) {
    bool fast = true, zero = true;
    for (int t : dterms[q]) {
      const float k0 = d->dihedral_term_prm[3 * t], per = d->dihedral_term_prm[3 * t + 2];
      if (k0 != 0.0f) zero = false;
      if (d->dihedral_term_harmonic[t] || !(per >= 1.0f && per <= (float)kMaxPer && per == floorf(per))) fast = false;
      else dmaxper[q] = std::max(dmaxper[q], (int)per);
    }
    dclass[q] = zero ? 2 : (fast ? 0 : 1);
  }
  // Lane filling: zero-k dihedrals leave the packed list and hand their 1-4 pair to the (cheaper) 1-4 list -- but a ragged
  // tail of that list costs a whole task for a few lanes (ALA10: 72 = 64 + 8).  When the tail fits into the free lanes of the
  // last dihedral chunk, that many zero-k dihedrals stay packed (zero coefficients, their 1-4 pair fused) instead.
  {
    int nfast = 0, nzero14 = 0;
    std::vector<char> has14(ND, 0);
    for (int k = 0; k < d->n14; ++k)
      for (int q = 0; q < ND; ++q) {
        const int32_t* a4 = d->dihedral_idx + 4 * q;
        const int i = d->idx14[2 * k], j = d->idx14[2 * k + 1];
        if (!has14[q] && ((a4[0] == i && a4[3] == j) || (a4[0] == j && a4[3] == i))) { has14[q] = 1; break; }
      }
    for (int q = 0; q < ND; ++q) { nfast += dclass[q] == 0; nzero14 += dclass[q] == 2 && has14[q]; }
    int tail = nzero14 % 64;
    const int room = (64 - nfast % 64) % 64;
    if (tail > 0 && tail <= room && !getenv("B200MD_NO_TAIL_FILL"))
      for (int q = ND - 1; q >= 0 && tail > 0; --q)
        if (dclass[q] == 2 && has14[q]) { dclass[q] = 0; --tail; }
  }
  std::vector<int> fastd, gend;
  for (int q = 0; q < ND; ++q) (dclass[q] == 0 ? fastd : gend).push_back(q);
  gend.erase(std::remove_if(gend.begin(), gend.end(), [&](int q) { return dclass[q] == 2; }), gend.end());
  const int nspec = (int)gend.size();

  // 1-4 pairs (forces.py:176-226): fused into the first packed dihedral whose end atoms they are, else own list
  std::vector<int> own14(ND, -1), list14;
  for (int k = 0; k < d->n14; ++k) {
    const int i = d->idx14[2 * k], j = d->idx14[2 * k + 1];
    int owner = -1;
    for (int q : fastd) {
      const int32_t* a4 = d->dihedral_idx + 4 * q;
      if (own14[q] < 0 && ((a4[0] == i && a4[3] == j) || (a4[0] == j && a4[3] == i))) { owner = q; break; }
    }
    if (owner >= 0) own14[owner] = k; else list14.push_back(k);
  }
  auto prm14 = [&](int k, float out[3]) {  // 12 A / scnb, -6 B / scnb, K qi qj / scee  (forces.py:197-226)
    const float* p4 = d->prm14 + 4 * k;
    const double qi = d->charge ? d->charge[d->idx14[2 * k]] : 0.0, qj = d->charge ? d->charge[d->idx14[2 * k + 1]] : 0.0;
    out[0] = (float)(12.0 * (double)p4[0] / p4[2]);
    out[1] = (float)(-6.0 * (double)p4[1] / p4[2]);
    out[2] = (float)(332.06371307417066 * qi * qj / p4[3]);
  };

  // item sequences (A / B sets of 32) per packed kind
  struct Packing { std::vector<int> seq; std::vector<int> lane_a, lane_b; int nchunks = 0, npairs = 0; std::vector<int> chunk_key; };
  auto pack_kind = [&](std::vector<int> items, const int32_t* idx, int arity, const std::vector<long long>& key) {
    Packing K;
    std::stable_sort(items.begin(), items.end(), [&](int x, int y) { return key[x] < key[y]; });
    const int M = (int)items.size();
    K.nchunks = (M + 63) / 64;
    std::vector<int> sets;
    for (int c = 0; c < K.nchunks; ++c) {
      const int r = std::min(64, M - 64 * c);
      const int na = r == 64 ? 32 : (r + 1) / 2;
      sets.push_back(na);
      sets.push_back(r - na);
    }
    K.seq = spread_banks(items, idx, arity, sets);
    K.lane_a.assign((size_t)K.nchunks * 32, -1);
    K.lane_b.assign((size_t)K.nchunks * 32, -1);
    K.chunk_key.assign(K.nchunks, 0);
    int pos = 0;
    for (int c = 0; c < K.nchunks; ++c) {
      for (int l = 0; l < sets[2 * c]; ++l) K.lane_a[(size_t)c * 32 + l] = K.seq[pos++];
      for (int l = 0; l < sets[2 * c + 1]; ++l) K.lane_b[(size_t)c * 32 + l] = K.seq[pos++];
      K.npairs = c * 32 + sets[2 * c];
    }
    return K;
  };
  std::vector<int> allb(d->nbonds), alla(d->nangles);
  for (int k = 0; k < d->nbonds; ++k) allb[k] = k;
  for (int k = 0; k < d->nangles; ++k) alla[k] = k;
  std::vector<long long> dkey(ND, 0);
  for (int q = 0; q < ND; ++q) dkey[q] = dmaxper[q];
  // the 1-4 list works on its own index space
  std::vector<int32_t> idx14l((size_t)2 * std::max<size_t>(list14.size(), 1));
  std::vector<int> all14(list14.size());
  for (size_t k = 0; k < list14.size(); ++k) { all14[k] = (int)k; idx14l[2 * k] = d->idx14[2 * list14[k]]; idx14l[2 * k + 1] = d->idx14[2 * list14[k] + 1]; }
  const Packing PB = pack_kind(allb, d->bond_idx, 2, std::vector<long long>(std::max(d->nbonds, 1), 0));
  const Packing PA = pack_kind(alla, d->angle_idx, 3, std::vector<long long>(std::max(d->nangles, 1), 0));
  const Packing PD = pack_kind(fastd, d->dihedral_idx, 4, dkey);
  const Packing PP = pack_kind(all14, idx14l.data(), 2, std::vector<long long>(std::max<size_t>(list14.size(), 1), 0));
  T.nbond2 = PB.npairs; T.nangle2 = PA.npairs; T.ndih2 = PD.npairs; T.n14_2 = PP.npairs;

  // general list: special dihedrals first, then the impropers
  const int NG = nspec + d->nimpropers;
  T.ngen = NG; T.nspec = nspec; T.gen_begin = 0; T.gen_end = NG;
  std::vector<int32_t> gen_idx4((size_t)4 * std::max(NG, 1));
  std::vector<int32_t> gen_owner;
  std::vector<float> gen_prm;
  std::vector<int32_t> gen_harm;
  for (int g = 0; g < nspec; ++g) {
    for (int c = 0; c < 4; ++c) gen_idx4[4 * g + c] = d->dihedral_idx[4 * gend[g] + c];
    for (int t : dterms[gend[g]]) {
      gen_owner.push_back(g); gen_harm.push_back(d->dihedral_term_harmonic[t]);
      for (int c = 0; c < 3; ++c) gen_prm.push_back(d->dihedral_term_prm[3 * t + c]);
    }
  }
  for (int k = 0; k < d->nimpropers; ++k)
    for (int c = 0; c < 4; ++c) gen_idx4[4 * (nspec + k) + c] = d->improper_idx[4 * k + c];
  for (int t = 0; t < d->nimproper_terms; ++t) {
    gen_owner.push_back(nspec + d->improper_term_owner[t]); gen_harm.push_back(d->improper_term_harmonic[t]);
    for (int c = 0; c < 3; ++c) gen_prm.push_back(d->improper_term_prm[3 * t + c]);
  }
  std::vector<uint2> gen_terms;
  std::vector<float4> gen_tp;
  build_terms(NG, (int)gen_owner.size(), gen_owner.data(), gen_prm.data(), gen_harm.data(), gen_terms, gen_tp);
  std::vector<uint32_t> gen_idx(std::max(NG, 1), 0);
  for (int g = 0; g < NG; ++g) gen_idx[g] = pack4(gen_idx4.data() + 4 * g, 4);

  // ---- atom-major force slots (slot mode, Forces.compute): contribution (item, position) of atom a gets the
  // next slot of a, in the fixed order bonds, angles, packed dihedrals, general torsions, 1-4 list => the per-atom
  // gather adds them in a fixed order.  Lanes without an item write to the trash slot `nslots`.
  struct KindRef { const std::vector<int>* seq; const int32_t* idx; int arity; };
  std::vector<int> genseq(NG);
  for (int g = 0; g < NG; ++g) genseq[g] = g;
  const KindRef kinds[5] = {{&PB.seq, d->bond_idx, 2}, {&PA.seq, d->angle_idx, 3}, {&PD.seq, d->dihedral_idx, 4},
                            {&genseq, gen_idx4.data(), 4}, {&PP.seq, idx14l.data(), 2}};
  std::vector<int> per_atom_count(kMaxAtoms, 0);
  for (const KindRef& K : kinds)
    for (int it : *K.seq) for (int c = 0; c < K.arity; ++c) per_atom_count[K.idx[K.arity * it + c]]++;
  std::vector<uint16_t> gstart(kMaxAtoms + 2, 0);
  int nslots = 0;
  // an atom's range is padded to a multiple of 4 slots (the padding is zeroed at staging and never written): the gather
  // reads it with LDS.128
  for (int a = 0; a < kMaxAtoms; ++a) { gstart[a] = (uint16_t)nslots; nslots += (per_atom_count[a] + 3) & ~3; }
  gstart[kMaxAtoms] = gstart[kMaxAtoms + 1] = (uint16_t)nslots;
  if (nslots > 65534) { delete topo; fail(B200MD_ERR_UNSUPPORTED, "too many bonded force contributions"); return nullptr; }
  T.nslots = nslots + 1;  // + the trash slot
  std::vector<int> cursor(gstart.begin(), gstart.begin() + kMaxAtoms);
  std::vector<std::vector<uint16_t>> slot_of(5);  // [kind][item * arity + c], indexed by ORIGINAL item id
  for (int ty = 0; ty < 5; ++ty) {
    const KindRef& K = kinds[ty];
    int maxid = -1;
    for (int it : *K.seq) maxid = std::max(maxid, it);
    slot_of[ty].assign((size_t)(maxid + 1) * K.arity, (uint16_t)nslots);
    for (int it : *K.seq) for (int c = 0; c < K.arity; ++c) slot_of[ty][(size_t)it * K.arity + c] = (uint16_t)cursor[K.idx[K.arity * it + c]]++;
  }
  auto pair16 = [](uint16_t lo, uint16_t hi) { return (uint32_t)lo | ((uint32_t)hi << 16); };
  const uint16_t trash = (uint16_t)nslots;
  auto slots4 = [&](int ty, int it, int arity) {  // (s0 | s1 << 16, s2 | s3 << 16) of one item, trash for a missing item
    uint16_t sl[4] = {trash, trash, trash, trash};
    if (it >= 0) for (int c = 0; c < arity; ++c) sl[c] = slot_of[ty][(size_t)it * arity + c];
    return make_uint2(pair16(sl[0], sl[1]), pair16(sl[2], sl[3]));
  };
  // a lane half without an item repeats the atoms of the lane's A item (finite geometry) with zero parameters
  auto atoms_of = [&](const int32_t* idx, int arity, int it, int fallback) { return pack4(idx + (size_t)arity * (it >= 0 ? it : fallback), arity); };

  std::vector<uint2> bond2_idx(std::max(PB.npairs, 1)), angle2_idx(std::max(PA.npairs, 1)), dih2_idx(std::max(PD.npairs, 1)), p142_idx(std::max(PP.npairs, 1));
  std::vector<uint4> bond2_slot(bond2_idx.size()), angle2_slot(angle2_idx.size()), dih2_slot(dih2_idx.size()), p142_slot(p142_idx.size());
  std::vector<float4> bond2_prm(bond2_idx.size()), angle2_prm(angle2_idx.size()), dih2_p14ab(dih2_idx.size()), p142_ab(p142_idx.size());
  std::vector<float2> dih2_k0(dih2_idx.size()), dih2_p14q(dih2_idx.size()), p142_q(p142_idx.size());
  const size_t dstride = dih2_idx.size();
  std::vector<float4> dih2_coef((size_t)kMaxPer * dstride, make_float4(0.f, 0.f, 0.f, 0.f));  // [n-1][pair]: n a_n (A,B), n b_n (A,B)
  auto u4 = [](uint2 a2, uint2 b2) { return make_uint4(a2.x, a2.y, b2.x, b2.y); };
  for (int p2 = 0; p2 < PB.npairs; ++p2) {
    const int A = PB.lane_a[p2], Bi = PB.lane_b[p2];
    bond2_idx[p2] = make_uint2(atoms_of(d->bond_idx, 2, A, A), atoms_of(d->bond_idx, 2, Bi, A));
    bond2_slot[p2] = u4(slots4(0, A, 2), slots4(0, Bi, 2));
    bond2_prm[p2] = make_float4(d->bond_prm[2 * A], Bi >= 0 ? d->bond_prm[2 * Bi] : 0.f, d->bond_prm[2 * A + 1], Bi >= 0 ? d->bond_prm[2 * Bi + 1] : 1.f);
  }
  for (int p2 = 0; p2 < PA.npairs; ++p2) {
    const int A = PA.lane_a[p2], Bi = PA.lane_b[p2];
    angle2_idx[p2] = make_uint2(atoms_of(d->angle_idx, 3, A, A), atoms_of(d->angle_idx, 3, Bi, A));
    angle2_slot[p2] = u4(slots4(1, A, 3), slots4(1, Bi, 3));
    angle2_prm[p2] = make_float4(d->angle_prm[2 * A], Bi >= 0 ? d->angle_prm[2 * Bi] : 0.f, d->angle_prm[2 * A + 1], Bi >= 0 ? d->angle_prm[2 * Bi + 1] : 1.f);
  }
  std::vector<int> dih_chunk_nmax(std::max(PD.nchunks, 1), 1);
  for (int p2 = 0; p2 < PD.npairs; ++p2) {
    const int it[2] = {PD.lane_a[p2], PD.lane_b[p2]};
    dih2_idx[p2] = make_uint2(atoms_of(d->dihedral_idx, 4, it[0], it[0]), atoms_of(d->dihedral_idx, 4, it[1], it[0]));
    dih2_slot[p2] = u4(slots4(2, it[0], 4), slots4(2, it[1], 4));
    float k0s[2] = {0.f, 0.f}, ab[2][2] = {{0.f, 0.f}, {0.f, 0.f}}, qq[2] = {0.f, 0.f};
    for (int h = 0; h < 2; ++h) {
      if (it[h] < 0) continue;
      dih_chunk_nmax[p2 / 32] = std::max(dih_chunk_nmax[p2 / 32], dmaxper[it[h]]);
      double K0 = 0.0, ca[kMaxPer] = {0}, cb[kMaxPer] = {0};
      for (int t : dterms[it[h]]) {  // k (1 + cos(n phi - phi0)) = k + k cos(phi0) cos(n phi) + k sin(phi0) sin(n phi)
        const double k0 = d->dihedral_term_prm[3 * t], phi0 = d->dihedral_term_prm[3 * t + 1];
        const int n = (int)d->dihedral_term_prm[3 * t + 2];
        K0 += k0; ca[n - 1] += n * k0 * cos(phi0); cb[n - 1] += n * k0 * sin(phi0);
      }
      k0s[h] = (float)K0;
      for (int n = 0; n < kMaxPer; ++n) {
        float4& cf = dih2_coef[(size_t)n * dstride + p2];
        (h ? cf.y : cf.x) = (float)ca[n];
        (h ? cf.w : cf.z) = (float)cb[n];
      }
      if (own14[it[h]] >= 0) { float o3[3]; prm14(own14[it[h]], o3); ab[h][0] = o3[0]; ab[h][1] = o3[1]; qq[h] = o3[2]; }
    }
    dih2_k0[p2] = make_float2(k0s[0], k0s[1]);
    dih2_p14ab[p2] = make_float4(ab[0][0], ab[1][0], ab[0][1], ab[1][1]);
    dih2_p14q[p2] = make_float2(qq[0], qq[1]);
  }
  for (int p2 = 0; p2 < PP.npairs; ++p2) {
    const int it[2] = {PP.lane_a[p2], PP.lane_b[p2]};
    p142_idx[p2] = make_uint2(atoms_of(idx14l.data(), 2, it[0], it[0]), atoms_of(idx14l.data(), 2, it[1], it[0]));
    p142_slot[p2] = u4(slots4(4, it[0], 2), slots4(4, it[1], 2));
    float o3[2][3] = {{0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}};
    for (int h = 0; h < 2; ++h) if (it[h] >= 0) prm14(list14[it[h]], o3[h]);
    p142_ab[p2] = make_float4(o3[0][0], o3[1][0], o3[0][1], o3[1][1]);
    p142_q[p2] = make_float2(o3[0][2], o3[1][2]);
  }
  std::vector<uint2> gen_slot(std::max(NG, 1));
  for (int g = 0; g < NG; ++g) gen_slot[g] = slots4(3, g, 4);
  T.dih_on = 1;
  T.p14_on = 1;
  T.dih2_stride = (int)dstride;

  // ---- static task schedule: longest-processing-time-first over the 4 warps -----------------------
  // Cost model = issue slots per warp of the force-only tasks, tuned on the GPU with tools/tune_costs.py (the
  // energies are a separate, evenly strided pass).  B200MD_COSTS = "pass,chunk,diaggroup,diag,bond,angle,dih,gen,p14"
  // overrides it for tuning runs.
  struct Task { uint32_t word; int cost; };
  int cms[1][9] = {{126, 144, 52, 30, 77, 230, 380, 400, 100}};  // profiles/r02_schedule_tuning.log
  const char* envs[1] = {getenv("B200MD_COSTS")};
  const int nblocks = (N + 31) / 32;
  std::vector<uint32_t> tasks_v[1];
  for (int sch = 0; sch < 1; ++sch) {
    int* cm = cms[sch];
    if (envs[sch]) sscanf(envs[sch], "%d,%d,%d,%d,%d,%d,%d,%d,%d", cm, cm + 1, cm + 2, cm + 3, cm + 4, cm + 5, cm + 6, cm + 7, cm + 8);
    std::vector<Task> all;
    if (d->npairs > 0) {
      for (int c = 4; c < T.nchunks; ++c) {  // off-diagonal: chunk of block J against the full blocks 0..J-1
        const int passes = c / 4;
        all.push_back({make_task(TASK_TILE, c, passes), passes * cm[0] + cm[1]});
      }
      for (int b = 0; b < nblocks; ++b) {  // diagonal blocks, full shell: tasks of up to 4 groups of 4 j atoms
        const int n = std::min(32, N - 32 * b), groups = (n + 3) / 4;
        for (int g0 = 0; g0 < groups; g0 += 4) {
          const int ng = std::min(4, groups - g0);
          all.push_back({make_task(TASK_DIAG, (uint32_t)b | ((uint32_t)g0 << 8), ng), ng * cm[2] + cm[3]});
        }
      }
    }
    // bonded chunks: 32 lanes x 2 items (packed kinds) or 32 items (general torsions)
    for (int c = 0; c < PB.nchunks; ++c) all.push_back({make_task(TASK_BOND, c, 0), cm[4]});
    for (int c = 0; c < PA.nchunks; ++c) all.push_back({make_task(TASK_ANGLE, c, 0), cm[5]});
    for (int c = 0; c < PD.nchunks; ++c) all.push_back({make_task(TASK_DIHEDRAL, c, dih_chunk_nmax[c]), cm[6] + 12 * dih_chunk_nmax[c]});
    for (int c = 0; c < (NG + 31) / 32; ++c) all.push_back({make_task(TASK_IMPROPER, c, 0), cm[7]});
    for (int c = 0; c < PP.nchunks; ++c) all.push_back({make_task(TASK_P14, c, 0), cm[8]});
    std::stable_sort(all.begin(), all.end(), [](const Task& a, const Task& b) { return a.cost > b.cost; });
    std::vector<uint32_t>& tasks = tasks_v[sch];
    tasks.assign((size_t)kWarps * kMaxTasks, 0);
    int load[kWarps] = {0, 0, 0, 0};
    std::vector<std::vector<uint32_t>> mine(kWarps);
    for (const Task& t : all) {
      int w = 0;
      for (int k = 1; k < kWarps; ++k) if (load[k] < load[w]) w = k;
      if ((int)mine[w].size() >= kMaxTasks - 1) { delete topo; fail(B200MD_ERR_UNSUPPORTED, "too many tasks per warp"); return nullptr; }
      mine[w].push_back(t.word);
      load[w] += t.cost;
    }
    // each warp runs its tasks kind by kind (run_tasks): tiles, diagonal blocks, dihedrals, general torsions, angles, bonds, 1-4
    const uint32_t kind_order[kNumKinds] = {TASK_TILE, TASK_DIAG, TASK_DIHEDRAL, TASK_IMPROPER, TASK_ANGLE, TASK_BOND, TASK_P14};
    for (int w = 0; w < kWarps; ++w) {
      int n = 0;
      uint2& k8 = T.kend8[w];
      k8 = make_uint2(0u, 0u);
      for (int k = 0; k < kNumKinds; ++k) {
        for (uint32_t word : mine[w]) if ((word & 0xff) == kind_order[k]) tasks[(size_t)w * kMaxTasks + n++] = word;
        (k < 4 ? k8.x : k8.y) |= (uint32_t)n << (8 * (k & 3));
      }
    }
  }

  // ---- rigid X-H bonds (include/b200md.h, b200md_integrate_opts.rigid_bonds) -------------------------
  std::vector<int4> cgroup(kMaxAtoms, make_int4(0, 0, 0, 0));
  std::vector<float4> cdist2(kMaxAtoms, make_float4(1.0f, 1.0f, 1.0f, 0.0f));
  std::vector<float4> cweight(kMaxAtoms, make_float4(1.0f, 1.0f, 1.0f, 1.0f));
  T.ncons = 0; T.ngroups = 0; T.cnmax = 0;
  {
    auto is_h = [&](int a) { return d->mass[a] > 0.5f && d->mass[a] < 3.5f; };
    std::vector<int> partner(N, -1);
    std::vector<float> req(N, 0.0f);
    for (int b = 0; b < d->nbonds; ++b) {  // a hydrogen is tied through its shortest bond row (Urey-Bradley rows are longer)
      const int i = d->bond_idx[2 * b], j = d->bond_idx[2 * b + 1];
      const float r0 = d->bond_prm[2 * b + 1];
      for (int side = 0; side < 2; ++side) {
        const int hh = side ? j : i, other = side ? i : j;
        if (is_h(hh) && (partner[hh] < 0 || r0 < req[hh])) { partner[hh] = other; req[hh] = r0; }
      }
    }
    std::vector<std::vector<int>> kids(N);
    for (int a = 0; a < N && T.ncons >= 0; ++a) {
      if (partner[a] < 0) continue;
      const int p = partner[a];
      if (is_h(p) || kids[p].size() >= 3 || !(req[a] > 0.0f)) { T.ncons = -1; break; }
      kids[p].push_back(a);
      T.ncons += 1;
    }
    if (T.ncons > 0) {
      std::vector<int> parents;
      for (int p = 0; p < N; ++p) if (!kids[p].empty()) parents.push_back(p);
      std::stable_sort(parents.begin(), parents.end(), [&](int a, int b) { return kids[a].size() > kids[b].size(); });
      T.ngroups = (int)parents.size();
      for (int gi = 0; gi < T.ngroups; ++gi) {
        const int p = parents[gi], n = (int)kids[p].size();
        int hh[3] = {p, p, p};
        float dd[3] = {1.0f, 1.0f, 1.0f}, ww[3] = {1.0f, 1.0f, 1.0f};
        for (int k = 0; k < n; ++k) { hh[k] = kids[p][k]; dd[k] = req[hh[k]] * req[hh[k]]; ww[k] = 1.0f / d->mass[hh[k]]; }
        cgroup[gi] = make_int4(p | (n << 8), hh[0], hh[1], hh[2]);
        cdist2[gi] = make_float4(dd[0], dd[1], dd[2], 0.0f);
        cweight[gi] = make_float4(1.0f / d->mass[p], ww[0], ww[1], ww[2]);
        const int w = gi / 32;
        const uint32_t cur = (T.cnmax >> (8 * w)) & 0xffu;
        if ((uint32_t)n > cur) T.cnmax = (T.cnmax & ~(0xffu << (8 * w))) | ((uint32_t)n << (8 * w));
      }
    }
  }

  // ---- proper-dihedral angle output (calc_phi) ------------------------------------------------------
  std::vector<uchar4> phi_idx;
  std::vector<uint16_t> phi_start(kMaxAtoms + 1, 0), phi_items;
  T.nphi = d->ndihedrals;
  if (T.nphi > 0x3fff) { delete topo; fail(B200MD_ERR_UNSUPPORTED, "more than 16383 dihedrals"); return nullptr; }
  for (int q = 0; q < d->ndihedrals; ++q) {
    const int32_t* a4 = d->dihedral_idx + 4 * q;
    phi_idx.push_back(make_uchar4((unsigned char)a4[0], (unsigned char)a4[1], (unsigned char)a4[2], (unsigned char)a4[3]));
  }
  for (int a = 0; a < kMaxAtoms; ++a) {
    phi_start[a] = (uint16_t)phi_items.size();
    for (int q = 0; q < d->ndihedrals; ++q)
      for (int role = 0; role < 4; ++role)
        if (d->dihedral_idx[4 * q + role] == a) phi_items.push_back((uint16_t)(q | (role << 14)));
  }
  phi_start[kMaxAtoms] = (uint16_t)phi_items.size();

  // ---- FP64 energy pass: plain item lists in topology order (energy_pass, b200md_forces.cuh) -----------
  std::vector<uint32_t> e_pair(std::max(d->npairs, 1), 0), e14_idx(std::max(d->n14, 1), 0), eb_idx(std::max(d->nbonds, 1), 0),
      ea_idx(std::max(d->nangles, 1), 0), et_idx(std::max(ND + d->nimpropers, 1), 0);
  std::vector<double2> e_ab((size_t)NT * NT);
  std::vector<double> e14_prm((size_t)3 * std::max(d->n14, 1), 0.0);
  std::vector<float2> eb_prm(std::max(d->nbonds, 1)), ea_prm(std::max(d->nangles, 1));
  for (int k = 0; k < NT * NT; ++k) e_ab[k] = make_double2((double)d->lj_a[k], (double)d->lj_b[k]);
  for (int k = 0; k < d->npairs; ++k) {
    const int i = d->pair_idx[2 * k], j = d->pair_idx[2 * k + 1];
    e_pair[k] = (uint32_t)i | ((uint32_t)j << 8) | ((uint32_t)(d->atom_type[i] * NT + d->atom_type[j]) << 16);
  }
  for (int k = 0; k < d->n14; ++k) {  // forces.py:197-226: A / scnb, B / scnb, K qi qj / scee
    const int i = d->idx14[2 * k], j = d->idx14[2 * k + 1];
    const float* p4 = d->prm14 + 4 * k;
    const double qi = d->charge ? d->charge[i] : 0.0, qj = d->charge ? d->charge[j] : 0.0;
    e14_idx[k] = (uint32_t)i | ((uint32_t)j << 8);
    e14_prm[3 * k] = (double)p4[0] / (double)p4[2];
    e14_prm[3 * k + 1] = (double)p4[1] / (double)p4[2];
    e14_prm[3 * k + 2] = 332.06371307417066 * qi * qj / (double)p4[3];
  }
  for (int k = 0; k < d->nbonds; ++k) { eb_idx[k] = pack4(d->bond_idx + 2 * k, 2); eb_prm[k] = make_float2(d->bond_prm[2 * k], d->bond_prm[2 * k + 1]); }
  for (int k = 0; k < d->nangles; ++k) { ea_idx[k] = pack4(d->angle_idx + 3 * k, 3); ea_prm[k] = make_float2(d->angle_prm[2 * k], d->angle_prm[2 * k + 1]); }
  // torsion terms grouped by their quadruple, slice order kept inside a group (forces.py:559-576); dihedrals, then impropers
  std::vector<int> et_start(ND + d->nimpropers + 1, 0);
  std::vector<float4> et_term(std::max(d->ndihedral_terms + d->nimproper_terms, 1));
  std::vector<double2> et_cs(et_term.size());
  {
    for (int q = 0; q < ND; ++q) et_idx[q] = pack4(d->dihedral_idx + 4 * q, 4);
    for (int q = 0; q < d->nimpropers; ++q) et_idx[ND + q] = pack4(d->improper_idx + 4 * q, 4);
    for (int t = 0; t < d->ndihedral_terms; ++t) ++et_start[d->dihedral_term_owner[t] + 1];
    for (int t = 0; t < d->nimproper_terms; ++t) ++et_start[ND + d->improper_term_owner[t] + 1];
    for (int q = 0; q < ND + d->nimpropers; ++q) et_start[q + 1] += et_start[q];
    std::vector<int> cur(et_start.begin(), et_start.end() - 1);
    auto put = [&](int owner, const float* prm3, int harmonic) {
      const float per = prm3[2];
      const bool integer = !harmonic && per >= 1.0f && per <= 64.0f && per == floorf(per);
      const int at = cur[owner]++;
      et_term[at] = make_float4(prm3[0], prm3[1], per, harmonic ? 1.0f : (integer ? 2.0f : 0.0f));
      et_cs[at] = make_double2(cos((double)prm3[1]), sin((double)prm3[1]));
    };
    for (int t = 0; t < d->ndihedral_terms; ++t) put(d->dihedral_term_owner[t], d->dihedral_term_prm + 3 * t, d->dihedral_term_harmonic[t]);
    for (int t = 0; t < d->nimproper_terms; ++t) put(ND + d->improper_term_owner[t], d->improper_term_prm + 3 * t, d->improper_term_harmonic[t]);
  }
  T.ne_pair = d->npairs; T.ne_14 = d->n14; T.ne_bond = d->nbonds; T.ne_angle = d->nangles; T.ne_dih = ND; T.ne_imp = d->nimpropers;

  // ---- upload --------------------------------------------------------------------------------------
  BlobBuilder B;
  const size_t o_ab = B.add(ab), o_q = B.add(charge), o_m = B.add(mass), o_ro = B.add(rowoff), o_mask = B.add(mask8), o_rot = B.add(diagmask);
  const size_t o_bi = B.add(bond2_idx), o_bp = B.add(bond2_prm), o_bs = B.add(bond2_slot);
  const size_t o_ai = B.add(angle2_idx), o_ap = B.add(angle2_prm), o_as = B.add(angle2_slot);
  const size_t o_di = B.add(dih2_idx), o_dc = B.add(dih2_coef), o_dk = B.add(dih2_k0), o_ds = B.add(dih2_slot), o_d14 = B.add(dih2_p14ab), o_dq = B.add(dih2_p14q);
  const size_t o_gi = B.add(gen_idx), o_gt = B.add(gen_terms), o_gp = B.add(gen_tp), o_gsl = B.add(gen_slot);
  const size_t o_pi = B.add(p142_idx), o_pp = B.add(p142_ab), o_pq = B.add(p142_q), o_ps = B.add(p142_slot);
  const size_t o_gs = B.add(gstart), o_tk = B.add(tasks_v[0]);
  const size_t o_ep = B.add(e_pair), o_eab = B.add(e_ab), o_e14i = B.add(e14_idx), o_e14p = B.add(e14_prm), o_ebi = B.add(eb_idx),
               o_ebp = B.add(eb_prm), o_eai = B.add(ea_idx), o_eap = B.add(ea_prm), o_eti = B.add(et_idx), o_ets = B.add(et_start),
               o_ett = B.add(et_term), o_etc = B.add(et_cs);
  const size_t o_cg = B.add(cgroup), o_cd = B.add(cdist2), o_cw = B.add(cweight);
  const size_t o_phi = B.add(phi_idx), o_phs = B.add(phi_start), o_phit = B.add(phi_items);
  auto cleanup = [&](const std::string& msg) { fail(B200MD_ERR_CUDA, msg); if (topo->blob) cudaFree(topo->blob); delete topo; };
  DeviceGuard guard(device);
  cudaError_t e = guard.err;
  if (e != cudaSuccess) { cleanup(std::string("cudaSetDevice: ") + cudaGetErrorString(e)); return nullptr; }
  e = cudaMalloc(&topo->blob, B.bytes.size());
  if (e != cudaSuccess) { topo->blob = nullptr; cleanup(std::string("cudaMalloc: ") + cudaGetErrorString(e)); return nullptr; }
  e = cudaMemcpy(topo->blob, B.bytes.data(), B.bytes.size(), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) { cleanup(std::string("cudaMemcpy: ") + cudaGetErrorString(e)); return nullptr; }
  char* base = (char*)topo->blob;
  T.ab = (const float2*)(base + o_ab); T.charge = (const float*)(base + o_q); T.mass = (const float*)(base + o_m);
  T.ab_row_off = (const int*)(base + o_ro); T.mask8 = (const uint8_t*)(base + o_mask); T.diagmask = (const uint32_t*)(base + o_rot);
  T.bond2_idx = (const uint2*)(base + o_bi); T.bond2_prm = (const float4*)(base + o_bp); T.bond2_slot = (const uint4*)(base + o_bs);
  T.angle2_idx = (const uint2*)(base + o_ai); T.angle2_prm = (const float4*)(base + o_ap); T.angle2_slot = (const uint4*)(base + o_as);
  T.dih2_idx = (const uint2*)(base + o_di); T.dih2_coef = (const float4*)(base + o_dc); T.dih2_k0 = (const float2*)(base + o_dk);
  T.dih2_slot = (const uint4*)(base + o_ds); T.dih2_p14ab = (const float4*)(base + o_d14); T.dih2_p14q = (const float2*)(base + o_dq);
  T.gen_idx = (const uint32_t*)(base + o_gi); T.gen_terms = (const uint2*)(base + o_gt); T.gen_term_prm = (const float4*)(base + o_gp);
  T.gen_slot = (const uint2*)(base + o_gsl);
  T.p142_idx = (const uint2*)(base + o_pi); T.p142_ab = (const float4*)(base + o_pp); T.p142_q = (const float2*)(base + o_pq);
  T.p142_slot = (const uint4*)(base + o_ps);
  T.gather_start = (const uint16_t*)(base + o_gs);
  T.tasks = (const uint32_t*)(base + o_tk);
  T.e_pair = (const uint32_t*)(base + o_ep); T.e_ab = (const double2*)(base + o_eab);
  T.e14_idx = (const uint32_t*)(base + o_e14i); T.e14_prm = (const double*)(base + o_e14p);
  T.eb_idx = (const uint32_t*)(base + o_ebi); T.eb_prm = (const float2*)(base + o_ebp);
  T.ea_idx = (const uint32_t*)(base + o_eai); T.ea_prm = (const float2*)(base + o_eap);
  T.et_idx = (const uint32_t*)(base + o_eti); T.et_start = (const int*)(base + o_ets);
  T.et_term = (const float4*)(base + o_ett); T.et_cs = (const double2*)(base + o_etc);
  T.phi_idx = (const uchar4*)(base + o_phi); T.phi_start = (const uint16_t*)(base + o_phs); T.phi_items = (const uint16_t*)(base + o_phit);
  T.cgroup = (const int4*)(base + o_cg); T.cdist2 = (const float4*)(base + o_cd); T.cweight = (const float4*)(base + o_cw);
  T.smem_bytes = (int)smem_layout(nullptr, nullptr, NT, T.nslots, false);
  T.smem_bytes_fixed = (int)smem_layout(nullptr, nullptr, NT, T.nslots, true);
  if (T.smem_bytes > 227 * 1024) { cleanup("topology needs more than 227 KB of shared memory per CTA"); g_last_error = "too much shared memory"; return nullptr; }
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, device);
  topo->num_sms = prop.multiProcessorCount;
  topo->occ_forces[0] = query_occupancy(forces_kernel<0>, T.smem_bytes);
  topo->occ_forces[1] = query_occupancy(forces_kernel<1>, T.smem_bytes);
  topo->occ_energy[0] = std::max(1, query_occupancy(energy_kernel<0>, 0));
  topo->occ_energy[1] = std::max(1, query_occupancy(energy_kernel<1>, 0));
  topo->occ_integrate[0] = query_occupancy(integrate_kernel<0, false>, T.smem_bytes_fixed);
  topo->occ_integrate[1] = query_occupancy(integrate_kernel<1, false>, T.smem_bytes_fixed);
  query_occupancy(integrate_kernel<0, true>, T.smem_bytes_fixed);  // also raises the dynamic shared-memory limit
  query_occupancy(integrate_kernel<1, true>, T.smem_bytes_fixed);
  if (topo->occ_forces[0] < 1 || topo->occ_integrate[0] < 1) { cleanup("kernel does not fit on an SM"); return nullptr; }
  g_last_error.clear();
  return topo;
}

// ---- solutes above 128 atoms: packing (b200md_large.cuh) ---------------------------------------------------------
static b200md_topology* create_large(const b200md_topology_desc* d, int device) {
  const int N = d->natoms, NT = d->ntypes;
  if (NT > 104) { fail(B200MD_ERR_UNSUPPORTED, "more than 104 LJ types: the type-pair tables must fit in shared memory"); return nullptr; }
  auto* topo = new b200md_topology();
  topo->device = device;
  topo->natoms = N;
  topo->large = true;
  topo->dev.natoms = N; topo->dev.ntypes = NT;
  topo->dev.nphi = d->ndihedrals;
  const int W = (N + 31) / 32, NP = 32 * W;
  topo->mask_words = W;
  // exclusion bit matrix from the pair list (= Forces.ava_idx): bit set <=> NOT evaluated; symmetric (full shell)
  std::vector<uint32_t> mask((size_t)NP * W, 0xffffffffu);
  for (int k = 0; k < d->npairs; ++k) {
    const int i = d->pair_idx[2 * k], j = d->pair_idx[2 * k + 1];
    if (i == j) continue;
    mask[(size_t)i * W + j / 32] &= ~(1u << (j % 32));
    mask[(size_t)j * W + i / 32] &= ~(1u << (i % 32));
  }
  topo->mask_rows.assign(mask.begin(), mask.begin() + (size_t)N * W);
  std::vector<int> atype(NP, 0);
  std::vector<float> charge(NP, 0.f), mass(NP, 1.f), t12a((size_t)NT * NT), t6b((size_t)NT * NT);
  for (int a = 0; a < N; ++a) {
    atype[a] = d->atom_type[a];
    charge[a] = d->charge ? d->charge[a] : 0.f;
    mass[a] = d->mass ? d->mass[a] : 1.f;
  }
  std::vector<double2> tab64((size_t)NT * NT);
  for (int k = 0; k < NT * NT; ++k) { t12a[k] = 12.0f * d->lj_a[k]; t6b[k] = 6.0f * d->lj_b[k]; tab64[k] = make_double2((double)d->lj_a[k], (double)d->lj_b[k]); }
  std::vector<int2> bidx(d->nbonds), i14(d->n14);
  std::vector<float2> bprm(d->nbonds), aprm(d->nangles);
  std::vector<int> aidx(d->angle_idx, d->angle_idx + 3 * (size_t)d->nangles);
  std::vector<float4> p14(d->n14);
  for (int k = 0; k < d->nbonds; ++k) { bidx[k] = make_int2(d->bond_idx[2 * k], d->bond_idx[2 * k + 1]); bprm[k] = make_float2(d->bond_prm[2 * k], d->bond_prm[2 * k + 1]); }
  for (int k = 0; k < d->nangles; ++k) aprm[k] = make_float2(d->angle_prm[2 * k], d->angle_prm[2 * k + 1]);
  for (int k = 0; k < d->n14; ++k) { i14[k] = make_int2(d->idx14[2 * k], d->idx14[2 * k + 1]); p14[k] = make_float4(d->prm14[4 * k], d->prm14[4 * k + 1], d->prm14[4 * k + 2], d->prm14[4 * k + 3]); }
  // torsion terms grouped by their quadruple, slice order kept inside a group (forces.py:559-576)
  auto group_terms = [&](int n, const int32_t* idx, int nt, const int32_t* owner, const float* prm, const int32_t* harm,
                         std::vector<int4>& q, std::vector<int>& start, std::vector<float4>& term, std::vector<double2>& cs) -> bool {
    q.resize(n); start.assign(n + 1, 0); term.resize(nt); cs.resize(nt);
    for (int k = 0; k < n; ++k) q[k] = make_int4(idx[4 * k], idx[4 * k + 1], idx[4 * k + 2], idx[4 * k + 3]);
    for (int t = 0; t < nt; ++t) { if (owner[t] < 0 || owner[t] >= n) return false; ++start[owner[t] + 1]; }
    for (int k = 0; k < n; ++k) start[k + 1] += start[k];
    std::vector<int> cur(start.begin(), start.end() - 1);
    for (int t = 0; t < nt; ++t) {
      const float per = prm[3 * t + 2];
      const bool integer = !harm[t] && per >= 1.0f && per <= 6.0f && per == floorf(per);
      const int at = cur[owner[t]]++;
      term[at] = make_float4(prm[3 * t], prm[3 * t + 1], per, harm[t] ? 1.0f : (integer ? 2.0f : 0.0f));
      cs[at] = make_double2(cos((double)prm[3 * t + 1]), sin((double)prm[3 * t + 1]));
    }
    return true;
  };
  std::vector<int4> dq, iq; std::vector<int> ds, is; std::vector<float4> dt, it; std::vector<double2> dcs, ics;
  if (!group_terms(d->ndihedrals, d->dihedral_idx, d->ndihedral_terms, d->dihedral_term_owner, d->dihedral_term_prm, d->dihedral_term_harmonic, dq, ds, dt, dcs) ||
      !group_terms(d->nimpropers, d->improper_idx, d->nimproper_terms, d->improper_term_owner, d->improper_term_prm, d->improper_term_harmonic, iq, is, it, ics)) {
    delete topo; fail(B200MD_ERR_INVALID_ARGUMENT, "torsion term owner out of range"); return nullptr;
  }
  // bonded forces by windows of kLWin atoms (b200md_large.cuh, large_bonded_forces_kernel): per window the items that
  // touch it, and per (item, role) the shared-memory slot of the contribution -- atom-major inside the window, in the fixed
  // order bonds, angles, dihedrals, 1-4 pairs, impropers
  const long long nitems = (long long)d->nbonds + d->nangles + d->ndihedrals + d->n14 + d->nimpropers;
  if (nitems >= (1ll << 31)) { delete topo; fail(B200MD_ERR_UNSUPPORTED, "more than 2^31 bonded items"); return nullptr; }
  const int nwin = (N + kLWin - 1) / kLWin;
  std::vector<int> gstart(N + 1, 0), win_start(nwin + 1, 0);
  std::vector<int4> win_atoms, win_kend(nwin, make_int4(0, 0, 0, 0));
  std::vector<float4> win_prm;
  std::vector<ushort4> win_slot;
  int win_slots_max = 1, halo_max = 0;
  std::vector<int> win_lo(nwin, 0), win_n(nwin, 0);
  {
    struct Kind { const int32_t* idx; int n, arity; };
    const Kind kinds[5] = {{d->bond_idx, d->nbonds, 2}, {d->angle_idx, d->nangles, 3}, {d->dihedral_idx, d->ndihedrals, 4},
                           {d->idx14, d->n14, 2}, {d->improper_idx, d->nimpropers, 4}};
    for (const Kind& K : kinds)
      for (int it = 0; it < K.n; ++it)
        for (int c = 0; c < K.arity; ++c) ++gstart[K.idx[K.arity * it + c] + 1];
    for (int a = 0; a < N; ++a) gstart[a + 1] += gstart[a];
    for (int w = 0; w < nwin; ++w) win_slots_max = std::max(win_slots_max, gstart[std::min(N, kLWin * (w + 1))] - gstart[kLWin * w]);
    if (win_slots_max >= 0xffff || (size_t)win_slots_max * 16 > 190 * 1024) {
      delete topo; fail(B200MD_ERR_UNSUPPORTED, "a window of 64 atoms carries more bonded contributions than fit in shared memory"); return nullptr;
    }
    std::vector<int> cur(gstart.begin(), gstart.end() - 1);
    std::vector<std::vector<std::pair<uint32_t, ushort4>>> per(nwin);
    uint32_t g = 0;
    for (const Kind& K : kinds)
      for (int it = 0; it < K.n; ++it, ++g) {
        int wins[4], nw = 0;
        unsigned short sl[4][4];
        for (int c = 0; c < K.arity; ++c) {
          const int a = K.idx[K.arity * it + c], w = a / kLWin;
          int k = 0;
          while (k < nw && wins[k] != w) ++k;
          if (k == nw) { wins[nw++] = w; for (int q = 0; q < 4; ++q) sl[k][q] = 0xffff; }
          sl[k][c] = (unsigned short)(cur[a]++ - gstart[kLWin * w]);
        }
        for (int k = 0; k < nw; ++k) per[wins[k]].push_back({g, make_ushort4(sl[k][0], sl[k][1], sl[k][2], sl[k][3])});
      }
    std::vector<int> lo(nwin, N), hi(nwin, -1);  // atoms the items of a window reference: [lo, hi]
    g = 0;
    for (const Kind& K : kinds)
      for (int it = 0; it < K.n; ++it, ++g) {
        int amin = N, amax = -1;
        for (int c = 0; c < K.arity; ++c) { amin = std::min(amin, (int)K.idx[K.arity * it + c]); amax = std::max(amax, (int)K.idx[K.arity * it + c]); }
        for (int c = 0; c < K.arity; ++c) {
          const int w = K.idx[K.arity * it + c] / kLWin;
          lo[w] = std::min(lo[w], amin); hi[w] = std::max(hi[w], amax);
        }
      }
    for (int w = 0; w < nwin; ++w) {
      win_start[w + 1] = win_start[w] + (int)per[w].size();
      int kend[4] = {win_start[w + 1], win_start[w + 1], win_start[w + 1], win_start[w + 1]};  // first entry of kinds 1 .. 4 (entries are in kind order)
      for (const auto& e : per[w]) {
        const int gi = (int)e.first, at = (int)win_slot.size();
        const int e1 = d->nbonds, e2 = e1 + d->nangles, e3 = e2 + d->ndihedrals, e4 = e3 + d->n14;
        const int kind = (gi >= e1) + (gi >= e2) + (gi >= e3) + (gi >= e4);
        for (int q = kind; q >= 1; --q) kend[q - 1] = std::min(kend[q - 1], at);
        int4 qa = make_int4(0, 0, 0, 0);
        float4 pr = make_float4(0.f, 0.f, 0.f, 0.f);
        if (kind == 0) {
          qa.x = bidx[gi].x; qa.y = bidx[gi].y; pr.x = bprm[gi].x; pr.y = bprm[gi].y;
        } else if (kind == 1) {
          const int a = gi - e1;
          qa.x = aidx[3 * a]; qa.y = aidx[3 * a + 1]; qa.z = aidx[3 * a + 2]; pr.x = aprm[a].x; pr.y = aprm[a].y;
        } else if (kind == 2) {
          const int k = gi - e2;
          qa = dq[k];
          memcpy(&pr.x, &ds[k], 4); memcpy(&pr.y, &ds[k + 1], 4);
        } else if (kind == 3) {
          const int k = gi - e3;
          qa.x = i14[k].x; qa.y = i14[k].y;
          pr.x = p14[k].x / p14[k].z; pr.y = p14[k].y / p14[k].z; pr.z = charge[qa.x] * charge[qa.y] / p14[k].w;
        } else {
          const int k = gi - e4;
          qa = iq[k];
          memcpy(&pr.x, &is[k], 4); memcpy(&pr.y, &is[k + 1], 4);
        }
        win_atoms.push_back(qa); win_prm.push_back(pr); win_slot.push_back(e.second);
      }
      win_kend[w] = make_int4(kend[0], kend[1], kend[2], kend[3]);
      if (hi[w] >= lo[w] && hi[w] - lo[w] + 1 <= kLHalo) { win_lo[w] = lo[w]; win_n[w] = hi[w] - lo[w] + 1; halo_max = std::max(halo_max, win_n[w]); }
    }
    if (win_slot.empty()) { win_atoms.push_back(make_int4(0, 0, 0, 0)); win_prm.push_back(make_float4(0.f, 0.f, 0.f, 0.f)); win_slot.push_back(make_ushort4(0xffff, 0xffff, 0xffff, 0xffff)); }
  }
  // atom-major lists of the dihedrals an atom sits in (dihedral << 2 | role): the dihedral-angle backward gathers by atom
  std::vector<int> phi_start(N + 1, 0), phi_items;
  {
    std::vector<int> cnt(N + 1, 0);
    for (int q = 0; q < d->ndihedrals; ++q)
      for (int role = 0; role < 4; ++role) ++cnt[d->dihedral_idx[4 * q + role] + 1];
    for (int a = 0; a < N; ++a) phi_start[a + 1] = phi_start[a] + cnt[a + 1];
    phi_items.assign(std::max(1, phi_start[N]), 0);
    std::vector<int> cur(phi_start.begin(), phi_start.end() - 1);
    for (int q = 0; q < d->ndihedrals; ++q)
      for (int role = 0; role < 4; ++role) phi_items[cur[d->dihedral_idx[4 * q + role]]++] = (q << 2) | role;
  }
  // rigid X-H bonds: groups of a heavy atom and its 1 .. 3 hydrogens (the rule of the on-chip path: a hydrogen is tied
  // through its shortest bond row)
  std::vector<int4> cgroup;
  std::vector<float4> cdist2, cweight;
  int ncons = 0, cmax = 0;
  {
    auto is_h = [&](int a) { return d->mass && d->mass[a] > 0.5f && d->mass[a] < 3.5f; };
    std::vector<int> partner(N, -1);
    std::vector<float> req(N, 0.0f);
    for (int b = 0; b < d->nbonds; ++b) {
      const int i = d->bond_idx[2 * b], j = d->bond_idx[2 * b + 1];
      const float r0 = d->bond_prm[2 * b + 1];
      for (int side = 0; side < 2; ++side) {
        const int hh = side ? j : i, other = side ? i : j;
        if (is_h(hh) && (partner[hh] < 0 || r0 < req[hh])) { partner[hh] = other; req[hh] = r0; }
      }
    }
    std::vector<std::vector<int>> kids(N);
    for (int a = 0; a < N && ncons >= 0; ++a) {
      if (partner[a] < 0) continue;
      const int pa = partner[a];
      if (is_h(pa) || kids[pa].size() >= 3 || !(req[a] > 0.0f)) { ncons = -1; break; }
      kids[pa].push_back(a);
      ncons += 1;
    }
    if (ncons > 0)
      for (int pa = 0; pa < N; ++pa) {
        const int n = (int)kids[pa].size();
        if (n == 0) continue;
        int hh[3] = {pa, pa, pa};
        float dd[3] = {1.0f, 1.0f, 1.0f}, ww[3] = {1.0f, 1.0f, 1.0f};
        for (int k = 0; k < n; ++k) { hh[k] = kids[pa][k]; dd[k] = req[hh[k]] * req[hh[k]]; ww[k] = 1.0f / d->mass[hh[k]]; }
        cgroup.push_back(make_int4(pa | (n << 24), hh[0], hh[1], hh[2]));
        cdist2.push_back(make_float4(dd[0], dd[1], dd[2], 0.0f));
        cweight.push_back(make_float4(1.0f / d->mass[pa], ww[0], ww[1], ww[2]));
        cmax = std::max(cmax, n);
      }
  }
  BlobBuilder B;
  const size_t o_cg = B.add(cgroup), o_cd = B.add(cdist2), o_cw = B.add(cweight);
  const size_t o_phs = B.add(phi_start), o_phit = B.add(phi_items);
  const size_t o_wst = B.add(win_start), o_wat = B.add(win_atoms), o_wpr = B.add(win_prm), o_wke = B.add(win_kend), o_wsl = B.add(win_slot), o_gst = B.add(gstart), o_wlo = B.add(win_lo), o_wn = B.add(win_n);
  const size_t o_type = B.add(atype), o_q = B.add(charge), o_m = B.add(mass), o_a = B.add(t12a), o_b = B.add(t6b), o_mask = B.add(mask),
               o_bi = B.add(bidx), o_bp = B.add(bprm), o_ai = B.add(aidx), o_ap = B.add(aprm), o_dq = B.add(dq), o_ds = B.add(ds),
               o_dt = B.add(dt), o_i14 = B.add(i14), o_p14 = B.add(p14), o_iq = B.add(iq), o_is = B.add(is), o_it = B.add(it), o_dcs = B.add(dcs), o_ics = B.add(ics), o_t64 = B.add(tab64);
  cudaDeviceProp prop;
  DeviceGuard guard(device);
  if (guard.err != cudaSuccess || cudaGetDeviceProperties(&prop, device) != cudaSuccess ||
      cudaMalloc(&topo->blob, B.bytes.size()) != cudaSuccess ||
      cudaMemcpy(topo->blob, B.bytes.data(), B.bytes.size(), cudaMemcpyHostToDevice) != cudaSuccess) {
    if (topo->blob) cudaFree(topo->blob);
    delete topo; fail(B200MD_ERR_CUDA, "device allocation of the packed topology failed"); return nullptr;
  }
  topo->num_sms = prop.multiProcessorCount;
  const char* base = (const char*)topo->blob;
  LargeTopo& L = topo->ldev;
  L.natoms = N; L.ntypes = NT; L.nwords = W;
  L.atom_type = (const int*)(base + o_type); L.charge = (const float*)(base + o_q); L.mass = (const float*)(base + o_m);
  L.tab12a = (const float*)(base + o_a); L.tab6b = (const float*)(base + o_b); L.mask = (const uint32_t*)(base + o_mask);
  L.tab64 = (const double2*)(base + o_t64);
  L.win_start = (const int*)(base + o_wst); L.win_atoms = (const int4*)(base + o_wat); L.win_prm = (const float4*)(base + o_wpr); L.win_kend = (const int4*)(base + o_wke); L.win_slot = (const ushort4*)(base + o_wsl);
  L.gstart = (const int*)(base + o_gst); L.nwin = nwin; L.win_slots_max = win_slots_max; L.halo_max = halo_max;
  L.win_lo = (const int*)(base + o_wlo); L.win_n = (const int*)(base + o_wn);
  L.phi_start = (const int*)(base + o_phs); L.phi_items = (const int*)(base + o_phit);
  L.ngroups = (int)cgroup.size(); L.ncons = ncons; L.cmax = cmax;
  topo->dev.ncons = ncons;  // (b200md_topology_num_constraints, the entry checks; the large-solute kernels read LargeTopo)
  L.cgroup = (const int4*)(base + o_cg); L.cdist2 = (const float4*)(base + o_cd); L.cweight = (const float4*)(base + o_cw);
  L.nblk = (W + kLBT - 1) / kLBT;  // pair rectangles over nblk blocks of bt <= kLBT tiles, as even as the tile count allows
  L.bt = (W + L.nblk - 1) / L.nblk;
  L.nbonds = d->nbonds; L.nangles = d->nangles; L.ndih = d->ndihedrals; L.n14 = d->n14; L.nimp = d->nimpropers;
  L.bond_idx = (const int2*)(base + o_bi); L.bond_prm = (const float2*)(base + o_bp);
  L.angle_idx = (const int*)(base + o_ai); L.angle_prm = (const float2*)(base + o_ap);
  L.dih_idx = (const int4*)(base + o_dq); L.dih_start = (const int*)(base + o_ds); L.dih_term = (const float4*)(base + o_dt);
  L.idx14 = (const int2*)(base + o_i14); L.prm14 = (const float4*)(base + o_p14);
  L.dih_cs = (const double2*)(base + o_dcs); L.imp_cs = (const double2*)(base + o_ics);
  L.imp_idx = (const int4*)(base + o_iq); L.imp_start = (const int*)(base + o_is); L.imp_term = (const float4*)(base + o_it);
  const int pair_bytes = large_pairs_dyn_bytes(L), bonded_bytes = large_bonded_dyn_bytes(L);
  if (pair_bytes > 200 * 1024 || bonded_bytes > 200 * 1024) {
    cudaFree(topo->blob); delete topo;
    fail(B200MD_ERR_UNSUPPORTED, "the type tables or the bonded slots of a 64-atom window exceed shared memory"); return nullptr;
  }
  if (pair_bytes > 32 * 1024) {
    cudaFuncSetAttribute(large_pairs_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, pair_bytes);
    cudaFuncSetAttribute(large_pairs_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, pair_bytes);
  }
  if (bonded_bytes > 32 * 1024) cudaFuncSetAttribute(large_bonded_forces_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bonded_bytes);
  {
    const int steps_bytes = large_steps_dyn_bytes(L);
    int coop = 0;
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, device);
    if (coop && steps_bytes <= 200 * 1024) {
      if (steps_bytes > 32 * 1024) {
        cudaFuncSetAttribute(large_steps_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, steps_bytes);
        cudaFuncSetAttribute(large_steps_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, steps_bytes);
      }
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&topo->occ_steps[0], large_steps_kernel<0>, kLThreads, steps_bytes) != cudaSuccess) topo->occ_steps[0] = 0;
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&topo->occ_steps[1], large_steps_kernel<1>, kLThreads, steps_bytes) != cudaSuccess) topo->occ_steps[1] = 0;
    }
    cudaGetLastError();
  }
  g_last_error.clear();
  return topo;
}

// Workspace of one large-solute call, carved from the topology's grow-only device buffer: the partial force rows, the tile
// boxes, the per-CTA energy partials, the per-replica energies / kinetic-energy partials / SMD record of the integrator.
struct LargeWork {
  cudaStream_t st;
  double* partial = nullptr; double* energies = nullptr; double* ekin = nullptr; double* smd = nullptr;
  float* bbox = nullptr;
  float* fpart = nullptr;  // [R][nblk + 1][natoms][3] partial force rows (large_pairs_kernel, large_bonded_forces_kernel)
  float* pos_old = nullptr;  // rigid bonds: x(t), the bond vectors SHAKE works along
  int nbonded_blocks = 0, npart = 0, nek = 0;
};

static int large_alloc(const b200md_topology* topo, int64_t R, bool integrator, bool smd, LargeWork& w, bool rigid = false) {
  const LargeTopo& L = topo->ldev;
  const long long items = (long long)L.nbonds + L.nangles + L.ndih + L.n14 + L.nimp;
  // bonded CTAs per replica: a function of the topology only (4 items per thread, at most 64 CTAs), so that the order of the
  // FP64 energy partials -- and with it every bit of a replica's result -- does not depend on the replica count of the call
  (void)R;
  w.nbonded_blocks = (int)std::max<long long>(1, std::min<long long>(64, (items + 4 * kLBondedThreads - 1) / (4 * kLBondedThreads)));
  w.npart = w.nbonded_blocks + L.nwords;
  w.nek = (L.natoms + 255) / 256;
  auto up = [](size_t x) { return (x + 255) & ~size_t(255); };
  const size_t b_f = up(sizeof(float) * 3 * (size_t)R * (L.nblk + 1) * L.natoms);
  const size_t b_p = up(sizeof(double) * (size_t)R * w.npart * B200MD_NUM_ENERGIES);
  const size_t b_e = integrator ? up(sizeof(double) * (size_t)R * B200MD_NUM_ENERGIES) : 0, b_k = integrator ? up(sizeof(double) * (size_t)R * w.nek) : 0;
  const size_t b_s = (integrator && smd) ? up(sizeof(double) * 2 * (size_t)R) : 0;
  const size_t b_b = up(sizeof(float) * 8 * (size_t)R * L.nwords);
  const size_t b_o = rigid ? up(sizeof(float) * 3 * (size_t)R * L.natoms) : 0;  // x(t) for SHAKE
  const size_t total = b_f + b_p + b_e + b_k + b_s + b_b + b_o;
  if (topo->ws != nullptr && topo->ws_stream != w.st) CUDA_TRY(cudaStreamSynchronize(topo->ws_stream));
  if (total > topo->ws_bytes) {
    if (topo->ws != nullptr) { CUDA_TRY(cudaStreamSynchronize(topo->ws_stream)); cudaFree(topo->ws); topo->ws = nullptr; topo->ws_bytes = 0; }
    CUDA_TRY(cudaMalloc(&topo->ws, total));
    topo->ws_bytes = total;
  }
  topo->ws_stream = w.st;
  char* base = (char*)topo->ws;
  w.fpart = (float*)base; base += b_f;
  w.partial = (double*)base; base += b_p;
  w.bbox = (float*)base; base += b_b;
  w.pos_old = rigid ? (float*)base : nullptr; base += b_o;
  if (integrator) {
    w.energies = (double*)base; base += b_e;
    w.ekin = (double*)base; base += b_k;
    if (smd) {
      w.smd = (double*)base;
      CUDA_TRY(cudaMemsetAsync(w.smd, 0, sizeof(double) * 2 * (size_t)R, w.st));
    }
  }
  return B200MD_OK;
}

// One Forces.compute over R replicas: tile boxes (cutoff mode), the force launch (partial rows), their sum into `forces`
// unless the caller takes it itself (the integrator), and on a report step the FP64 energies beside them on a second stream.
static int large_forces(const b200md_topology* topo, const ForceParams& P, int64_t R, const float* pos, const float* box,
                        float* forces, double* energies, uint32_t* status, double smd_center, bool with_energy, LargeWork& w,
                        bool bbox_ready = false, bool sum_rows = true) {
  const LargeTopo& L = topo->ldev;
  cudaStream_t st = w.st;
  const int gy = (int)std::min<int64_t>(R, 65535);
  static const bool overlap_on = !(getenv("B200MD_LARGE_OVERLAP") && atoi(getenv("B200MD_LARGE_OVERLAP")) == 0);
  if (overlap_on && topo->aux == nullptr) {
    CUDA_TRY(cudaStreamCreateWithFlags(&topo->aux, cudaStreamNonBlocking));
    CUDA_TRY(cudaEventCreateWithFlags(&topo->ev_fork, cudaEventDisableTiming));
    CUDA_TRY(cudaEventCreateWithFlags(&topo->ev_join, cudaEventDisableTiming));
    CUDA_TRY(cudaEventCreateWithFlags(&topo->ev_bonded, cudaEventDisableTiming));
  }
  const bool pairs_on = (P.terms & (B200MD_T_ELECTROSTATICS | B200MD_T_LJ | B200MD_T_REPULSION | B200MD_T_REPULSIONCG)) != 0;
  const float* bbox = nullptr;
  if (P.has_cutoff && pairs_on && L.nwords > 4) {  // tile bounding boxes for the cutoff skips (vacuum replicas use them)
    if (!bbox_ready) {  // (the integrator's fused half-kick kernel has written them already)
      large_bbox_kernel<<<dim3((L.nwords + 3) / 4, gy), 128, 0, st>>>(L, R, pos, w.bbox);
      g_launches.fetch_add(1);
    }
    bbox = w.bbox;
  }
  // The bonded windows and the pair rectangles write different slots of the partial rows and only read the positions: the
  // bonded launch (latency-bound, 64 registers) runs on a second stream beside the pair launch (packed-FP32-bound, 96
  // registers), and on a report step the FP64 energy kernels follow it there.  Fork / join by events, so the caller's
  // stream still orders the call as a whole.
  static const bool split_on = overlap_on && !(getenv("B200MD_LARGE_SPLIT") && atoi(getenv("B200MD_LARGE_SPLIT")) == 0);
  cudaStream_t sb = overlap_on ? topo->aux : st;
  // stream of the bonded force launch: the second one, except on a report step, where that stream carries the FP64 energy
  // kernels (the longer chain: 400 us against 233 us of force kernels for 256 x 1 040 atoms, plain) and the caller's stream both force kernels
  cudaStream_t sf = (split_on && !with_energy) ? topo->aux : st;
  const int npart = pairs_on ? w.npart : w.nbonded_blocks;
  if (overlap_on && (with_energy || (split_on && pairs_on))) {
    CUDA_TRY(cudaEventRecord(topo->ev_fork, st));
    CUDA_TRY(cudaStreamWaitEvent(sb, topo->ev_fork, 0));
  } else {
    sf = st;
  }
  large_bonded_forces_kernel<<<dim3(L.nwin, gy), kLBondedThreads, large_bonded_dyn_bytes(L), sf>>>(L, P, R, pos, box, w.fpart, status, P.smd_centers,
                                                                                                   smd_center, w.smd);
  g_launches.fetch_add(1);
  if (sf != st) CUDA_TRY(cudaEventRecord(topo->ev_bonded, sf));
  if (with_energy) {
    large_bonded_energy_kernel<<<dim3(w.nbonded_blocks, gy), kLBondedThreads, 0, sb>>>(L, P, R, pos, box, w.partial, npart,
                                                                                      P.smd_centers, smd_center);
    g_launches.fetch_add(1);
    if (pairs_on) {
      if (P.has_cutoff) large_pair_energy_kernel<1><<<dim3(L.nwords, gy), kLThreads, 0, sb>>>(L, P, R, pos, box, w.partial, npart, w.nbonded_blocks, bbox);
      else large_pair_energy_kernel<0><<<dim3(L.nwords, gy), kLThreads, 0, sb>>>(L, P, R, pos, box, w.partial, npart, w.nbonded_blocks, bbox);
      g_launches.fetch_add(1);
    }
    if (overlap_on) CUDA_TRY(cudaEventRecord(topo->ev_join, sb));
  }
  if (pairs_on) {
    const dim3 grid(L.nblk * (L.nblk + 1) / 2, gy);
    const int dyn_bytes = large_pairs_dyn_bytes(L);
    if (P.has_cutoff) large_pairs_kernel<1><<<grid, kLThreads, dyn_bytes, st>>>(L, P, R, pos, box, w.fpart, bbox);
    else large_pairs_kernel<0><<<grid, kLThreads, dyn_bytes, st>>>(L, P, R, pos, box, w.fpart, bbox);
    g_launches.fetch_add(1);
  }
  if (sf != st) CUDA_TRY(cudaStreamWaitEvent(st, topo->ev_bonded, 0));
  if (sum_rows) {
    const long long n = (long long)R * L.natoms * 3;
    large_sum_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(L, R, w.fpart, pairs_on ? 0 : L.nblk, forces);
    g_launches.fetch_add(1);
  }
  if (with_energy) {
    if (overlap_on) CUDA_TRY(cudaStreamWaitEvent(st, topo->ev_join, 0));  // join: the energy partials are complete
    const long long n = (long long)R * B200MD_NUM_ENERGIES;
    large_finish_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(R, w.partial, npart, energies);
    g_launches.fetch_add(1);
  }
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

static int large_forces_compute(const b200md_topology* topo, const ForceParams& P, int64_t R, const float* pos, const float* box,
                                float* forces, double* energies, uint32_t* status, void* stream) {
  LargeWork w;
  w.st = (cudaStream_t)stream;
  int rc = large_alloc(topo, R, false, false, w);
  if (rc) return rc;
  if (status) CUDA_TRY(cudaMemsetAsync(status, 0, sizeof(uint32_t) * (size_t)R, w.st));
  return large_forces(topo, P, R, pos, box, forces, energies, status, (double)P.smd_c0, energies != nullptr, w);
}

// Integrator.step(niter) for large solutes: the upstream sequence as separate launches per step (the state lives in HBM)
static int large_integrate(const b200md_topology* topo, const ForceParams& P, const StepParams& I, int64_t R, float* pos,
                           float* vel, float* forces, const float* box, double* observables, float* snapshots,
                           uint32_t* status, void* stream, bool rigid) {
  const LargeTopo& L = topo->ldev;
  LargeWork w;
  w.st = (cudaStream_t)stream;
  int rc = large_alloc(topo, R, true, P.smd_on != 0, w, rigid);
  if (rc) return rc;
  cudaStream_t st = w.st;
  if (status) CUDA_TRY(cudaMemsetAsync(status, 0, sizeof(uint32_t) * (size_t)R, st));
  const long long natom = (long long)R * L.natoms;
  const size_t frame = sizeof(float) * 3 * (size_t)natom;
  // Between two force evaluations the second half-kick of step s, the first half-kick of step s + 1 and the tile bounding
  // boxes are ONE launch (large_vv2_kernel with fuse_next) -- except where pos must stay x(t + dt): snapshot steps and the
  // last step of the call.
  const bool pairs_on = (P.terms & (B200MD_T_ELECTROSTATICS | B200MD_T_LJ | B200MD_T_REPULSION | B200MD_T_REPULSIONCG)) != 0;
  const bool want_bbox = P.has_cutoff && L.nwords > 4 && pairs_on;
  bool fused_prev = false;
  auto is_report = [&](int s) { return (I.energy_every > 0 && (s + 1) % I.energy_every == 0) || s == I.nsteps - 1; };
  auto is_snap = [&](int s) { return I.snapshot_every > 0 && snapshots != nullptr && (s + 1) % I.snapshot_every == 0; };
  // Few replicas: runs of steps between two report steps go to ONE cooperative launch (large_steps_kernel) while the units
  // of a step fill at most half of its resident grid (measured, cutoff mode: one 3 328-atom replica 30.5 -> 27.9 us per
  // step; 16 replicas 48.8 -> 60.1 us, 64 replicas 113 -> 153 us the other way: with more work per step the per-step
  // launches, each kernel at its own register budget, are faster).
  const int mode = P.has_cutoff ? 1 : 0;
  const long long units = (long long)((pairs_on ? L.nblk * (L.nblk + 1) / 2 : 0) + L.nwin) * R;
  const long long resident = (long long)topo->num_sms * topo->occ_steps[mode];
  bool persist = resident > 0 && 2 * units <= resident;
  if (const char* e = getenv("B200MD_LARGE_PERSISTENT")) persist = atoi(e) != 0 && resident > 0;  // (read per call: tests toggle it)
  // Rigid X-H bonds: SHAKE sits between the drift and the force evaluation, RATTLE behind the second half-kick, so the
  // half-kicks of neighbouring steps are not fused and every step is its own launches (first half-kick, SHAKE, forces,
  // second half-kick, RATTLE).
  if (rigid) persist = false;
  const unsigned gcons = rigid ? (unsigned)std::min<long long>(((long long)R * L.ngroups + 127) / 128, (long long)topo->num_sms * 32) : 0;
  for (int s = 0; s < I.nsteps; ++s) {
    const unsigned long long gstep = I.step0 + (unsigned long long)s;
    const bool report = is_report(s);
    if (!fused_prev) {
      large_vv1_kernel<<<(unsigned)((natom + 255) / 256), 256, 0, st>>>(L, I.dt, R, pos, vel, forces, rigid ? w.pos_old : nullptr);
      g_launches.fetch_add(1);
      if (rigid) {
        large_shake_kernel<<<gcons, 128, 0, st>>>(L, R, w.pos_old, pos, vel, 1.0f / I.dt, 1, status);
        g_launches.fetch_add(1);
      }
    }
    bool boxes_ready = fused_prev && want_bbox;
    if (persist && !report) {
      int e = s;
      while (e < I.nsteps && !is_report(e)) { ++e; if (is_snap(e - 1)) break; }  // steps [s, e): no report, a snapshot only after the last
      if (want_bbox && !boxes_ready) {
        large_bbox_kernel<<<dim3((L.nwords + 3) / 4, (unsigned)std::min<int64_t>(R, 65535)), 128, 0, st>>>(L, R, pos, w.bbox);
        g_launches.fetch_add(1);
        boxes_ready = true;
      }
      const long long chunks = (long long)((L.natoms + kLThreads - 1) / kLThreads) * R;
      const unsigned grid = (unsigned)std::max<long long>(1, std::min<long long>(resident, std::max<long long>(units, chunks)));
      float* bb = want_bbox ? w.bbox : nullptr;
      long long Rl = R;
      int s_begin = s, s_end = e, snaps_on = snapshots != nullptr ? 1 : 0, pon = pairs_on ? 1 : 0;
      LargeTopo Lc = L; ForceParams Pc = P; StepParams Ic = I;
      void* args[] = {&Lc, &Pc, &Ic, &Rl, &pos, &vel, &forces, &box, &w.fpart, &bb, &status, &w.smd, &s_begin, &s_end, &snaps_on, &pon};
      const void* fn = mode ? (const void*)large_steps_kernel<1> : (const void*)large_steps_kernel<0>;
      if (cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(kLThreads), args, (size_t)large_steps_dyn_bytes(L), st) == cudaSuccess) {
        g_launches.fetch_add(1);
        fused_prev = e < I.nsteps && !is_snap(e - 1);
        if (is_snap(e - 1)) {
          const long long slot = e / I.snapshot_every - 1;
          CUDA_TRY(cudaMemcpyAsync((char*)snapshots + slot * frame, pos, frame, cudaMemcpyDeviceToDevice, st));
        }
        s = e - 1;
        continue;
      }
      cudaGetLastError();  // the grid does not fit this context (MPS share, a capture ...): the per-step launches do the same work
      persist = false;
    }
    const double center = (double)P.smd_c0 + (double)P.smd_rate * (double)(gstep + 1);
    rc = large_forces(topo, P, R, pos, box, forces, w.energies, status, center, report, w, boxes_ready, false);
    if (rc) return rc;
    const bool snap = I.snapshot_every > 0 && snapshots != nullptr && (s + 1) % I.snapshot_every == 0;
    const bool fuse = s + 1 < I.nsteps && !snap && !rigid;
    large_vv2_kernel<<<dim3(w.nek, (unsigned)std::min<int64_t>(R, 65535)), 256, 0, st>>>(L, I, gstep, R, vel, forces, pos, (report && !rigid) ? w.ekin : nullptr,
                                                                                       status, fuse ? 1 : 0, (fuse && want_bbox) ? w.bbox : nullptr, w.fpart, pairs_on ? 0 : L.nblk);
    fused_prev = fuse;
    g_launches.fetch_add(1);
    if (rigid) {
      large_rattle_kernel<<<gcons, 128, 0, st>>>(L, R, pos, vel);
      g_launches.fetch_add(1);
      if (report) {  // the kinetic energy of the record is that of the constrained velocities
        large_ekin_kernel<<<dim3(w.nek, (unsigned)std::min<int64_t>(R, 65535)), 256, 0, st>>>(L, R, vel, w.ekin);
        g_launches.fetch_add(1);
      }
    }
    if (report && observables != nullptr) {
      const long long slot = I.energy_every > 0 ? s / I.energy_every : 0;
      large_observe_kernel<<<(unsigned)((R + 127) / 128), 128, 0, st>>>(R, w.energies, w.ekin, w.nek, w.smd,
                                                                        observables + slot * R * B200MD_NUM_OBSERVABLES);
      g_launches.fetch_add(1);
    }
    if (I.snapshot_every > 0 && snapshots != nullptr && (s + 1) % I.snapshot_every == 0) {
      const long long slot = (s + 1) / I.snapshot_every - 1;
      CUDA_TRY(cudaMemcpyAsync((char*)snapshots + slot * frame, pos, frame, cudaMemcpyDeviceToDevice, st));
    }
  }
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

extern "C" void b200md_topology_destroy(b200md_topology* topo) {
  if (!topo) return;
  DeviceGuard guard(topo->device);
  if (topo->blob) cudaFree(topo->blob);
  if (topo->ws) cudaFree(topo->ws);
  if (topo->aux) cudaStreamDestroy(topo->aux);
  if (topo->ev_fork) cudaEventDestroy(topo->ev_fork);
  if (topo->ev_join) cudaEventDestroy(topo->ev_join);
  if (topo->ev_bonded) cudaEventDestroy(topo->ev_bonded);
  delete topo;
}

extern "C" int b200md_constrain(const b200md_topology* topo, int64_t R, float* pos, float* vel, uint32_t* status,
                                void* stream) {
  if (!topo || !pos) return fail(B200MD_ERR_INVALID_ARGUMENT, "null argument");
  if (R < 0) return fail(B200MD_ERR_INVALID_ARGUMENT, "negative replica count");
  if (topo->dev.ncons < 0)
    return fail(B200MD_ERR_UNSUPPORTED, "rigid_bonds: a hydrogen is bonded to a hydrogen, or a heavy atom carries more than 3");
  if (R == 0 || topo->dev.ncons == 0) return B200MD_OK;
  B200MD_ON_DEVICE(topo->device);
  if (topo->large) {  // one thread per (replica, group): four SHAKE sweeps about the present bond vectors, then RATTLE
    const LargeTopo& L = topo->ldev;
    if (status) CUDA_TRY(cudaMemsetAsync(status, 0, sizeof(uint32_t) * (size_t)R, (cudaStream_t)stream));
    const unsigned gb = (unsigned)std::min<long long>(((long long)R * L.ngroups + 127) / 128, (long long)topo->num_sms * 32);
    large_shake_kernel<<<gb, 128, 0, (cudaStream_t)stream>>>(L, R, pos, pos, nullptr, 0.0f, 4, status);
    if (vel) large_rattle_kernel<<<gb, 128, 0, (cudaStream_t)stream>>>(L, R, pos, vel);
    g_launches.fetch_add(vel ? 2 : 1);
    CUDA_TRY(cudaGetLastError());
    return B200MD_OK;
  }
  const int grid = (int)std::min<long long>(R, (long long)topo->num_sms * 8);
  constrain_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(topo->dev, R, pos, vel, status);
  g_launches.fetch_add(1);
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

extern "C" int b200md_topology_ndihedrals(const b200md_topology* topo) { return topo ? (topo->large ? topo->ldev.ndih : topo->dev.nphi) : -1; }

extern "C" int b200md_dihedral_angles(const b200md_topology* topo, int64_t R, const float* pos, const float* box,
                                      float* phi, void* stream) {
  if (!topo || !pos || !phi) return fail(B200MD_ERR_INVALID_ARGUMENT, "null argument");
  if (R < 0) return fail(B200MD_ERR_INVALID_ARGUMENT, "negative replica count");
  if (topo->large) {
    if (R == 0 || topo->ldev.ndih == 0) return B200MD_OK;
    B200MD_ON_DEVICE(topo->device);
    const long long n = (long long)R * topo->ldev.ndih;
    large_dihedral_angles_kernel<<<(unsigned)std::min<long long>((n + 255) / 256, (long long)topo->num_sms * 32), 256, 0, (cudaStream_t)stream>>>(topo->ldev, R, pos, box, phi);
    g_launches.fetch_add(1);
    CUDA_TRY(cudaGetLastError());
    return B200MD_OK;
  }
  if (R == 0 || topo->dev.nphi == 0) return B200MD_OK;
  B200MD_ON_DEVICE(topo->device);
  const int grid = (int)std::min<long long>(R, (long long)topo->num_sms * 16);
  dihedral_angles_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(topo->dev, R, pos, box, phi);
  g_launches.fetch_add(1);
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

extern "C" int b200md_dihedral_angles_backward(const b200md_topology* topo, int64_t R, const float* pos, const float* box,
                                               const float* grad_phi, float* grad_pos, void* stream) {
  if (!topo || !pos || !grad_phi || !grad_pos) return fail(B200MD_ERR_INVALID_ARGUMENT, "null argument");
  if (R < 0) return fail(B200MD_ERR_INVALID_ARGUMENT, "negative replica count");
  if (topo->large) {
    if (R == 0) return B200MD_OK;
    B200MD_ON_DEVICE(topo->device);
    const long long n = (long long)R * topo->ldev.natoms;
    large_dihedral_angles_backward_kernel<<<(unsigned)std::min<long long>((n + 255) / 256, (long long)topo->num_sms * 32), 256, 0, (cudaStream_t)stream>>>(
        topo->ldev, R, pos, box, grad_phi, grad_pos);
    g_launches.fetch_add(1);
    CUDA_TRY(cudaGetLastError());
    return B200MD_OK;
  }
  if (topo->dev.nphi > 512) return fail(B200MD_ERR_UNSUPPORTED, "dihedral-angle backward holds at most 512 dihedrals per replica");
  if (R == 0) return B200MD_OK;
  B200MD_ON_DEVICE(topo->device);
  const int grid = (int)std::min<long long>(R, (long long)topo->num_sms * 16);
  dihedral_angles_backward_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(topo->dev, R, pos, box, grad_phi, grad_pos);
  g_launches.fetch_add(1);
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

extern "C" int b200md_topology_num_constraints(const b200md_topology* topo) { return topo ? topo->dev.ncons : -1; }

extern "C" int b200md_topology_exclusion_mask(const b200md_topology* topo, uint32_t* out) {
  if (!topo || !out) return fail(B200MD_ERR_INVALID_ARGUMENT, "null argument");
  memcpy(out, topo->mask_rows.data(), topo->mask_rows.size() * sizeof(uint32_t));
  return B200MD_OK;
}

extern "C" int b200md_topology_natoms(const b200md_topology* topo) { return topo ? topo->natoms : -1; }
extern "C" int b200md_topology_mask_words(const b200md_topology* topo) { return topo ? topo->mask_words : -1; }

// ---- launch helpers ---------------------------------------------------------------------------------
static int make_force_params(const b200md_topology* topo, const b200md_force_opts* o, const b200md_smd_opts* smd,
                             bool integrator, ForceParams* P) {
  if (!o) return fail(B200MD_ERR_INVALID_ARGUMENT, "force options are required");
  // forces.py:258-316 evaluates "lj", "repulsion" (A r^-12, :416-431) and "repulsioncg" (B r^-6, :434-448) over the
  // same pair list; this ABI has ONE pair-table slot and reports its energy in B200MD_E_LJ, so at most one of the
  // three may be enabled per call
  const int nfamily = !!(o->terms & B200MD_T_LJ) + !!(o->terms & B200MD_T_REPULSION) + !!(o->terms & B200MD_T_REPULSIONCG);
  if (nfamily > 1)
    return fail(B200MD_ERR_UNSUPPORTED, "at most one of lj / repulsion / repulsioncg per call (they share the pair-table slot)");
  memset(P, 0, sizeof(*P));
  P->terms = o->terms;
  P->has_cutoff = o->cutoff > 0.0f;
  P->rc = o->cutoff;
  P->rc2 = o->cutoff * o->cutoff;
  const bool sw = P->has_cutoff && o->switch_dist > 0.0f && (o->terms & B200MD_T_LJ);  // forces.py:402 (evaluate_LJ only)
  P->rs = sw ? o->switch_dist : INFINITY;
  P->inv_sw = sw ? 1.0f / (o->cutoff - o->switch_dist) : 0.0f;
  P->rc_d = (double)o->cutoff;
  P->rs_d = sw ? (double)o->switch_dist : (double)INFINITY;
  P->inv_sw_d = sw ? 1.0 / ((double)o->cutoff - (double)o->switch_dist) : 0.0;
  if (o->rfa) {
    if (!P->has_cutoff) return fail(B200MD_ERR_INVALID_ARGUMENT, "rfa needs a cutoff");
    const double eps = o->solvent_dielectric, rc = o->cutoff, denom = 2.0 * eps + 1.0;  // forces.py:466-468
    P->krf_d = (1.0 / (rc * rc * rc)) * (eps - 1.0) / denom;
    P->crf_d = (1.0 / rc) * (3.0 * eps) / denom;
    P->krf = (float)P->krf_d;
    P->crf = (float)P->crf_d;
  }
  P->lj_on = (o->terms & B200MD_T_LJ) ? 1.0f : 0.0f;
  P->tab_a = (o->terms & (B200MD_T_LJ | B200MD_T_REPULSION)) ? 1.0f : 0.0f;
  P->tab_b = (o->terms & B200MD_T_LJ) ? -1.0f : ((o->terms & B200MD_T_REPULSIONCG) ? 1.0f : 0.0f);
  P->el_on = (o->terms & B200MD_T_ELECTROSTATICS) ? 1.0f : 0.0f;
  if (smd && smd->enabled) {
    if (smd->atom1 < 0 || smd->atom1 >= topo->natoms || smd->atom2 < 0 || smd->atom2 >= topo->natoms ||
        smd->atom1 == smd->atom2)
      return fail(B200MD_ERR_INVALID_ARGUMENT, "SMD atoms out of range");
    P->smd_on = 1; P->smd_a1 = smd->atom1; P->smd_a2 = smd->atom2;
    P->smd_fk = smd->fk; P->smd_c0 = smd->center0; P->smd_rate = smd->center_rate;
    P->smd_centers = integrator ? nullptr : smd->centers;
  }
  return B200MD_OK;
}

// DevTopo copy with the per-call term switches applied: a disabled bonded term gets a zero count, so its
// chunks return immediately; its slots are zeroed once at staging and never written, so the gather adds 0.
static DevTopo apply_terms(const DevTopo& in, uint32_t terms) {
  DevTopo T = in;
  const bool p14 = (terms & B200MD_T_14) && (terms & (B200MD_T_LJ | B200MD_T_ELECTROSTATICS));
  if (!(terms & B200MD_T_BONDS)) T.nbond2 = 0;
  if (!(terms & B200MD_T_ANGLES)) T.nangle2 = 0;
  T.dih_on = (terms & B200MD_T_DIHEDRALS) ? 1 : 0;
  T.p14_on = p14 ? 1 : 0;
  if (!T.dih_on && !T.p14_on) T.ndih2 = 0;
  if (!p14) T.n14_2 = 0;
  // general list = [special dihedrals | impropers]
  T.gen_begin = T.dih_on ? 0 : T.nspec;
  T.gen_end = (terms & B200MD_T_IMPROPERS) ? T.ngen : T.nspec;
  return T;
}

extern "C" int b200md_forces_compute(const b200md_topology* topo, const b200md_force_opts* opts,
                                     const b200md_smd_opts* smd, int64_t R, const float* pos, const float* box,
                                     float* forces, double* energies, uint32_t* status, void* stream) {
  if (!topo || !pos || !forces) return fail(B200MD_ERR_INVALID_ARGUMENT, "null argument");
  if (R < 0) return fail(B200MD_ERR_INVALID_ARGUMENT, "negative replica count");
  if (R == 0) return B200MD_OK;
  ForceParams P;
  int rc = make_force_params(topo, opts, smd, false, &P);
  if (rc) return rc;
  if (topo->large) {
    B200MD_ON_DEVICE(topo->device);
    return large_forces_compute(topo, P, R, pos, box, forces, energies, status, stream);
  }
  DevTopo T = apply_terms(topo->dev, opts->terms);
  B200MD_ON_DEVICE(topo->device);
  const int mode = P.has_cutoff ? 1 : 0;
  const long long cap = (long long)topo->num_sms * topo->occ_forces[mode];
  const int grid = (int)std::min<long long>(R, cap);
  cudaStream_t st = (cudaStream_t)stream;
  if (mode == 0)
    forces_kernel<0><<<grid, kThreads, T.smem_bytes, st>>>(T, P, R, pos, box, forces, status);
  else
    forces_kernel<1><<<grid, kThreads, T.smem_bytes, st>>>(T, P, R, pos, box, forces, status);
  g_launches.fetch_add(1);
  if (energies != nullptr) {  // FP64 energies: a second kernel on the same stream (energy_kernel)
    const int egrid = (int)std::min<long long>(R, (long long)topo->num_sms * topo->occ_energy[mode]);
    if (mode == 0) energy_kernel<0><<<egrid, kThreads, 0, st>>>(T, P, R, pos, box, R, energies, B200MD_NUM_ENERGIES, B200MD_NUM_ENERGIES);
    else energy_kernel<1><<<egrid, kThreads, 0, st>>>(T, P, R, pos, box, R, energies, B200MD_NUM_ENERGIES, B200MD_NUM_ENERGIES);
    g_launches.fetch_add(1);
  }
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

extern "C" int b200md_integrate(const b200md_topology* topo, const b200md_force_opts* fopts,
                                const b200md_smd_opts* smd, const b200md_integrate_opts* io, int64_t R, float* pos,
                                float* vel, float* forces, const float* box, double* observables, float* snapshots,
                                uint32_t* status, void* stream) {
  if (!topo || !io || !pos || !vel || !forces) return fail(B200MD_ERR_INVALID_ARGUMENT, "null argument");
  if (R < 0 || io->nsteps < 0) return fail(B200MD_ERR_INVALID_ARGUMENT, "negative replica or step count");
  if (R == 0 || io->nsteps == 0) return B200MD_OK;
  if (io->rigid_bonds != 0 && topo->dev.ncons < 0)
    return fail(B200MD_ERR_UNSUPPORTED, "rigid_bonds: a hydrogen is bonded to a hydrogen, or a heavy atom carries more than 3");
  ForceParams P;
  int rc = make_force_params(topo, fopts, smd, true, &P);
  if (rc) return rc;
  DevTopo T = apply_terms(topo->dev, fopts->terms);
  StepParams I{};
  I.dt = io->dt; I.gamma = io->gamma; I.temperature = io->temperature;
  I.seed = io->seed; I.step0 = io->step0; I.replica0 = io->replica0;
  I.nsteps = io->nsteps; I.energy_every = io->energy_every; I.snapshot_every = io->snapshot_every;
  B200MD_ON_DEVICE(topo->device);
  if (topo->large)
    return large_integrate(topo, P, I, R, pos, vel, forces, box, observables, snapshots, status, stream,
                           io->rigid_bonds != 0 && topo->ldev.ncons > 0);
  const int mode = P.has_cutoff ? 1 : 0;
  cudaStream_t st = (cudaStream_t)stream;
  const bool rigid = io->rigid_bonds != 0 && T.ncons > 0;
  // One wave of persistent CTAs draws (slice, replica) units from a ticket counter (integrate_kernel).  Slice length:
  // about 40 rounds of units over the resident CTAs, so that the tail of a launch is a few percent of it (1 024 replicas x
  // 1 000 steps: 8.45 / 8.50 / 8.44 / 8.33 / 8.29e7 replica-steps/s at slices of 16 / 32 / 64 / 86 / 128 steps), within
  // [8, 128] steps (parking a replica costs ~3 us against ~8 us per step: 8.73 / 8.85 / 8.91 / 8.93e7 replica-steps/s at
  // 16 / 32 / 64 / 128 for 8 192 replicas x 1 000 steps); B200MD_SLICE overrides it for experiments.
  const long long cap = (long long)topo->num_sms * topo->occ_integrate[mode];
  const int grid = (int)std::min<long long>(R, cap);
  long long slice = (long long)io->nsteps * R / (cap * 40);
  slice = std::max<long long>(8, std::min<long long>(128, slice));
  if (const char* e = getenv("B200MD_SLICE")) { const long long v = atoll(e); if (v > 0) slice = v; }
  slice = std::min<long long>(slice, io->nsteps);
  I.slice_len = (int)slice;
  I.nslices = (int)((io->nsteps + slice - 1) / slice);
  long long nrec = 1;
  {  // ticket counter, per-replica slice counts and (SMD) carried work: grow-only workspace of the handle, zeroed per launch
    auto up = [](size_t x) { return (x + 255) & ~size_t(255); };
    const size_t b_t = 256, b_d = I.nslices > 1 ? up(sizeof(unsigned int) * (size_t)R) : 0;
    const size_t b_w = (I.nslices > 1 && P.smd_on) ? up(sizeof(double) * (size_t)R) : 0;
    // deferred energies (energy_kernel behind the launch): positions of every record, unless that buffer would pass
    // B200MD_DEFER_ENERGY_MB (default 8192; 0: always evaluate the energies inside the fused kernel)
    nrec = io->energy_every > 0 ? (io->nsteps + io->energy_every - 1) / io->energy_every : 1;
    size_t b_r = up(sizeof(float) * 3 * (size_t)T.natoms * (size_t)R * (size_t)nrec);
    long long budget_mb = 8192;
    if (const char* e = getenv("B200MD_DEFER_ENERGY_MB")) budget_mb = atoll(e);
    if (observables == nullptr || (long long)(b_r >> 20) >= budget_mb) b_r = 0;
    const size_t total = b_t + b_d + b_w + b_r;
    if (topo->ws != nullptr && topo->ws_stream != st) CUDA_TRY(cudaStreamSynchronize(topo->ws_stream));
    if (total > topo->ws_bytes) {
      if (topo->ws != nullptr) { CUDA_TRY(cudaStreamSynchronize(topo->ws_stream)); cudaFree(topo->ws); topo->ws = nullptr; topo->ws_bytes = 0; }
      CUDA_TRY(cudaMalloc(&topo->ws, total));
      topo->ws_bytes = total;
    }
    topo->ws_stream = st;
    char* wbase = (char*)topo->ws;
    I.ticket = (unsigned long long*)wbase;
    I.slice_done = b_d ? (unsigned int*)(wbase + b_t) : nullptr;
    I.smd_work = b_w ? (double*)(wbase + b_t + b_d) : nullptr;
    I.report_pos = b_r ? (float*)(wbase + b_t + b_d + b_w) : nullptr;
    CUDA_TRY(cudaMemsetAsync(wbase, 0, b_t + b_d, st));
  }
#define B200MD_LAUNCH_INTEGRATE(M, RG) \
  integrate_kernel<M, RG><<<grid, kThreads, T.smem_bytes_fixed, st>>>(T, P, I, R, pos, vel, forces, box, observables, snapshots, status)
  if (mode == 0) { if (rigid) B200MD_LAUNCH_INTEGRATE(0, true); else B200MD_LAUNCH_INTEGRATE(0, false); }
  else { if (rigid) B200MD_LAUNCH_INTEGRATE(1, true); else B200MD_LAUNCH_INTEGRATE(1, false); }
#undef B200MD_LAUNCH_INTEGRATE
  g_launches.fetch_add(1);
  if (I.report_pos != nullptr) {  // the force-field energies of every record, FP64, in one launch behind the steps
    const long long nev = nrec * R;
    const int egrid = (int)std::min<long long>(nev, (long long)topo->num_sms * topo->occ_energy[mode]);
    if (mode == 0) energy_kernel<0><<<egrid, kThreads, 0, st>>>(T, P, nev, I.report_pos, box, R, observables, B200MD_NUM_OBSERVABLES, B200MD_E_EXTERNAL);
    else energy_kernel<1><<<egrid, kThreads, 0, st>>>(T, P, nev, I.report_pos, box, R, observables, B200MD_NUM_OBSERVABLES, B200MD_E_EXTERNAL);
    g_launches.fetch_add(1);
  }
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

// ---- b200md_integrate without a single allocation -------------------------------------------------------------------
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  b200md_integrate takes the scratch of
// the time-sliced scheduler (ticket counter, per-replica slice counts, carried SMD work) from a grow-only cudaMalloc'd
// workspace of the topology handle; this variant keeps it in a file-scope device array of fixed size:
//   [ticket: 256 B][slice_done: 4 B x B200MD_NOALLOC_MAX_REPLICAS][smd_work: 8 B x B200MD_NOALLOC_MAX_REPLICAS] = 786 688 B,
// resident on every GPU that loads the library until the process exits.  In host code `g_noalloc_scratch` is a host-side
// stand-in, NOT a device address: the address the kernel gets comes from cudaGetSymbolAddress, resolved once per device
// with that device current.  Every call shares the array and finds the previous call's contents in it: the ticket and the
// slice counts are cleared (cudaMemsetAsync on the caller's stream, not an allocation), the SMD work of a replica is
// written by its first slice before a later one reads it.  Calls that OVERLAP on the GPU (two streams, two host threads)
// race on the array -- serialise them.  Energies of report steps are evaluated inside the fused kernel (the deferred
// energy_kernel path of b200md_integrate needs a buffer that grows with replicas x records).
constexpr size_t kNoAllocTicketBytes = 256;
constexpr size_t kNoAllocDoneBytes = sizeof(unsigned int) * (size_t)B200MD_NOALLOC_MAX_REPLICAS;
constexpr size_t kNoAllocWorkBytes = sizeof(double) * (size_t)B200MD_NOALLOC_MAX_REPLICAS;
constexpr size_t kNoAllocBytes = kNoAllocTicketBytes + kNoAllocDoneBytes + kNoAllocWorkBytes;
static_assert((kNoAllocTicketBytes + kNoAllocDoneBytes) % 8 == 0, "smd_work must stay 8-byte aligned");
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];
constexpr int kNoAllocMaxDevices = 64;
static std::mutex g_noalloc_mutex;
static unsigned char* g_noalloc_ptr[kNoAllocMaxDevices] = {nullptr};

extern "C" int b200md_integrate_noalloc(const b200md_topology* topo, const b200md_force_opts* fopts,
                                        const b200md_smd_opts* smd, const b200md_integrate_opts* io, int64_t R, float* pos,
                                        float* vel, float* forces, const float* box, double* observables, float* snapshots,
                                        uint32_t* status, void* stream) {
  if (!topo || !io || !pos || !vel || !forces) return fail(B200MD_ERR_INVALID_ARGUMENT, "null argument");
  if (R < 0 || io->nsteps < 0) return fail(B200MD_ERR_INVALID_ARGUMENT, "negative replica or step count");
  if (R > (int64_t)B200MD_NOALLOC_MAX_REPLICAS)
    return fail(B200MD_ERR_UNSUPPORTED, "b200md_integrate_noalloc: nreplicas = " + std::to_string((long long)R) +
                                            " exceeds B200MD_NOALLOC_MAX_REPLICAS = " + std::to_string(B200MD_NOALLOC_MAX_REPLICAS) +
                                            " (the fixed device scratch array); split the batch or use b200md_integrate");
  if (topo->large)
    return fail(B200MD_ERR_UNSUPPORTED, "b200md_integrate_noalloc: solutes above 128 atoms keep their state in an allocated workspace; use b200md_integrate");
  if (topo->device < 0 || topo->device >= kNoAllocMaxDevices)
    return fail(B200MD_ERR_UNSUPPORTED, "b200md_integrate_noalloc: device index beyond 63");
  if (R == 0 || io->nsteps == 0) return B200MD_OK;
  if (io->rigid_bonds != 0 && topo->dev.ncons < 0)
    return fail(B200MD_ERR_UNSUPPORTED, "rigid_bonds: a hydrogen is bonded to a hydrogen, or a heavy atom carries more than 3");
  ForceParams P;
  int rc = make_force_params(topo, fopts, smd, true, &P);
  if (rc) return rc;
  DevTopo T = apply_terms(topo->dev, fopts->terms);
  StepParams I{};
  I.dt = io->dt; I.gamma = io->gamma; I.temperature = io->temperature;
  I.seed = io->seed; I.step0 = io->step0; I.replica0 = io->replica0;
  I.nsteps = io->nsteps; I.energy_every = io->energy_every; I.snapshot_every = io->snapshot_every;
  B200MD_ON_DEVICE(topo->device);
  unsigned char* scratch = nullptr;
  {  // set-up, once per device, with that device current: the DEVICE address of the scratch array
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (g_noalloc_ptr[topo->device] == nullptr) {
      void* p = nullptr;
      const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
      if (e != cudaSuccess || p == nullptr)
        return fail(B200MD_ERR_CUDA, std::string("cudaGetSymbolAddress(g_noalloc_scratch): ") + cudaGetErrorString(e));
      g_noalloc_ptr[topo->device] = (unsigned char*)p;
    }
    scratch = g_noalloc_ptr[topo->device];
  }
  const int mode = P.has_cutoff ? 1 : 0;
  cudaStream_t st = (cudaStream_t)stream;
  const bool rigid = io->rigid_bonds != 0 && T.ncons > 0;
  const long long cap = (long long)topo->num_sms * topo->occ_integrate[mode];
  const int grid = (int)std::min<long long>(R, cap);
  long long slice = (long long)io->nsteps * R / (cap * 40);  // as b200md_integrate
  slice = std::max<long long>(8, std::min<long long>(128, slice));
  slice = std::min<long long>(slice, io->nsteps);
  I.slice_len = (int)slice;
  I.nslices = (int)((io->nsteps + slice - 1) / slice);
  I.ticket = (unsigned long long*)scratch;
  I.slice_done = (unsigned int*)(scratch + kNoAllocTicketBytes);
  I.smd_work = (double*)(scratch + kNoAllocTicketBytes + kNoAllocDoneBytes);
  I.report_pos = nullptr;  // energies inside the fused kernel
  // the array keeps the last call's contents: the ticket and the slice counts of this call's replicas must start at zero
  CUDA_TRY(cudaMemsetAsync(scratch, 0, kNoAllocTicketBytes + sizeof(unsigned int) * (size_t)R, st));
#define B200MD_LAUNCH_INTEGRATE(M, RG) \
  integrate_kernel<M, RG><<<grid, kThreads, T.smem_bytes_fixed, st>>>(T, P, I, R, pos, vel, forces, box, observables, snapshots, status)
  if (mode == 0) { if (rigid) B200MD_LAUNCH_INTEGRATE(0, true); else B200MD_LAUNCH_INTEGRATE(0, false); }
  else { if (rigid) B200MD_LAUNCH_INTEGRATE(1, true); else B200MD_LAUNCH_INTEGRATE(1, false); }
#undef B200MD_LAUNCH_INTEGRATE
  g_launches.fetch_add(1);
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

extern "C" int b200md_forces_compute_host(const b200md_topology* topo, const b200md_force_opts* opts,
                                          const b200md_smd_opts* smd, int64_t R, const float* pos_host,
                                          const float* box_host, float* forces_host, double* energies_host,
                                          uint32_t* status_host) {
  if (!topo || !pos_host || !forces_host || !energies_host) return fail(B200MD_ERR_INVALID_ARGUMENT, "null argument");
  if (R <= 0) return R == 0 ? B200MD_OK : fail(B200MD_ERR_INVALID_ARGUMENT, "negative replica count");
  if (smd && smd->enabled && smd->centers) return fail(B200MD_ERR_UNSUPPORTED, "per-replica SMD centres need device memory");
  B200MD_ON_DEVICE(topo->device);
  const size_t n3 = (size_t)R * topo->natoms * 3;
  float *dpos = nullptr, *dbox = nullptr, *dfrc = nullptr;
  double* dene = nullptr;
  uint32_t* dst = nullptr;
  int rc = B200MD_OK;
  cudaError_t e = cudaSuccess;
  auto ok = [&](cudaError_t err) { if (err != cudaSuccess && e == cudaSuccess) e = err; return err == cudaSuccess; };
  ok(cudaMalloc(&dpos, n3 * 4)) && ok(cudaMalloc(&dfrc, n3 * 4)) && ok(cudaMalloc(&dene, (size_t)R * 8 * 8)) &&
      ok(cudaMalloc(&dst, (size_t)R * 4));
  if (e == cudaSuccess && box_host) ok(cudaMalloc(&dbox, (size_t)R * 9 * 4)) && ok(cudaMemcpy(dbox, box_host, (size_t)R * 9 * 4, cudaMemcpyHostToDevice));
  if (e == cudaSuccess) ok(cudaMemcpy(dpos, pos_host, n3 * 4, cudaMemcpyHostToDevice));
  if (e == cudaSuccess) {
    rc = b200md_forces_compute(topo, opts, smd, R, dpos, dbox, dfrc, dene, dst, nullptr);
    if (rc == B200MD_OK) {
      ok(cudaMemcpy(forces_host, dfrc, n3 * 4, cudaMemcpyDeviceToHost));
      ok(cudaMemcpy(energies_host, dene, (size_t)R * 8 * 8, cudaMemcpyDeviceToHost));
      std::vector<uint32_t> st((size_t)R);
      ok(cudaMemcpy(st.data(), dst, (size_t)R * 4, cudaMemcpyDeviceToHost));
      if (status_host) memcpy(status_host, st.data(), (size_t)R * 4);
      if (e == cudaSuccess)
        for (uint32_t f : st) {
          if (f & B200MD_FLAG_NAN_POS) { rc = fail(B200MD_ERR_NAN_COORDINATES, "Found NaN coordinates."); break; }
        }
    }
  }
  cudaFree(dpos); cudaFree(dbox); cudaFree(dfrc); cudaFree(dene); cudaFree(dst);
  if (e != cudaSuccess) return fail(B200MD_ERR_CUDA, cudaGetErrorString(e));
  return rc;
}

extern "C" int b200md_fire_update(int device, int64_t R, int32_t natoms, float* pos, float* vel, const float* forces,
                                  float* state, const b200md_fire_opts* opts, void* stream) {
  if (!pos || !vel || !forces || !state || !opts) return fail(B200MD_ERR_INVALID_ARGUMENT, "null argument");
  if (R < 0 || natoms <= 0 || (long long)natoms * 3 > 0x7fffffffll) return fail(B200MD_ERR_INVALID_ARGUMENT, "bad replica or atom count");
  if (R == 0) return B200MD_OK;
  B200MD_ON_DEVICE(device);
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  const int grid = (int)std::min<long long>(R, (long long)sms * 8);
  fire_update_kernel<<<grid, kFireThreads, 0, (cudaStream_t)stream>>>(R, natoms * 3, pos, vel, forces, (float4*)state, *opts);
  g_launches.fetch_add(1);
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

extern "C" double b200md_fp32_peak_tflops(int device, int iters) {
  DeviceGuard guard(device);
  if (guard.err != cudaSuccess) { fail(B200MD_ERR_CUDA, "cudaSetDevice failed"); return -1.0; }
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, device);
  const int blocks = prop.multiProcessorCount * 8, threads = 256;
  float* out = nullptr;
  if (cudaMalloc(&out, (size_t)blocks * threads * sizeof(float)) != cudaSuccess) return -1.0;
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  fma_peak_kernel<<<blocks, threads>>>(out, 64, 1.000001f, 1e-7f);  // warm-up
  double best = -1.0;
  for (int rep = 0; rep < 3; ++rep) {
    cudaEventRecord(a);
    fma_peak_kernel<<<blocks, threads>>>(out, iters, 1.000001f, 1e-7f);
    cudaEventRecord(b);
    if (cudaEventSynchronize(b) != cudaSuccess) { best = -1.0; break; }
    float ms = 0.f;
    cudaEventElapsedTime(&ms, a, b);
    const double flops = 2.0 * 64.0 * (double)iters * blocks * threads;
    best = std::max(best, flops / (ms * 1e-3) * 1e-12);
    g_launches.fetch_add(1);
  }
  cudaEventDestroy(a); cudaEventDestroy(b);
  cudaFree(out);
  return best;
}

extern "C" int64_t b200md_launch_count(void) { return g_launches.load(); }
extern "C" const char* b200md_last_error(void) { return g_last_error.c_str(); }
extern "C" const char* b200md_version(void) { return "b200md 0.1 (sm_100a)"; }
