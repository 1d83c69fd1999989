// b200md.cu -- kernels + C-ABI (include/b200md.h) of the B200-native TorchMD force-field/integrator path.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -shared -Xcompiler -fPIC
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <string>
#include <vector>

#include "b200md_forces.cuh"

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
      S.ta[t * kTabStride + j] = v.x * P.lj_on;
      S.tb[t * kTabStride + j] = -(v.y * P.lj_on);
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
  if (tid < 8) S.smd[tid] = 0.0f;
  // slots of disabled / dropped terms are never written: keep them at zero for the gather
  if (FIXED) {
    for (int k = tid; k < 3 * kMaxAtoms; k += kThreads) S.fbi[k] = 0;
  } else {
    for (int k = tid; k < 3 * T.nslots; k += kThreads) S.fb[k] = 0.0f;
  }
  // chunks without a tile task (no non-bonded term enabled) must read as zero in the gather
  for (int k = tid; k < 3 * kMaxAtoms; k += kThreads) S.fj[k] = 0.0f;
  for (int k = tid; k < kWarps * 3 * kMaxAtoms; k += kThreads) S.fi[k] = 0.0f;  // i-side partials: armed here, re-armed by the gather
}

// Forces.compute (forces.py:82-344) for a batch of replicas: one CTA per replica (grid-stride).
template <int MODE>
__global__ void __launch_bounds__(kThreads, 4)
forces_kernel(const DevTopo T, const ForceParams P, const long long R, const float* __restrict__ pos,
              const float* __restrict__ box, float* __restrict__ forces, double* __restrict__ energies,
              uint32_t* __restrict__ status) {
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
  for (long long r = blockIdx.x; r < R; r += gridDim.x) {
    __syncthreads();  // previous replica fully consumed (and tables staged)
    uint32_t flag = 0;
    if (tid < N) {
      const float* p = pos + (r * N + tid) * 3;
      const float x = __ldg(p), y = __ldg(p + 1), z = __ldg(p + 2);
      if (x != x || y != y || z != z) flag |= B200MD_FLAG_NAN_POS;
      S.pos[tid] = make_float4(x, y, z, q);
      S.sx[tid] = x; S.sx[kMaxAtoms + tid] = y; S.sx[2 * kMaxAtoms + tid] = z; S.sx[3 * kMaxAtoms + tid] = q;
    }
    if (box != nullptr && tid < 3 && __ldg(box + r * 9 + tid * 4) != 0.0f) flag |= B200MD_FLAG_PBC_UNSUPPORTED;
    if (status != nullptr) {
      if (tid == 0) status[r] = 0;
    }
    __syncthreads();
    if (flag && status != nullptr) atomicOr(status + r, flag);
    Energies E;
    E.clear();
    run_tasks<MODE, true, false>(S, T, P, warp, lane, E, flag);
    if (P.smd_on && tid == kThreads - 1) smd_task(S, P, P.smd_centers ? __ldg(P.smd_centers + r) : P.smd_c0);
    __syncthreads();
    if (tid < N) {
      float fx, fy, fz;
      gather_forces<false>(S, P, tid, fx, fy, fz);
      float* f = forces + (r * N + tid) * 3;
      f[0] = fx; f[1] = fy; f[2] = fz;
    }
    if (P.smd_on && tid == 0) E.e[B200MD_E_EXTERNAL] += S.smd_energy[0];
    reduce_energies(S, E, warp, lane);
    if (tid == 0) {
#pragma unroll
      for (int k = 0; k < B200MD_NUM_ENERGIES; ++k) energies[r * B200MD_NUM_ENERGIES + k] = E.e[k];
    }
  }
}

// Integrator.step(niter) of upstream TorchMD fused with Forces.compute: positions, velocities and
// forces of a replica stay in registers / shared memory for all nsteps; HBM sees only the initial
// load, the final store, the observable records and the optional snapshots.
template <int MODE>
__global__ void __launch_bounds__(kThreads, 5)
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
  for (long long r = blockIdx.x; r < R; r += gridDim.x) {
    __syncthreads();
    float x = 0.f, y = 0.f, z = 0.f, vx = 0.f, vy = 0.f, vz = 0.f, fx = 0.f, fy = 0.f, fz = 0.f;
    uint32_t flag = 0;
    const long long base = (r * N + tid) * 3;
    if (live) {
      x = pos[base]; y = pos[base + 1]; z = pos[base + 2];
      vx = vel[base]; vy = vel[base + 1]; vz = vel[base + 2];
      fx = forces[base]; fy = forces[base + 1]; fz = forces[base + 2];
      if (x != x || y != y || z != z) flag |= B200MD_FLAG_NAN_POS;
    }
    if (box != nullptr && tid < 3 && __ldg(box + r * 9 + tid * 4) != 0.0f) flag |= B200MD_FLAG_PBC_UNSUPPORTED;
    if (status != nullptr && tid == 0) status[r] = 0;
    for (int s = 0; s < I.nsteps; ++s) {
      const unsigned long long gstep = I.step0 + (unsigned long long)s;
      // _first_VV: pos += vel dt + 0.5 a dt^2 ; vel += 0.5 dt a
      vx = fmaf(hdt_m, fx, vx); vy = fmaf(hdt_m, fy, vy); vz = fmaf(hdt_m, fz, vz);
      x = fmaf(dt, vx, x); y = fmaf(dt, vy, y); z = fmaf(dt, vz, z);
      if (live) {
        S.pos[tid] = make_float4(x, y, z, q);
        S.sx[tid] = x; S.sx[kMaxAtoms + tid] = y; S.sx[2 * kMaxAtoms + tid] = z; S.sx[3 * kMaxAtoms + tid] = q;
      }
      S.vel[tid] = make_float4(vx, vy, vz, 0.0f);  // parked: frees 6 registers across the task phase
      __syncthreads();
      const bool report = I.energy_every > 0 ? ((s + 1) % I.energy_every == 0) : (s == I.nsteps - 1);
      Energies E;
      if (report) {
        E.clear();
        run_tasks<MODE, true, true>(S, T, P, warp, lane, E, flag);
      } else {
        run_tasks<MODE, false, true>(S, T, P, warp, lane, E, flag);
      }
      if (P.smd_on && tid == kThreads - 1) smd_task(S, P, (float)((double)P.smd_c0 + (double)P.smd_rate * (double)(gstep + 1)));
      __syncthreads();
      {
        const float4 pp = S.pos[tid], vv = S.vel[tid];
        x = pp.x; y = pp.y; z = pp.z; vx = vv.x; vy = vv.y; vz = vv.z;
      }
      if (live) gather_forces<true>(S, P, tid, fx, fy, fz);
      if (langevin) {  // vel += -gamma vel dt + vcoeff N(0,1)
        float n[3];
        normal3(I.seed, (unsigned long long)(I.replica0 + r), (uint32_t)tid, gstep, n);
        vx += fmaf(vcoef, n[0], -vdamp * vx); vy += fmaf(vcoef, n[1], -vdamp * vy); vz += fmaf(vcoef, n[2], -vdamp * vz);
      }
      // _second_VV
      vx = fmaf(hdt_m, fx, vx); vy = fmaf(hdt_m, fy, vy); vz = fmaf(hdt_m, fz, vz);
      if (report) {
        if (live) E.e[B200MD_OBS_EKIN] = 0.5 * (double)mass * ((double)vx * vx + (double)vy * vy + (double)vz * vz);
        if (P.smd_on && tid == 0) { E.e[B200MD_E_EXTERNAL] += S.smd_energy[0]; E.e[B200MD_OBS_E2E_DIST] = (double)S.smd[4]; }
        reduce_energies(S, E, warp, lane);
        if (tid == 0 && observables != nullptr) {
          const long long slot = I.energy_every > 0 ? (s + 1) / I.energy_every - 1 : 0;
          double* o = observables + (slot * R + r) * B200MD_NUM_OBSERVABLES;
#pragma unroll
          for (int k = 0; k < B200MD_NUM_OBSERVABLES; ++k) o[k] = E.e[k];
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
    __syncthreads();
    if (flag && status != nullptr) atomicOr(status + r, flag);
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

struct b200md_topology {
  int device = 0;
  int natoms = 0;
  int num_sms = 0;
  DevTopo dev{};
  void* blob = nullptr;
  std::vector<uint32_t> mask_rows;  // [natoms][4], bit j set <=> pair (i, j) not evaluated
  int occ_forces[2] = {0, 0}, occ_integrate[2] = {0, 0};
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

// Orders the items of one bonded kind so that the 32 lanes of a chunk touch distinct shared-memory banks at every
// atom position: the fused integrator accumulates bonded forces with shared-memory atomics (bank = atom mod 32),
// which serialise on bank and on address collisions (4.6 wavefronts per ATOMS in topology order, profiles/r01_v11),
// and the LDS.128 position gathers want the 8 lanes of a quarter warp on distinct atoms mod 8.  `order` must
// already be grouped by `key` (items of equal key stay contiguous, so chunks keep running uniform term loops).
std::vector<int> spread_banks(const std::vector<int>& order, const int32_t* idx, int arity, const std::vector<long long>& key) {
  std::vector<int> out;
  std::vector<char> used(order.size(), 0);
  uint32_t busy[4] = {0, 0, 0, 0}, busy8[4] = {0, 0, 0, 0};
  int fill = 0;
  size_t g0 = 0;
  while (g0 < order.size()) {
    size_t g1 = g0;
    while (g1 < order.size() && key[order[g1]] == key[order[g0]]) ++g1;
    for (size_t left = g1 - g0; left > 0; --left) {
      if (fill == 32) { fill = 0; for (int k = 0; k < 4; ++k) busy[k] = 0; }
      if (fill % 8 == 0) for (int k = 0; k < 4; ++k) busy8[k] = 0;
      long best = -1;
      int bestc = 1 << 30;
      for (size_t m = g0; m < g1 && bestc > 0; ++m) {
        if (used[m]) continue;
        int c = 0;
        for (int k = 0; k < arity; ++k) {
          const int a = idx[arity * order[m] + k];
          c += 3 * ((busy[k] >> (a & 31)) & 1u) + ((busy8[k] >> (a & 7)) & 1u);
        }
        if (c < bestc) { bestc = c; best = (long)m; }
      }
      used[best] = 1;
      out.push_back(order[best]);
      for (int k = 0; k < arity; ++k) {
        const int a = idx[arity * order[best] + k];
        busy[k] |= 1u << (a & 31);
        busy8[k] |= 1u << (a & 7);
      }
      ++fill;
    }
    g0 = g1;
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

extern "C" b200md_topology* b200md_topology_create(const b200md_topology_desc* d, int device) {
  if (!d || d->natoms <= 0 || d->ntypes <= 0) { fail(B200MD_ERR_INVALID_ARGUMENT, "empty topology"); return nullptr; }
  if (d->natoms > kMaxAtoms) {
    fail(B200MD_ERR_TOO_MANY_ATOMS, "natoms > 128: the one-CTA-per-replica kernels hold a replica on chip");
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

  auto* topo = new b200md_topology();
  topo->device = device;
  topo->natoms = N;
  DevTopo& T = topo->dev;
  T.natoms = N; T.ntypes = NT; T.nchunks = (N + kChunk - 1) / kChunk;
  T.nbonds = d->nbonds; T.nangles = d->nangles; T.ndih = d->ndihedrals; T.nimp = d->nimpropers; T.n14 = d->n14;

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
  // Dihedral items are processed in an internal order sorted by (number of terms, branch kind, largest
  // periodicity), so the 32 lanes of a chunk run the same term loop; par.dihedrals itself is untouched.
  std::vector<int> dperm(d->ndihedrals), dinv(d->ndihedrals);
  {
    std::vector<long long> key(d->ndihedrals, 0);
    std::vector<int> nterm(d->ndihedrals, 0), maxper(d->ndihedrals, 0), special(d->ndihedrals, 0);
    for (int t = 0; t < d->ndihedral_terms; ++t) {
      const int o = d->dihedral_term_owner[t];
      if (o < 0 || o >= d->ndihedrals) { delete topo; fail(B200MD_ERR_INVALID_ARGUMENT, "dihedral term owner out of range"); return nullptr; }
      const float per = d->dihedral_term_prm[3 * t + 2];
      nterm[o]++;
      const bool fast = !d->dihedral_term_harmonic[t] && per >= 1.0f && per <= 8.0f && per == floorf(per);
      if (!fast) special[o] = 1; else maxper[o] = std::max(maxper[o], (int)per);
    }
    for (int k = 0; k < d->ndihedrals; ++k) key[k] = ((long long)special[k] << 40) | ((long long)nterm[k] << 20) | maxper[k];
    for (int k = 0; k < d->ndihedrals; ++k) dperm[k] = k;
    std::stable_sort(dperm.begin(), dperm.end(), [&](int a, int b) { return key[a] < key[b]; });
    dperm = spread_banks(dperm, d->dihedral_idx, 4, key);
    for (int k = 0; k < d->ndihedrals; ++k) dinv[dperm[k]] = k;
  }
  std::vector<int32_t> dihedral_idx_s((size_t)4 * std::max(d->ndihedrals, 1)), dihedral_owner_s(std::max(d->ndihedral_terms, 1));
  for (int k = 0; k < d->ndihedrals; ++k)
    for (int c = 0; c < 4; ++c) dihedral_idx_s[4 * k + c] = d->dihedral_idx[4 * dperm[k] + c];
  for (int t = 0; t < d->ndihedral_terms; ++t) dihedral_owner_s[t] = dinv[d->dihedral_term_owner[t]];
  // bonds and angles in bank-spread order too (energies are sums, the slot order stays fixed: results are deterministic)
  std::vector<int32_t> bond_idx_p((size_t)2 * std::max(d->nbonds, 1)), angle_idx_p((size_t)3 * std::max(d->nangles, 1));
  std::vector<float> bond_prm_p((size_t)2 * std::max(d->nbonds, 1)), angle_prm_p((size_t)2 * std::max(d->nangles, 1));
  {
    std::vector<int> ob(d->nbonds), oa(d->nangles);
    for (int k = 0; k < d->nbonds; ++k) ob[k] = k;
    for (int k = 0; k < d->nangles; ++k) oa[k] = k;
    ob = spread_banks(ob, d->bond_idx, 2, std::vector<long long>(d->nbonds, 0));
    oa = spread_banks(oa, d->angle_idx, 3, std::vector<long long>(d->nangles, 0));
    for (int k = 0; k < d->nbonds; ++k) {
      for (int c = 0; c < 2; ++c) { bond_idx_p[2 * k + c] = d->bond_idx[2 * ob[k] + c]; bond_prm_p[2 * k + c] = d->bond_prm[2 * ob[k] + c]; }
    }
    for (int k = 0; k < d->nangles; ++k) {
      for (int c = 0; c < 3; ++c) angle_idx_p[3 * k + c] = d->angle_idx[3 * oa[k] + c];
      for (int c = 0; c < 2; ++c) angle_prm_p[2 * k + c] = d->angle_prm[2 * oa[k] + c];
    }
  }
  std::vector<uint32_t> bond_idx(d->nbonds), angle_idx(d->nangles), dih_idx(d->ndihedrals), imp_idx(d->nimpropers);
  std::vector<float2> bond_prm(d->nbonds), angle_prm(d->nangles);
  for (int k = 0; k < d->nbonds; ++k) {
    bond_idx[k] = pack4(bond_idx_p.data() + 2 * k, 2);
    bond_prm[k] = make_float2(bond_prm_p[2 * k], bond_prm_p[2 * k + 1]);
  }
  for (int k = 0; k < d->nangles; ++k) {
    angle_idx[k] = pack4(angle_idx_p.data() + 3 * k, 3);
    angle_prm[k] = make_float2(angle_prm_p[2 * k], angle_prm_p[2 * k + 1]);
  }
  for (int k = 0; k < d->ndihedrals; ++k) dih_idx[k] = pack4(dihedral_idx_s.data() + 4 * k, 4);
  for (int k = 0; k < d->nimpropers; ++k) imp_idx[k] = pack4(d->improper_idx + 4 * k, 4);
  // 1-4 pairs (forces.py:176-226): fused into the first dihedral item whose end atoms they are; pairs
  // without such a dihedral stay in their own ("orphan") list.  Stored as (12 A/scnb, 6 B/scnb, K qi qj/scee).
  std::vector<float4> dih_p14(std::max(d->ndihedrals, 1), make_float4(0.f, 0.f, 0.f, 0.f)), p14_prm;
  std::vector<uint32_t> p14_idx;
  std::vector<int> orphan_src;
  for (int k = 0; k < d->n14; ++k) {
    const int i = d->idx14[2 * k], j = d->idx14[2 * k + 1];
    const float* p = d->prm14 + 4 * k;  // A, B, scnb, scee (forces.py:197-200)
    const double qi = d->charge ? d->charge[i] : 0.0, qj = d->charge ? d->charge[j] : 0.0;
    float4 prm = make_float4((float)(12.0 * (double)p[0] / p[2]), (float)(6.0 * (double)p[1] / p[2]),
                             (float)(332.06371307417066 * qi * qj / p[3]), 1.0f);
    int owner = -1;
    for (int q = 0; q < d->ndihedrals && owner < 0; ++q) {
      const int32_t* a = dihedral_idx_s.data() + 4 * q;
      if (dih_p14[q].w == 0.0f && ((a[0] == i && a[3] == j) || (a[0] == j && a[3] == i))) owner = q;
    }
    if (owner >= 0) {
      dih_p14[owner] = prm;
    } else {
      p14_idx.push_back(pack4(d->idx14 + 2 * k, 2));
      p14_prm.push_back(prm);
      orphan_src.push_back(k);
    }
  }
  T.n14 = (int)p14_idx.size();
  T.dih_on = 1;
  T.p14_on = 1;
  std::vector<uint2> dih_terms, imp_terms;
  std::vector<float4> dih_tp, imp_tp;
  build_terms(d->ndihedrals, d->ndihedral_terms, dihedral_owner_s.data(), d->dihedral_term_prm,
              d->dihedral_term_harmonic, dih_terms, dih_tp);
  build_terms(d->nimpropers, d->nimproper_terms, d->improper_term_owner, d->improper_term_prm,
              d->improper_term_harmonic, imp_terms, imp_tp);

  // ---- atom-major force slots: contribution (item, position) of atom a gets the next slot of a -----
  // (fixed order bonds, angles, dihedrals, impropers, orphan 1-4 => deterministic sums in the gather)
  std::vector<int> orphan_atoms(2 * p14_idx.size());
  for (size_t k = 0; k < p14_idx.size(); ++k) {
    orphan_atoms[2 * k] = d->idx14[2 * orphan_src[k]];
    orphan_atoms[2 * k + 1] = d->idx14[2 * orphan_src[k] + 1];
  }
  const int counts[6] = {0, d->nbonds, d->nangles, d->ndihedrals, d->nimpropers, (int)p14_idx.size()};
  const int arity[6] = {0, 2, 3, 4, 4, 2};
  const int32_t* idxs[6] = {nullptr, bond_idx_p.data(), angle_idx_p.data(), dihedral_idx_s.data(), d->improper_idx, orphan_atoms.data()};
  std::vector<int> per_atom_count(kMaxAtoms, 0);
  for (int ty = 1; ty < 6; ++ty)
    for (int it = 0; it < counts[ty] * arity[ty]; ++it) per_atom_count[idxs[ty][it]]++;
  std::vector<uint16_t> gstart(kMaxAtoms + 2, 0);
  int nslots = 0;
  for (int a = 0; a < kMaxAtoms; ++a) { gstart[a] = (uint16_t)nslots; nslots += per_atom_count[a]; }
  gstart[kMaxAtoms] = gstart[kMaxAtoms + 1] = (uint16_t)nslots;
  if (nslots > 65535) { delete topo; fail(B200MD_ERR_UNSUPPORTED, "too many bonded force contributions"); return nullptr; }
  T.nslots = nslots;
  std::vector<int> cursor(gstart.begin(), gstart.begin() + kMaxAtoms);
  std::vector<std::vector<uint16_t>> slot_of(6);
  for (int ty = 1; ty < 6; ++ty) {
    slot_of[ty].resize((size_t)counts[ty] * arity[ty]);
    for (int it = 0; it < counts[ty] * arity[ty]; ++it) slot_of[ty][it] = (uint16_t)cursor[idxs[ty][it]]++;
  }
  auto pair16 = [](uint16_t lo, uint16_t hi) { return (uint32_t)lo | ((uint32_t)hi << 16); };
  std::vector<uint32_t> bond_slot(d->nbonds), p14_slot(p14_idx.size());
  std::vector<uint2> angle_slot(d->nangles), dih_slot(std::max(d->ndihedrals, 1)), imp_slot(std::max(d->nimpropers, 1));
  for (int k = 0; k < d->nbonds; ++k) bond_slot[k] = pair16(slot_of[1][2 * k], slot_of[1][2 * k + 1]);
  for (int k = 0; k < d->nangles; ++k)
    angle_slot[k] = make_uint2(pair16(slot_of[2][3 * k], slot_of[2][3 * k + 1]), slot_of[2][3 * k + 2]);
  for (int k = 0; k < d->ndihedrals; ++k)
    dih_slot[k] = make_uint2(pair16(slot_of[3][4 * k], slot_of[3][4 * k + 1]), pair16(slot_of[3][4 * k + 2], slot_of[3][4 * k + 3]));
  for (int k = 0; k < d->nimpropers; ++k)
    imp_slot[k] = make_uint2(pair16(slot_of[4][4 * k], slot_of[4][4 * k + 1]), pair16(slot_of[4][4 * k + 2], slot_of[4][4 * k + 3]));
  for (size_t k = 0; k < p14_idx.size(); ++k) p14_slot[k] = pair16(slot_of[5][2 * k], slot_of[5][2 * k + 1]);

  // ---- static task schedule: longest-processing-time-first over the 4 warps ----------------------
  struct Task { uint32_t word; int cost; };
  std::vector<Task> all;
  // cost model of the schedule (issue slots per warp); B200MD_COSTS="pass,chunk,diagstep,diag,bond,angle,dih,imp,p14"
  // overrides it for tuning runs
  int cm[9] = {150, 120, 52, 30, 60, 140, 330, 300, 50};
  if (const char* env = getenv("B200MD_COSTS")) sscanf(env, "%d,%d,%d,%d,%d,%d,%d,%d,%d", cm, cm + 1, cm + 2, cm + 3, cm + 4, cm + 5, cm + 6, cm + 7, cm + 8);
  const int nblocks = (N + 31) / 32;
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
  const int chunk_cost[6] = {0, cm[4], cm[5], cm[6], cm[7], cm[8]};
  for (int ty = 1; ty < 6; ++ty)
    for (int c = 0; c < (counts[ty] + 31) / 32; ++c) all.push_back({make_task(ty, c, 0), chunk_cost[ty]});
  std::stable_sort(all.begin(), all.end(), [](const Task& a, const Task& b) { return a.cost > b.cost; });
  std::vector<uint32_t> tasks((size_t)kWarps * kMaxTasks, 0);
  int load[kWarps] = {0, 0, 0, 0};
  std::vector<std::vector<uint32_t>> mine(kWarps);
  for (const Task& t : all) {
    int w = 0;
    for (int k = 1; k < kWarps; ++k) if (load[k] < load[w]) w = k;
    if ((int)mine[w].size() >= kMaxTasks - 1) { delete topo; fail(B200MD_ERR_UNSUPPORTED, "too many tasks per warp"); return nullptr; }
    mine[w].push_back(t.word);
    load[w] += t.cost;
  }
  // each warp runs its tasks kind by kind (run_tasks): tiles, diagonal blocks, dihedrals, impropers, angles, bonds, 1-4
  const uint32_t kind_order[kNumKinds] = {TASK_TILE, TASK_DIAG, TASK_DIHEDRAL, TASK_IMPROPER, TASK_ANGLE, TASK_BOND, TASK_P14};
  for (int w = 0; w < kWarps; ++w) {
    int n = 0;
    for (int k = 0; k < kNumKinds; ++k) {
      for (uint32_t word : mine[w]) if ((word & 0xff) == kind_order[k]) tasks[(size_t)w * kMaxTasks + n++] = word;
      T.kend[w][k] = n;
    }
    T.kend[w][kNumKinds] = n;
  }

  // ---- upload --------------------------------------------------------------------------------------
  BlobBuilder B;
  const size_t o_ab = B.add(ab), o_q = B.add(charge), o_m = B.add(mass), o_ro = B.add(rowoff), o_mask = B.add(mask8), o_rot = B.add(diagmask);
  const size_t o_bi = B.add(bond_idx), o_bp = B.add(bond_prm), o_bs = B.add(bond_slot);
  const size_t o_ai = B.add(angle_idx), o_ap = B.add(angle_prm), o_as = B.add(angle_slot);
  const size_t o_di = B.add(dih_idx), o_dt = B.add(dih_terms), o_dp = B.add(dih_tp), o_ds = B.add(dih_slot), o_d14 = B.add(dih_p14);
  const size_t o_ii = B.add(imp_idx), o_it = B.add(imp_terms), o_ip = B.add(imp_tp), o_is = B.add(imp_slot);
  const size_t o_pi = B.add(p14_idx), o_pp = B.add(p14_prm), o_ps = B.add(p14_slot);
  const size_t o_gs = B.add(gstart), o_tk = B.add(tasks);
  auto cleanup = [&](const std::string& msg) { fail(B200MD_ERR_CUDA, msg); if (topo->blob) cudaFree(topo->blob); delete topo; };
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) { cleanup(std::string("cudaSetDevice: ") + cudaGetErrorString(e)); return nullptr; }
  e = cudaMalloc(&topo->blob, B.bytes.size());
  if (e != cudaSuccess) { topo->blob = nullptr; cleanup(std::string("cudaMalloc: ") + cudaGetErrorString(e)); return nullptr; }
  e = cudaMemcpy(topo->blob, B.bytes.data(), B.bytes.size(), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) { cleanup(std::string("cudaMemcpy: ") + cudaGetErrorString(e)); return nullptr; }
  char* base = (char*)topo->blob;
  T.ab = (const float2*)(base + o_ab); T.charge = (const float*)(base + o_q); T.mass = (const float*)(base + o_m);
  T.ab_row_off = (const int*)(base + o_ro); T.mask8 = (const uint8_t*)(base + o_mask); T.diagmask = (const uint32_t*)(base + o_rot);
  T.bond_idx = (const uint32_t*)(base + o_bi); T.bond_prm = (const float2*)(base + o_bp); T.bond_slot = (const uint32_t*)(base + o_bs);
  T.angle_idx = (const uint32_t*)(base + o_ai); T.angle_prm = (const float2*)(base + o_ap); T.angle_slot = (const uint2*)(base + o_as);
  T.dih_idx = (const uint32_t*)(base + o_di); T.dih_terms = (const uint2*)(base + o_dt); T.dih_term_prm = (const float4*)(base + o_dp);
  T.dih_slot = (const uint2*)(base + o_ds); T.dih_p14 = (const float4*)(base + o_d14);
  T.imp_idx = (const uint32_t*)(base + o_ii); T.imp_terms = (const uint2*)(base + o_it); T.imp_term_prm = (const float4*)(base + o_ip);
  T.imp_slot = (const uint2*)(base + o_is);
  T.p14_idx = (const uint32_t*)(base + o_pi); T.p14_prm = (const float4*)(base + o_pp); T.p14_slot = (const uint32_t*)(base + o_ps);
  T.gather_start = (const uint16_t*)(base + o_gs);
  T.tasks = (const uint32_t*)(base + o_tk);
  T.smem_bytes = (int)smem_layout(nullptr, nullptr, NT, nslots, false);
  T.smem_bytes_fixed = (int)smem_layout(nullptr, nullptr, NT, nslots, true);
  if (T.smem_bytes > 227 * 1024) { cleanup("topology needs more than 227 KB of shared memory per CTA"); g_last_error = "too much shared memory"; return nullptr; }
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, device);
  topo->num_sms = prop.multiProcessorCount;
  topo->occ_forces[0] = query_occupancy(forces_kernel<0>, T.smem_bytes);
  topo->occ_forces[1] = query_occupancy(forces_kernel<1>, T.smem_bytes);
  topo->occ_integrate[0] = query_occupancy(integrate_kernel<0>, T.smem_bytes_fixed);
  topo->occ_integrate[1] = query_occupancy(integrate_kernel<1>, T.smem_bytes_fixed);
  if (topo->occ_forces[0] < 1 || topo->occ_integrate[0] < 1) { cleanup("kernel does not fit on an SM"); return nullptr; }
  g_last_error.clear();
  return topo;
}

extern "C" void b200md_topology_destroy(b200md_topology* topo) {
  if (!topo) return;
  cudaSetDevice(topo->device);
  if (topo->blob) cudaFree(topo->blob);
  delete topo;
}

extern "C" int b200md_topology_exclusion_mask(const b200md_topology* topo, uint32_t* out) {
  if (!topo || !out) return fail(B200MD_ERR_INVALID_ARGUMENT, "null argument");
  memcpy(out, topo->mask_rows.data(), topo->mask_rows.size() * sizeof(uint32_t));
  return B200MD_OK;
}

extern "C" int b200md_topology_natoms(const b200md_topology* topo) { return topo ? topo->natoms : -1; }

// ---- launch helpers ---------------------------------------------------------------------------------
static int make_force_params(const b200md_topology* topo, const b200md_force_opts* o, const b200md_smd_opts* smd,
                             bool integrator, ForceParams* P) {
  if (!o) return fail(B200MD_ERR_INVALID_ARGUMENT, "force options are required");
  if (o->terms & (B200MD_T_REPULSION | B200MD_T_REPULSIONCG))
    return fail(B200MD_ERR_UNSUPPORTED, "repulsion / repulsioncg terms are not implemented");
  memset(P, 0, sizeof(*P));
  P->terms = o->terms;
  P->has_cutoff = o->cutoff > 0.0f;
  P->rc = o->cutoff;
  P->rc2 = o->cutoff * o->cutoff;
  const bool sw = P->has_cutoff && o->switch_dist > 0.0f;  // forces.py:402
  P->rs = sw ? o->switch_dist : INFINITY;
  P->inv_sw = sw ? 1.0f / (o->cutoff - o->switch_dist) : 0.0f;
  if (o->rfa) {
    if (!P->has_cutoff) return fail(B200MD_ERR_INVALID_ARGUMENT, "rfa needs a cutoff");
    const double eps = o->solvent_dielectric, rc = o->cutoff, denom = 2.0 * eps + 1.0;  // forces.py:466-468
    P->krf = (float)((1.0 / (rc * rc * rc)) * (eps - 1.0) / denom);
    P->crf = (float)((1.0 / rc) * (3.0 * eps) / denom);
  }
  P->lj_on = (o->terms & B200MD_T_LJ) ? 1.0f : 0.0f;
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
  if (!(terms & B200MD_T_BONDS)) T.nbonds = 0;
  if (!(terms & B200MD_T_ANGLES)) T.nangles = 0;
  if (!(terms & B200MD_T_IMPROPERS)) T.nimp = 0;
  T.dih_on = (terms & B200MD_T_DIHEDRALS) ? 1 : 0;
  T.p14_on = p14 ? 1 : 0;
  if (!T.dih_on && !T.p14_on) T.ndih = 0;
  if (!p14) T.n14 = 0;
  return T;
}

extern "C" int b200md_forces_compute(const b200md_topology* topo, const b200md_force_opts* opts,
                                     const b200md_smd_opts* smd, int64_t R, const float* pos, const float* box,
                                     float* forces, double* energies, uint32_t* status, void* stream) {
  if (!topo || !pos || !forces || !energies) return fail(B200MD_ERR_INVALID_ARGUMENT, "null argument");
  if (R < 0) return fail(B200MD_ERR_INVALID_ARGUMENT, "negative replica count");
  if (R == 0) return B200MD_OK;
  ForceParams P;
  int rc = make_force_params(topo, opts, smd, false, &P);
  if (rc) return rc;
  DevTopo T = apply_terms(topo->dev, opts->terms);
  CUDA_TRY(cudaSetDevice(topo->device));
  const int mode = P.has_cutoff ? 1 : 0;
  const long long cap = (long long)topo->num_sms * topo->occ_forces[mode];
  const int grid = (int)std::min<long long>(R, cap);
  cudaStream_t st = (cudaStream_t)stream;
  if (mode == 0)
    forces_kernel<0><<<grid, kThreads, T.smem_bytes, st>>>(T, P, R, pos, box, forces, energies, status);
  else
    forces_kernel<1><<<grid, kThreads, T.smem_bytes, st>>>(T, P, R, pos, box, forces, energies, status);
  g_launches.fetch_add(1);
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
  ForceParams P;
  int rc = make_force_params(topo, fopts, smd, true, &P);
  if (rc) return rc;
  DevTopo T = apply_terms(topo->dev, fopts->terms);
  StepParams I;
  I.dt = io->dt; I.gamma = io->gamma; I.temperature = io->temperature;
  I.seed = io->seed; I.step0 = io->step0; I.replica0 = io->replica0;
  I.nsteps = io->nsteps; I.energy_every = io->energy_every; I.snapshot_every = io->snapshot_every;
  CUDA_TRY(cudaSetDevice(topo->device));
  const int mode = P.has_cutoff ? 1 : 0;
  // long launches: one CTA per replica, so the hardware block scheduler hands a finished SM its next replica
  // (a static grid-stride split of 8 192 replicas over 740 resident CTAs ends with a 12th round on 52 CTAs);
  // short launches keep the grid-stride loop to stage the tables once per resident CTA
  const long long cap = (long long)topo->num_sms * topo->occ_integrate[mode];
  const int grid = (int)((io->nsteps >= 32 && R < (1ll << 30)) ? R : std::min<long long>(R, cap));
  cudaStream_t st = (cudaStream_t)stream;
  if (mode == 0)
    integrate_kernel<0><<<grid, kThreads, T.smem_bytes_fixed, st>>>(T, P, I, R, pos, vel, forces, box, observables, snapshots, status);
  else
    integrate_kernel<1><<<grid, kThreads, T.smem_bytes_fixed, st>>>(T, P, I, R, pos, vel, forces, box, observables, snapshots, status);
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
  CUDA_TRY(cudaSetDevice(topo->device));
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
          if (f & B200MD_FLAG_PBC_UNSUPPORTED) { rc = fail(B200MD_ERR_UNSUPPORTED, "non-zero box: periodic wrapping is not implemented"); break; }
        }
    }
  }
  cudaFree(dpos); cudaFree(dbox); cudaFree(dfrc); cudaFree(dene); cudaFree(dst);
  if (e != cudaSuccess) return fail(B200MD_ERR_CUDA, cudaGetErrorString(e));
  return rc;
}

extern "C" double b200md_fp32_peak_tflops(int device, int iters) {
  if (cudaSetDevice(device) != cudaSuccess) { fail(B200MD_ERR_CUDA, "cudaSetDevice failed"); return -1.0; }
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
