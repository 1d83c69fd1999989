// b200md_kernels.cuh -- device code of the B200-native TorchMD force-field + integrator path.
//
// Mapping: one CTA of 128 threads (4 warps) per replica.  Positions (+ charge) live in
// shared memory as float4[128] and structure-of-arrays; the off-diagonal part of the non-bonded half
// matrix is cut into "j-chunks" of 8 atoms: a warp that owns a chunk sweeps the 32-lane i-blocks below
// it in packed FP32 (two j atoms per FFMA2), the i-side force of a pass goes to the warp's shared-memory
// partial row, the 8 x 3 j-force partials are butterfly-reduced with warp shuffles once per chunk; the
// diagonal 32 x 32 blocks are evaluated full shell.  Bonded terms run two items per lane in packed FP32;
// their force vectors go to private atom-major slots summed in a fixed order (single-shot
// Forces.compute) or to 2^-17 fixed-point shared-memory integer atomics (fused integrator): every
// reduction is associative or fixed-order, hence bit-reproducible.  Energies are NOT part of the force
// tasks: a separate FP64 pass (energy_pass, b200md_forces.cuh) evaluates every term in double from the
// fp32 positions whenever energies are wanted (Forces.compute, report steps of the integrator).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200md.h"

namespace b200md {

constexpr int kThreads = 128;
constexpr int kWarps = 4;
constexpr int kMaxAtoms = B200MD_MAX_ATOMS;  // 128
constexpr int kChunk = 8;                    // atoms per j-chunk
constexpr int kMaxChunks = kMaxAtoms / kChunk;
constexpr int kAbStride = kMaxAtoms + 1;     // float2 per LJ-table row of the global (HBM) table
constexpr int kTabStride = kMaxAtoms + 4;    // floats per row of the shared-memory 12A / -6B tables: rows of
                                             // different types start 4 banks apart, so the per-lane LDS.128 of a
                                             // quarter warp (<= 8 types) is conflict-free
constexpr int kMaxTasks = 64;                // per warp
constexpr float kElecFactor = 332.06371307417066f;  // forces.py:373-376
constexpr float kBoltzman = 0.001987191f;           // upstream torchmd.integrator

// task kinds (one task = one warp-wide unit of work, scheduled statically on the host)
enum TaskKind : uint32_t { TASK_TILE = 0, TASK_BOND = 1, TASK_ANGLE = 2, TASK_DIHEDRAL = 3, TASK_IMPROPER = 4, TASK_P14 = 5, TASK_DIAG = 6 };
// execution order of the kinds inside a warp's task list
constexpr int kNumKinds = 7;
constexpr int kMaxPer = 6;  // largest periodicity of the packed torsion series
__host__ __device__ inline uint32_t make_task(uint32_t kind, uint32_t chunk, uint32_t passes) {
  return kind | (chunk << 8) | (passes << 24);
}

// torsion term modes (float4 term = k, per | marker, c0, s0)
constexpr float kTermHarmonic = -1.0f;  // CHARMM harmonic branch (forces.py:570-576): c0 = phi0
constexpr float kTermGeneral = -2.0f;   // AMBER branch with a non-integer periodicity: c0 = phi0, s0 = per

// Device view of a packed topology.  All pointers point into one device allocation.
struct DevTopo {
  int natoms, ntypes, nchunks;
  const float2* ab;         // [ntypes][kAbStride]  (12 A, 6 B) of (type t, atom j)
  const float* charge;      // [128]
  const float* mass;        // [128]
  const int* ab_row_off;    // [128] byte offset of atom's row in the shared-memory 12A / -6B tables
  const uint8_t* mask8;     // [kMaxChunks][128] bit s: pair (i, 8*chunk+s) is NOT evaluated
  const uint32_t* diagmask;  // [128] bit j: pair (i, 32*block(i) + j) is not evaluated in the diagonal block
  // packed bonded kinds: lane p of chunk c (pair index 32c + p) evaluates two items (A, B) in f32x2
  int nbond2, nangle2, ndih2, n14_2;      // lane pairs in use per kind
  int dih_on, p14_on;                     // proper-torsion part / fused 1-4 part of the packed dihedral items
  const uint2* bond2_idx;  const float4* bond2_prm;  const uint4* bond2_slot;    // atoms 4 x u8 (A, B); (kA, kB, r0A, r0B)
  const uint2* angle2_idx; const float4* angle2_prm; const uint4* angle2_slot;   // (kA, kB, theta0A, theta0B)
  const uint2* dih2_idx;   const float4* dih2_coef;  const float2* dih2_k0;      // coef[n-1][pair] = n a_n (A, B), n b_n (A, B)
  const uint4* dih2_slot;  const float4* dih2_p14ab; const float2* dih2_p14q;    // fused 1-4: 12A' (A, B), -6B' (A, B); K qi qj'
  int dih2_stride;
  const uint2* p142_idx;   const float4* p142_ab;    const float2* p142_q;   const uint4* p142_slot;
  // general torsions (one item per lane): special dihedrals [0, nspec) then impropers [nspec, ngen)
  int ngen, nspec, gen_begin, gen_end;
  const uint32_t* gen_idx;   const uint2* gen_terms;  const float4* gen_term_prm;  const uint2* gen_slot;
  int nslots;
  const uint16_t* gather_start;  // [129]: slots of atom a are [gather_start[a], gather_start[a+1])
  const uint32_t* tasks;    // [kWarps][kMaxTasks] static schedule of the force tasks (issue-slot cost model)
  uint2 kend8[kWarps];      // per warp: end index of the k-th run of its kind-sorted task list, one byte per kind
  // FP64 energy pass (energy_pass): plain item lists in topology order, one item per thread
  int ne_pair, ne_14, ne_bond, ne_angle, ne_dih, ne_imp;
  const uint32_t* e_pair;   // [ne_pair] i | j << 8 | (type_i * ntypes + type_j) << 16: Forces.ava_idx, row-major
  const double2* e_ab;      // [ntypes^2] (A, B) of par.A / par.B (forces.py:379-383)
  const uint32_t* e14_idx;  // [ne_14] i | j << 8
  const double* e14_prm;    // [3 ne_14] A / scnb, B / scnb, K qi qj / scee (forces.py:197-226)
  const uint32_t* eb_idx;   const float2* eb_prm;   // bonds: i | j << 8; (k, r0)
  const uint32_t* ea_idx;   const float2* ea_prm;   // angles: 3 x u8; (k, theta0)
  const uint32_t* et_idx;   const int* et_start;    // torsions, dihedrals [0, ne_dih) then impropers: 4 x u8; terms [start[d], start[d + 1])
  const float4* et_term;    const double2* et_cs;   // (k, phi0, per, kind: 0 AMBER, 1 harmonic, 2 AMBER with integer per); (cos phi0, sin phi0)
  int smem_bytes, smem_bytes_fixed;
  // proper-dihedral angle output (Energy.calc_phi of the ML callers) and its backward
  int nphi;                      // = par.dihedrals rows
  const uchar4* phi_idx;         // [nphi] atoms of dihedral d
  const uint16_t* phi_start;     // [129] items of atom a are [phi_start[a], phi_start[a + 1])
  const uint16_t* phi_items;     // [4 nphi] dihedral | role << 14 (role = position of the atom in the quadruple)
  // rigid X-H bonds (b200md_integrate_opts.rigid_bonds): star groups, one per heavy atom that carries hydrogens,
  // sorted by size (largest first) and handled by thread g of the CTA
  const int4* cgroup;     // [128] (parent | nH << 8, h0, h1, h2)
  const float4* cdist2;   // [128] squared lengths of the group's constraints
  const float4* cweight;  // [128] 1 / mass of the parent and of its hydrogens
  int ncons;              // total; -1: a group the kernel cannot hold (see include/b200md.h)
  int ngroups;
  uint32_t cnmax;         // byte w: largest group among the lanes of warp w (selects the 1 x 1 or 3 x 3 solve)
};

struct ForceParams {
  uint32_t terms;
  int has_cutoff;
  float rc, rc2, rs, inv_sw;  // cutoff, cutoff^2, switch_dist (or +inf), 1/(rc-rs)
  float krf, crf;             // reaction-field constants (0 when rfa is off)
  float lj_on, el_on;         // 1-4 block switches (forces.py:176-226 tests "lj" / "electrostatics")
  float tab_a, tab_b;         // pair tables = (tab_a * 12A, tab_b * 6B): lj (1, -1), repulsion (1, 0), repulsioncg (0, +1)
  int smd_on, smd_a1, smd_a2;
  float smd_fk, smd_c0, smd_rate;
  const float* smd_centers;
  // the same constants in double for the FP64 energy pass (computed on the host from the float options)
  double rc_d, rs_d, inv_sw_d, krf_d, crf_d;
};

struct StepParams {
  float dt, gamma, temperature;
  unsigned long long seed, step0;
  long long replica0;
  int nsteps, energy_every, snapshot_every;
  // time slicing (b200md.cu, integrate_kernel): a work unit is `slice_len` consecutive steps of one replica; units are
  // handed out slice-major from a global ticket counter, a replica's state parks in the caller's pos / vel / forces
  // arrays between its slices
  int slice_len, nslices;
  unsigned long long* ticket;   // [1], zeroed before the launch
  unsigned int* slice_done;     // [R] slices of replica r that are complete, zeroed before the launch
  double* smd_work;             // [R] work of the moving restraint carried from slice to slice (SMD only)
  // deferred energies: on a report step the kernel writes the positions here ([nreports][R][natoms][3]) and energy_kernel
  // evaluates the 7 force-field terms of all records behind the launch; nullptr: FP64 energy pass inside the kernel
  float* report_pos;
};

// ---- shared-memory carve-up -----------------------------------------------------------------
struct Smem {
  float4* pos;        // [128] (x, y, z, q) per atom: bonded items and the i side of the pair tiles
  float* sx;          // [4][128] the same as x[], y[], z[], q[]: two j atoms arrive as one packed f32x2 operand
  float* ta;          // [ntypes][kTabStride] 12 A of (type t, atom j)
  float* tb;          // [ntypes][kTabStride] -6 B of (type t, atom j)
  int* ab_row_off;    // [128]
  uint8_t* mask8;     // [kMaxChunks*128]
  uint32_t* diagmask;  // [128]
  float* fi;          // [kWarps][3][128]
  float* fj;          // [3][128]
  float* fb;          // [3][fbs]: atom-major bonded force slots (slot mode), one plane per component
  int fbs;            // plane stride = nslots rounded up to 32
  int* fbi;           // [3][128] fixed-point bonded force accumulators (fixed mode)
  float4* vel;        // [128] velocities parked during the task phase of the integrator
  uint16_t* gstart;   // [130]
  uint32_t* tasks;    // [kWarps*kMaxTasks]
  double* ered;       // [kWarps][kNumReduced]
  float* smd;         // [8]: force on atom2 (xyz), -, distance
  double* smd_energy; // [2]: restraint energy of this step, work accumulated since the first step of the launch
  float* box;         // [8]: -box diagonal (x, y, z), 0, 1 / box diagonal (0 where the entry is 0), 0
  double* pd;         // [4][128] x, y, z, q in double for the energy pass: OVERLAYS vel + sx (free between the force
                      // gather of a step and the staging of the next one)
  unsigned long long* tick;  // [1] ticket of the unit this CTA works on (integrator)
};

__host__ __device__ inline size_t align16(size_t x) { return (x + 15) & ~size_t(15); }

// Every array except the two whose size depends on the topology sits at a compile-time offset, so shared
// addresses fold into LDS/STS immediates (a run-time layout cost 11 % of all executed instructions in
// address re-materialisation, profiles/r01_v4).  Order: fixed arrays, LJ table (ntypes rows), slot buffer.
constexpr size_t kOffPos = 0;
constexpr size_t kOffRowOff = kOffPos + sizeof(float4) * kMaxAtoms;
constexpr size_t kOffMask8 = kOffRowOff + sizeof(int) * kMaxAtoms;
constexpr size_t kOffDiagMask = kOffMask8 + kMaxChunks * kMaxAtoms;
constexpr size_t kOffFi = kOffDiagMask + sizeof(uint32_t) * kMaxAtoms;
constexpr size_t kOffFj = kOffFi + sizeof(float) * kWarps * 3 * kMaxAtoms;
constexpr size_t kOffGstart = kOffFj + sizeof(float) * 3 * kMaxAtoms;
constexpr size_t kOffTasks = kOffGstart + 272;
// observables that are block-reduced (8 energy slots, Ekin, end-to-end distance); the SMD work lives in S.smd_energy[1]
constexpr int kNumReduced = 10;
constexpr size_t kOffEred = kOffTasks + sizeof(uint32_t) * kWarps * kMaxTasks;
constexpr size_t kOffSmd = kOffEred + sizeof(double) * kWarps * kNumReduced;
constexpr size_t kOffSmdEnergy = kOffSmd + 32;
constexpr size_t kOffBox = kOffSmdEnergy + 16;
constexpr size_t kOffTick = kOffBox + 32;
constexpr size_t kOffFbi = kOffTick + 16;
constexpr size_t kOffVel = kOffFbi + sizeof(int) * 3 * kMaxAtoms;
constexpr size_t kOffSx = kOffVel + sizeof(float4) * kMaxAtoms;
constexpr size_t kOffAb = kOffSx + sizeof(float) * 4 * kMaxAtoms;
static_assert(kOffAb % 16 == 0 && kOffFi % 16 == 0 && kOffEred % 16 == 0 && kOffVel % 16 == 0, "shared-memory layout must stay 16-byte aligned");
static_assert(kOffAb - kOffVel == sizeof(double) * 4 * kMaxAtoms, "the FP64 position overlay needs exactly the vel + sx area");

__host__ __device__ inline size_t smem_layout(Smem* s, char* base, int ntypes, int nslots, bool fixed) {
  const size_t tab = align16(sizeof(float) * (size_t)ntypes * kTabStride);
  const size_t off_fb = kOffAb + 2 * tab;
  if (s) {
    s->pos = (float4*)(base + kOffPos);
    s->ab_row_off = (int*)(base + kOffRowOff);
    s->mask8 = (uint8_t*)(base + kOffMask8);
    s->diagmask = (uint32_t*)(base + kOffDiagMask);
    s->fi = (float*)(base + kOffFi);
    s->fj = (float*)(base + kOffFj);
    s->gstart = (uint16_t*)(base + kOffGstart);
    s->tasks = (uint32_t*)(base + kOffTasks);
    s->ered = (double*)(base + kOffEred);
    s->smd = (float*)(base + kOffSmd);
    s->smd_energy = (double*)(base + kOffSmdEnergy);
    s->box = (float*)(base + kOffBox);
    s->tick = (unsigned long long*)(base + kOffTick);
    s->pd = (double*)(base + kOffVel);
    s->fbi = (int*)(base + kOffFbi);
    s->vel = (float4*)(base + kOffVel);
    s->sx = (float*)(base + kOffSx);
    s->ta = (float*)(base + kOffAb);
    s->tb = (float*)(base + kOffAb + tab);
    s->fb = (float*)(base + off_fb);
    s->fbs = ((nslots > 0 ? nslots : 1) + 31) & ~31;
  }
  // 12 bytes per slot (three planes) instead of a float4: ALA10's 1 612 slots then leave room for 4 CTAs per SM
  return fixed ? off_fb : off_fb + 3 * sizeof(float) * (size_t)(((nslots > 0 ? nslots : 1) + 31) & ~31);
}

// ---- small math helpers ---------------------------------------------------------------------
__device__ __forceinline__ float fast_rsqrt(float x) {
  float y;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float fast_sqrt(float x) {
  float y;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float fast_rcp(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// rsqrt with one Newton step: relative error ~1 ulp (used where the value is raised to high powers)
__device__ __forceinline__ float rsqrt_nr(float x) {
  float y = fast_rsqrt(x);
  float t = x * y;
  float u = fmaf(-0.5f * t, y, 1.5f);
  return y * u;
}
// ---- packed FP32 (sm_100a FFMA2 / FADD2 / FMUL2): two lanes of work per issue slot ---------------------
// One instruction does two IEEE fp32 operations on a 64-bit register pair.  The FP32 pipe takes two cycles for
// it (same flops as two scalar instructions) but only ONE issue slot, which leaves room for the loads,
// predicates and MUFU of the pair loop (profiles/r01_microbench_ffma2.txt).
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk(u64 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 sub2(u64 a, u64 b) { u64 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ u64 dup2(float v) { return pk(v, v); }
__device__ __forceinline__ float hsum2(u64 v) { float a, b; upk(v, a, b); return __fadd_rn(a, b); }

// wrap_dist (forces.py:358-363): d - box * round(d / box), round-half-even through the 1.5 * 2^23 trick (|d / box| < 2^22).
// nb = -box, ib = 1 / box (0 for a zero box entry: that axis is left alone)
constexpr float kRoundMagic = 12582912.0f;
__device__ __forceinline__ u64 wrap2(u64 d, float nb, float ib) {
  const u64 m = dup2(kRoundMagic);
  return fma2(dup2(nb), sub2(add2(mul2(d, dup2(ib)), m), m), d);
}
__device__ __forceinline__ float wrap1(float d, float nb, float ib) {
  return __fmaf_rn(nb, __fsub_rn(__fadd_rn(__fmul_rn(d, ib), kRoundMagic), kRoundMagic), d);
}

struct f3 { float x, y, z; };
__device__ __forceinline__ f3 sub3(float4 a, float4 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ f3 cross3(f3 a, f3 b) {
  return {fmaf(a.y, b.z, -(a.z * b.y)), fmaf(a.z, b.x, -(a.x * b.z)), fmaf(a.x, b.y, -(a.y * b.x))};
}
__device__ __forceinline__ float dot3(f3 a, f3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }

// ---- Philox4x32-10 + Box-Muller (same stream as oracle_normal3) ---------------------------------
__device__ __forceinline__ void philox4x32_10(unsigned long long seed, uint32_t c0, uint32_t c1, uint32_t c2,
                                              uint32_t c3, uint32_t out[4]) {
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
__device__ __forceinline__ void normal3(unsigned long long seed, unsigned long long replica, uint32_t atom,
                                        unsigned long long step, float n[3]) {
  uint32_t u[4];
  philox4x32_10(seed, atom, (uint32_t)replica, (uint32_t)step, (uint32_t)(step >> 32), u);
  const float s = 1.0f / 16777216.0f;
  float u0 = ((float)(u[0] >> 8) + 0.5f) * s, u1 = ((float)(u[1] >> 8) + 0.5f) * s;
  float u2 = ((float)(u[2] >> 8) + 0.5f) * s, u3 = ((float)(u[3] >> 8) + 0.5f) * s;
  // Box-Muller on the SFU: r = sqrt(-2 ln u) through MUFU.LG2 / MUFU.SQRT, cos / sin(2 pi u) through MUFU.COS / MUFU.SIN
  // on the argument reduced to [-pi, pi) (cos(x - pi) = -cos x).  Absolute errors ~1e-6 on a unit-variance kick that is
  // scaled by vcoeff ~ 1e-2: far below the fp32 rounding of the velocities it is added to.
  const float two_pi = 6.283185307179586f;
  float r0 = fast_sqrt(-2.0f * __logf(u0)), r1 = fast_sqrt(-2.0f * __logf(u2));
  const float a1 = two_pi * (u1 - 0.5f), a3 = two_pi * (u3 - 0.5f);
  n[0] = -r0 * __cosf(a1); n[1] = -r0 * __sinf(a1); n[2] = -r1 * __cosf(a3);
}

}  // namespace b200md
