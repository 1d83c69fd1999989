// b200md_forces.cuh -- per-replica force evaluation (device functions shared by the single-shot and
// the fused persistent kernel).  Reference semantics: TorchMD_Fix/forces.py Forces.compute :82-344.
#pragma once
#include "b200md_kernels.cuh"

namespace b200md {

// Per-thread energy accumulators (FP64; FP32 partial sums are folded in once per pass / item).
struct Energies {
  double e[B200MD_NUM_OBSERVABLES];  // 8 energy slots + Ekin + end-to-end distance
  __device__ __forceinline__ void clear() {
#pragma unroll
    for (int k = 0; k < B200MD_NUM_OBSERVABLES; ++k) e[k] = 0.0;
  }
};

// ---- one non-bonded pair (forces.py:258-316 with evaluate_LJ :379-413, evaluate_electrostatics
// :451-489).  MODE 0: no cutoff, plain Coulomb.  MODE 1: cutoff with optional reaction field and
// LJ switching.  Returns s = -(dE/dr)/r so that F_i += s * d, F_j -= s * d with d = r_i - r_j.
//
// The force path is written with explicit round-to-nearest intrinsics so that the ENERGY and the
// force-only instantiations produce bit-identical forces (no compiler-chosen FMA contraction): a
// trajectory does not depend on how often energies are reported.  (a12, nb6) = (12 A, -6 B).
// Energies come from the same r^-1, r^-6 (a Newton step on rsqrt was tried and changed nothing measurable:
// the energy error is set by the fp32 rounding of dx, dy, dz, not by MUFU.RSQ).
template <int MODE, bool ENERGY>
__device__ __forceinline__ float pair_eval(float dx, float dy, float dz, float kqi, float qj, float a12, float nb6,
                                           bool excl, const ForceParams& P, float& pel, float& plj) {
  const float r2 = __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmul_rn(dx, dx)));
  float rinv = fast_rsqrt(r2);
  if (MODE == 1) excl = excl || !(r2 <= P.rc2);
  rinv = excl ? 0.0f : rinv;
  const float r2inv = __fmul_rn(rinv, rinv);
  const float r6 = __fmul_rn(__fmul_rn(r2inv, r2inv), r2inv);
  const float qq = __fmul_rn((MODE == 1 && excl) ? 0.0f : kqi, qj);
  const float glj1 = __fmaf_rn(a12, r6, nb6);  // 12 A r^-6 - 6 B;  glj1 * r6 = -r dE_lj/dr
  if (MODE == 0) {
    const float gc = __fmul_rn(qq, rinv);  // K qi qj / r = -r dE_c/dr
    if (ENERGY) {
      pel += gc;
      plj += __fmul_rn(__fmaf_rn(__fmul_rn(a12, 1.0f / 12.0f), r6, __fmul_rn(nb6, 1.0f / 6.0f)), r6);
    }
    return __fmul_rn(__fmaf_rn(glj1, r6, gc), r2inv);
  } else {
    const float gc = __fmul_rn(qq, __fmaf_rn(-2.0f * P.krf, r2, rinv));  // -r * forces.py:478
    float glj = __fmul_rn(glj1, r6);
    const float r = __fmul_rn(r2, rinv);
    // forces.py:401-411 (pot*deriv/dist as written).  t is clamped at 0 below the switch distance (and for
    // switch_dist = None, where rs = +inf makes it NaN): sv = 1, sd = 0 there, which leaves glj and elj untouched.
    const float t = fmaxf(__fmul_rn(r - P.rs, P.inv_sw), 0.0f);
    const float t2 = __fmul_rn(t, t);
    const float sv = __fmaf_rn(__fmul_rn(t2, t), __fmaf_rn(t, __fmaf_rn(t, -6.0f, 15.0f), -10.0f), 1.0f);
    const float nsd = __fmul_rn(__fmul_rn(t2, __fmaf_rn(t, __fmaf_rn(t, 30.0f, -60.0f), 30.0f)), P.inv_sw);
    float elj = __fmul_rn(__fmaf_rn(__fmul_rn(a12, 1.0f / 12.0f), r6, __fmul_rn(nb6, 1.0f / 6.0f)), r6);
    glj = __fmaf_rn(sv, glj, __fmul_rn(elj, nsd));
    elj = __fmul_rn(elj, sv);
    if (ENERGY) {
      pel += __fmul_rn(qq, __fadd_rn(rinv, __fmaf_rn(P.krf, r2, -P.crf)));  // forces.py:476
      plj += elj;
    }
    return __fmul_rn(__fadd_rn(glj, gc), r2inv);
  }
}

// Two pairs of one i atom (j, j+1) in packed f32x2: the same operations in the same order as pair_eval, one issue
// slot per two fp32 results.  Only rsqrt and the exclusion selects are per half.
struct PairConst {  // loop-invariant packed constants
  u64 c12, c6, m2krf, krf, ncrf, rs, inv_sw, cm6, c15, cm10, one, c30, cm60;
};
template <int MODE>
__device__ __forceinline__ PairConst make_pair_const(const ForceParams& P) {
  PairConst C;
  C.c12 = dup2(1.0f / 12.0f); C.c6 = dup2(1.0f / 6.0f);
  if (MODE == 1) {
    C.m2krf = dup2(-2.0f * P.krf); C.krf = dup2(P.krf); C.ncrf = dup2(-P.crf); C.rs = dup2(P.rs); C.inv_sw = dup2(P.inv_sw);
    C.cm6 = dup2(-6.0f); C.c15 = dup2(15.0f); C.cm10 = dup2(-10.0f); C.one = dup2(1.0f); C.c30 = dup2(30.0f); C.cm60 = dup2(-60.0f);
  }
  return C;
}
template <int MODE, bool ENERGY>
__device__ __forceinline__ u64 pair_eval2(u64 dx, u64 dy, u64 dz, float kqi, u64 kq2, u64 qj, u64 a12, u64 nb6, bool e0,
                                          bool e1, const ForceParams& P, const PairConst& C, u64& pel, u64& plj) {
  const u64 r2 = fma2(dz, dz, fma2(dy, dy, mul2(dx, dx)));
  float r2a, r2b;
  upk(r2, r2a, r2b);
  float ia = fast_rsqrt(r2a), ib = fast_rsqrt(r2b);
  if (MODE == 1) { e0 = e0 || !(r2a <= P.rc2); e1 = e1 || !(r2b <= P.rc2); }
  ia = e0 ? 0.0f : ia; ib = e1 ? 0.0f : ib;
  const u64 rinv = pk(ia, ib);
  const u64 r2inv = mul2(rinv, rinv);
  const u64 r6 = mul2(mul2(r2inv, r2inv), r2inv);
  const u64 qq = mul2(MODE == 1 ? pk(e0 ? 0.0f : kqi, e1 ? 0.0f : kqi) : kq2, qj);
  const u64 glj1 = fma2(a12, r6, nb6);
  if (MODE == 0) {
    const u64 gc = mul2(qq, rinv);
    if (ENERGY) {
      pel = add2(pel, gc);
      plj = add2(plj, mul2(fma2(mul2(a12, C.c12), r6, mul2(nb6, C.c6)), r6));
    }
    return mul2(fma2(glj1, r6, gc), r2inv);
  } else {
    const u64 gc = mul2(qq, fma2(C.m2krf, r2, rinv));
    u64 glj = mul2(glj1, r6);
    const u64 r = mul2(r2, rinv);
    float ta, tb;
    upk(mul2(sub2(r, C.rs), C.inv_sw), ta, tb);
    const u64 t = pk(fmaxf(ta, 0.0f), fmaxf(tb, 0.0f));
    const u64 t2 = mul2(t, t);
    const u64 sv = fma2(mul2(t2, t), fma2(t, fma2(t, C.cm6, C.c15), C.cm10), C.one);
    const u64 nsd = mul2(mul2(t2, fma2(t, fma2(t, C.c30, C.cm60), C.c30)), C.inv_sw);
    u64 elj = mul2(fma2(mul2(a12, C.c12), r6, mul2(nb6, C.c6)), r6);
    glj = fma2(sv, glj, mul2(elj, nsd));
    elj = mul2(elj, sv);
    if (ENERGY) {
      pel = add2(pel, mul2(qq, add2(rinv, fma2(C.krf, r2, C.ncrf))));
      plj = add2(plj, elj);
    }
    return mul2(add2(glj, gc), r2inv);
  }
}

// ---- TASK_TILE: j-chunk `chunk` (8 atoms) against i-blocks 0..passes-1 (32 lanes each) ----------
// Lane = i atom; the 8 j atoms arrive as 4 packed pairs (x[], y[], z[], q[] and the 12A / -6B rows are
// structure-of-arrays, so one LDS.128 delivers two f32x2 operands).  The pass loop is rolled (one body in the
// instruction cache); the i-side force of a pass is added to this warp's shared-memory partial.
template <int MODE, bool ENERGY>
__device__ __forceinline__ void tile_task(const Smem& S, const ForceParams& P, int chunk, int passes, int warp,
                                          int lane, Energies& E) {
  u64 fjx[kChunk / 2], fjy[kChunk / 2], fjz[kChunk / 2];
#pragma unroll
  for (int s = 0; s < kChunk / 2; ++s) fjx[s] = fjy[s] = fjz[s] = 0ull;
  const PairConst C = make_pair_const<MODE>(P);
  const float* sj = S.sx + chunk * kChunk;
  float* fiw = S.fi + warp * 3 * kMaxAtoms + lane;
  // software-pipelined pass prologue: the i atom, its LJ row offset and its 8 mask bits of pass p+1 are loaded
  // while pass p computes (their shared-memory latency was exposed at the top of every pass)
  float4 pi_n = S.pos[lane];
  int roff_n = S.ab_row_off[lane];
  uint32_t m8_n = S.mask8[chunk * kMaxAtoms + lane];
#pragma unroll 1
  for (int p = 0; p < passes; ++p) {
    asm volatile("" ::: "memory");  // re-read the 8 broadcast j atoms every pass instead of pinning 32 registers
    const float4 pi = pi_n;
    const float kqi = __fmul_rn(pi.w, kElecFactor);
    const u64 xi = dup2(pi.x), yi = dup2(pi.y), zi = dup2(pi.z), kq2 = dup2(kqi);
    const char* rowa = (const char*)S.ta + roff_n + chunk * kChunk * (int)sizeof(float);
    const char* rowb = (const char*)S.tb + roff_n + chunk * kChunk * (int)sizeof(float);
    const uint32_t m8 = m8_n;
    {
      const int in = (p + 1 < passes ? p + 1 : p) * 32 + lane;
      pi_n = S.pos[in];
      roff_n = S.ab_row_off[in];
      m8_n = S.mask8[chunk * kMaxAtoms + in];
    }
    u64 pel = 0ull, plj = 0ull, ax = 0ull, ay = 0ull, az = 0ull;
#pragma unroll
    for (int h = 0; h < kChunk / 4; ++h) {
      const ulonglong2 xj = *(const ulonglong2*)(sj + 4 * h), yj = *(const ulonglong2*)(sj + kMaxAtoms + 4 * h),
                       zj = *(const ulonglong2*)(sj + 2 * kMaxAtoms + 4 * h), qj = *(const ulonglong2*)(sj + 3 * kMaxAtoms + 4 * h);
      const ulonglong2 aj = *(const ulonglong2*)(rowa + 16 * h), bj = *(const ulonglong2*)(rowb + 16 * h);
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const int s = 2 * h + k;
        const u64 dx = sub2(xi, k ? xj.y : xj.x), dy = sub2(yi, k ? yj.y : yj.x), dz = sub2(zi, k ? zj.y : zj.x);
        const u64 sc = pair_eval2<MODE, ENERGY>(dx, dy, dz, kqi, kq2, k ? qj.y : qj.x, k ? aj.y : aj.x, k ? bj.y : bj.x,
                                                (m8 >> (2 * s)) & 1u, (m8 >> (2 * s + 1)) & 1u, P, C, pel, plj);
        ax = fma2(sc, dx, ax); ay = fma2(sc, dy, ay); az = fma2(sc, dz, az);
        fjx[s] = fma2(sc, dx, fjx[s]); fjy[s] = fma2(sc, dy, fjy[s]); fjz[s] = fma2(sc, dz, fjz[s]);
      }
    }
    float* f = fiw + p * 32;
    f[0] = __fadd_rn(f[0], hsum2(ax)); f[kMaxAtoms] = __fadd_rn(f[kMaxAtoms], hsum2(ay)); f[2 * kMaxAtoms] = __fadd_rn(f[2 * kMaxAtoms], hsum2(az));
    if (ENERGY) {
      float a, b;
      upk(pel, a, b); E.e[B200MD_E_ELECTROSTATICS] += (double)a + (double)b;
      upk(plj, a, b); E.e[B200MD_E_LJ] += (double)a + (double)b;
    }
  }
  float fj[kChunk][3];
#pragma unroll
  for (int s = 0; s < kChunk / 2; ++s) {
    upk(fjx[s], fj[2 * s][0], fj[2 * s + 1][0]);
    upk(fjy[s], fj[2 * s][1], fj[2 * s + 1][1]);
    upk(fjz[s], fj[2 * s][2], fj[2 * s + 1][2]);
  }
  // Transposing butterfly: 24 per-lane partials -> per-atom sums.  After the xor-16/8/4 stages lane l
  // holds atom (l >> 2) & 7 of the chunk; xor-2/1 finish the sum.  Fixed order => deterministic.
  const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4;
  float v4[4][3], v2[2][3], v1[3];
#pragma unroll
  for (int k = 0; k < 4; ++k)
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float send = b4 ? fj[k][c] : fj[k + 4][c];
      const float keep = b4 ? fj[k + 4][c] : fj[k][c];
      v4[k][c] = __fadd_rn(keep, __shfl_xor_sync(0xffffffffu, send, 16));
    }
#pragma unroll
  for (int k = 0; k < 2; ++k)
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float send = b3 ? v4[k][c] : v4[k + 2][c];
      const float keep = b3 ? v4[k + 2][c] : v4[k][c];
      v2[k][c] = __fadd_rn(keep, __shfl_xor_sync(0xffffffffu, send, 8));
    }
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float send = b2 ? v2[0][c] : v2[1][c];
    const float keep = b2 ? v2[1][c] : v2[0][c];
    float v = __fadd_rn(keep, __shfl_xor_sync(0xffffffffu, send, 4));
    v = __fadd_rn(v, __shfl_xor_sync(0xffffffffu, v, 2));
    v = __fadd_rn(v, __shfl_xor_sync(0xffffffffu, v, 1));
    v1[c] = v;
  }
  if ((lane & 3) == 0) {
    const int j = chunk * kChunk + (lane >> 2);
    S.fj[0 * kMaxAtoms + j] = -v1[0];
    S.fj[1 * kMaxAtoms + j] = -v1[1];
    S.fj[2 * kMaxAtoms + j] = -v1[2];
  }
}

// ---- TASK_DIAG: pairs inside one 32-atom block, full shell -----------------------------------------------
// Lane l owns atom 32*block + l and meets `ng` groups of 4 atoms of its own block as broadcast j atoms, exactly
// like a tile pass but WITHOUT a j side: every unordered pair is evaluated by both of its lanes (identical
// r^2, hence identical s and identical energy terms), so there is no partner force to hand back, no shuffle and
// no butterfly, and the whole step stays in packed FP32.  17 packed instructions per two pair evaluations against
// 26 scalar ones + 3 shuffles per pair of the rotation scheme it replaces (profiles/r01_v10: 46 issue slots per
// rotation step).  Energies are halved.  diagmask bit j: pair (lane, j) is not evaluated (self, exclusion, padding).
template <int MODE, bool ENERGY>
__device__ __forceinline__ void tile_diag_task(const Smem& S, const ForceParams& P, int block, int g0, int ng, int warp,
                                               int lane, Energies& E) {
  const int i0 = block * 32;
  const PairConst C = make_pair_const<MODE>(P);
  const float4 pi = S.pos[i0 + lane];
  const float kqi = __fmul_rn(pi.w, kElecFactor);
  const u64 xi = dup2(pi.x), yi = dup2(pi.y), zi = dup2(pi.z), kq2 = dup2(kqi);
  const int roff = S.ab_row_off[i0 + lane] + (i0 + 4 * g0) * (int)sizeof(float);
  const char* rowa = (const char*)S.ta + roff;
  const char* rowb = (const char*)S.tb + roff;
  const float* sj = S.sx + i0 + 4 * g0;
  uint32_t dm = S.diagmask[i0 + lane] >> (4 * g0);
  u64 pel = 0ull, plj = 0ull, ax = 0ull, ay = 0ull, az = 0ull;
#pragma unroll 1
  for (int g = 0; g < ng; ++g) {
    const ulonglong2 xj = *(const ulonglong2*)(sj + 4 * g), yj = *(const ulonglong2*)(sj + kMaxAtoms + 4 * g),
                     zj = *(const ulonglong2*)(sj + 2 * kMaxAtoms + 4 * g), qj = *(const ulonglong2*)(sj + 3 * kMaxAtoms + 4 * g);
    const ulonglong2 aj = *(const ulonglong2*)(rowa + 16 * g), bj = *(const ulonglong2*)(rowb + 16 * g);
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const u64 dx = sub2(xi, k ? xj.y : xj.x), dy = sub2(yi, k ? yj.y : yj.x), dz = sub2(zi, k ? zj.y : zj.x);
      const u64 sc = pair_eval2<MODE, ENERGY>(dx, dy, dz, kqi, kq2, k ? qj.y : qj.x, k ? aj.y : aj.x, k ? bj.y : bj.x,
                                              (dm >> (2 * k)) & 1u, (dm >> (2 * k + 1)) & 1u, P, C, pel, plj);
      ax = fma2(sc, dx, ax); ay = fma2(sc, dy, ay); az = fma2(sc, dz, az);
    }
    dm >>= 4;
  }
  float* f = S.fi + warp * 3 * kMaxAtoms + i0 + lane;
  f[0] = __fadd_rn(f[0], hsum2(ax)); f[kMaxAtoms] = __fadd_rn(f[kMaxAtoms], hsum2(ay)); f[2 * kMaxAtoms] = __fadd_rn(f[2 * kMaxAtoms], hsum2(az));
  if (ENERGY) {
    float a, b;
    upk(pel, a, b); E.e[B200MD_E_ELECTROSTATICS] += 0.5 * ((double)a + (double)b);
    upk(plj, a, b); E.e[B200MD_E_LJ] += 0.5 * ((double)a + (double)b);
  }
}

// ---- bonded items: where a force vector goes ---------------------------------------------------------
// FIXED = false (single-shot Forces.compute): private atom-major 12-byte slots, summed per atom in a fixed
//   order after the barrier -- exact fp32 sums for arbitrarily large forces (generated / clashing frames).
// FIXED = true (fused integrator): 2^-17 kcal/mol/A fixed point, one shared-memory integer atomic per
//   component.  Integer addition is associative, so the sum is still order-independent and bit-reproducible,
//   and the slot buffer, its stores and the gather loop disappear.  A contribution above kFixedLimit raises the
//   per-replica range flag (the run has exploded long before 16 384 kcal/mol/A are reached on one atom).
constexpr float kFixedScale = 131072.0f, kFixedInv = 1.0f / 131072.0f, kFixedLimit = 2048.0f;

template <bool FIXED>
__device__ __forceinline__ void emit(const Smem& S, uint32_t slot, uint32_t atom, float x, float y, float z,
                                     uint32_t& flag) {
  if (FIXED) {
    if (fmaxf(fabsf(x), fmaxf(fabsf(y), fabsf(z))) > kFixedLimit) flag |= B200MD_FLAG_FORCE_RANGE;
    atomicAdd(S.fbi + atom, __float2int_rn(x * kFixedScale));
    atomicAdd(S.fbi + kMaxAtoms + atom, __float2int_rn(y * kFixedScale));
    atomicAdd(S.fbi + 2 * kMaxAtoms + atom, __float2int_rn(z * kFixedScale));
  } else {
    float* p = S.fb + 3 * slot;
    p[0] = x; p[1] = y; p[2] = z;
  }
}

// evaluate_bonds, forces.py:111-132, 492-501
template <bool ENERGY, bool FIXED>
__device__ __forceinline__ void bond_task(const Smem& S, const DevTopo& T, const ForceParams& P, int item, Energies& E,
                                          uint32_t& flag) {
  if (item >= T.nbonds) return;
  const uint32_t q = __ldg(T.bond_idx + item), sl = FIXED ? 0u : __ldg(T.bond_slot + item);
  const float2 prm = __ldg(T.bond_prm + item);
  const float4 pa = S.pos[q & 0xff], pb = S.pos[(q >> 8) & 0xff];
  const f3 d = sub3(pa, pb);
  const float r2 = dot3(d, d);
  const float rinv = fast_rsqrt(r2), dist = r2 * rinv;
  const bool keep = !P.has_cutoff || dist <= P.rc;  // forces.py:117-125
  const float x = dist - prm.y;
  const float c = keep ? 2.0f * prm.x * x * rinv : 0.0f;  // dE/dr / r
  emit<FIXED>(S, sl & 0xffff, q & 0xff, -c * d.x, -c * d.y, -c * d.z, flag);
  emit<FIXED>(S, sl >> 16, (q >> 8) & 0xff, c * d.x, c * d.y, c * d.z, flag);
  if (ENERGY && keep) {
    // k (r - r0)^2 cancels ~40:1 in (r - r0): the energy (report steps only) takes r in FP64
    const double ddx = (double)pa.x - (double)pb.x, ddy = (double)pa.y - (double)pb.y, ddz = (double)pa.z - (double)pb.z;
    const double xd = sqrt(ddx * ddx + ddy * ddy + ddz * ddz) - (double)prm.y;
    E.e[B200MD_E_BONDS] += (double)prm.x * xd * xd;
  }
}

// evaluate_angles, forces.py:134-145, 504-537
template <bool ENERGY, bool FIXED>
__device__ __forceinline__ void angle_task(const Smem& S, const DevTopo& T, int item, Energies& E, uint32_t& flag) {
  if (item >= T.nangles) return;
  const uint32_t q = __ldg(T.angle_idx + item);
  const uint2 sl = FIXED ? make_uint2(0u, 0u) : __ldg(T.angle_slot + item);
  const float2 prm = __ldg(T.angle_prm + item);
  const float4 p0 = S.pos[q & 0xff], p1 = S.pos[(q >> 8) & 0xff], p2 = S.pos[(q >> 16) & 0xff];
  const f3 r21 = sub3(p0, p1), r23 = sub3(p2, p1);
  const float n21i = fast_rsqrt(dot3(r21, r21)), n23i = fast_rsqrt(dot3(r23, r23));
  float ct = dot3(r23, r21) * n21i * n23i;
  ct = fminf(fmaxf(ct, -1.0f), 1.0f);
  const float th = acosf(ct), dth = th - prm.y;
  if (ENERGY) E.e[B200MD_E_ANGLES] += (double)(prm.x * dth * dth);
  const float st2 = fmaf(-ct, ct, 1.0f);
  const float coef = (st2 > 0.0f) ? -2.0f * prm.x * dth * fast_rsqrt(st2) : 0.0f;  // 0 where sin(theta) == 0
  const float a0 = coef * n21i, a2 = coef * n23i;
  const float c21 = ct * n21i, c23 = ct * n23i;
  const float f0x = a0 * fmaf(c21, r21.x, -r23.x * n23i), f0y = a0 * fmaf(c21, r21.y, -r23.y * n23i),
              f0z = a0 * fmaf(c21, r21.z, -r23.z * n23i);
  const float f2x = a2 * fmaf(c23, r23.x, -r21.x * n21i), f2y = a2 * fmaf(c23, r23.y, -r21.y * n21i),
              f2z = a2 * fmaf(c23, r23.z, -r21.z * n21i);
  emit<FIXED>(S, sl.x & 0xffff, q & 0xff, f0x, f0y, f0z, flag);
  emit<FIXED>(S, sl.x >> 16, (q >> 8) & 0xff, -(f0x + f2x), -(f0y + f2y), -(f0z + f2z), flag);
  emit<FIXED>(S, sl.y & 0xffff, (q >> 16) & 0xff, f2x, f2y, f2z, flag);
}

// One LJ + Coulomb pair with its own (12 A', 6 B', K qi qj') parameters: forces.py:176-226 (1-4 pairs are
// never cut, switched or reaction-field corrected).  Returns s with F_a += s d, F_b -= s d, d = r_a - r_b.
template <bool ENERGY>
__device__ __forceinline__ float pair14(f3 d, float4 prm, const ForceParams& P, Energies& E) {
  const float rinv = rsqrt_nr(dot3(d, d));
  const float r2inv = rinv * rinv, r6 = r2inv * r2inv * r2inv;
  const float a12 = prm.x * P.lj_on, b6 = prm.y * P.lj_on, qq = prm.z * P.el_on;
  const float ec = qq * rinv;
  const float glj = fmaf(a12, r6, -b6) * r6;
  if (ENERGY) {
    E.e[B200MD_E_LJ] += (double)(fmaf(a12 * (1.0f / 12.0f), r6, -b6 * (1.0f / 6.0f)) * r6);
    E.e[B200MD_E_ELECTROSTATICS] += (double)ec;
  }
  return (glj + ec) * r2inv;
}

// evaluate_torsion, forces.py:147-174 / 228-255, 540-602 (dihedrals and impropers share it), with the
// 1-4 pair on the end atoms (forces.py:176-226) fused into the dihedral item that owns it.
// phi = -atan2(sinPhi, cosPhi) is never formed for AMBER terms with integer periodicity: cos/sin(n phi)
// come from (cos phi + i sin phi)^n and the phase enters through its precomputed cosine / sine.
template <bool ENERGY, bool FIXED>
__device__ __forceinline__ void torsion_task(const Smem& S, const ForceParams& P, bool torsion_on, bool fuse14,
                                             int item, int count, const uint32_t* idx, const uint2* terms,
                                             const float4* term_prm, const uint2* slots, const float4* p14,
                                             double& Eacc, Energies& E, uint32_t& flag) {
  if (item >= count) return;
  const uint32_t q = __ldg(idx + item);
  const uint2 sl = FIXED ? make_uint2(0u, 0u) : __ldg(slots + item);
  const float4 p0 = S.pos[q & 0xff], p1 = S.pos[(q >> 8) & 0xff], p2 = S.pos[(q >> 16) & 0xff], p3 = S.pos[q >> 24];
  float f0x = 0.f, f0y = 0.f, f0z = 0.f, f1x = 0.f, f1y = 0.f, f1z = 0.f, f2x = 0.f, f2y = 0.f, f2z = 0.f,
        f3x = 0.f, f3y = 0.f, f3z = 0.f;
  if (torsion_on) {
    const f3 r12 = sub3(p0, p1), r23 = sub3(p1, p2), r34 = sub3(p2, p3);
    const f3 cA = cross3(r12, r23), cB = cross3(r23, r34);
    const float nA2 = dot3(cA, cA), nB2 = dot3(cB, cB), n23sq = dot3(r23, r23);
    const float rnA = fast_rsqrt(nA2), rnB = fast_rsqrt(nB2), rn23 = fast_rsqrt(n23sq), n23 = n23sq * rn23;
    // cosPhi = A.B / (|A||B|);  sinPhi = (C.B)/(|C||B|) with C = r23 x A  =>  |r23| (r12.B) / (|A||B|)
    const float rab = rnA * rnB;
    const float cosPhi = dot3(cA, cB) * rab, sinPhi = n23 * dot3(r12, cB) * rab;
    const float hinv = fast_rsqrt(fmaf(cosPhi, cosPhi, sinPhi * sinPhi));
    const float c = cosPhi * hinv, s = -sinPhi * hinv;  // cos / sin of phi = -atan2(sinPhi, cosPhi)
    const uint2 tr = __ldg(terms + item);
    float pot = 0.0f, coeff = 0.0f;
    for (uint32_t t = tr.x; t < tr.x + tr.y; ++t) {
      const float4 tp = __ldg(term_prm + t);  // k0, per | marker, c0, s0
      if (tp.y > 0.0f) {                      // AMBER, integer periodicity (forces.py:565-569)
        float cn = c, sn = s;
        for (float k = 1.5f; k < tp.y; k += 1.0f) {
          const float t0 = fmaf(cn, c, -sn * s);
          sn = fmaf(cn, s, sn * c);
          cn = t0;
        }
        pot += tp.x * (1.0f + fmaf(cn, tp.z, sn * tp.w));      // k (1 + cos(n phi - phi0))
        coeff -= tp.y * tp.x * fmaf(sn, tp.z, -cn * tp.w);     // -n k sin(n phi - phi0)
      } else {
        const float phi = atan2f(s, c);
        if (tp.y == kTermHarmonic) {          // CHARMM harmonic (forces.py:570-576)
          float ad = phi - tp.z;
          const float pi_f = 3.14159265358979323846f;
          if (ad < -pi_f) ad += 2.0f * pi_f; else if (ad > pi_f) ad -= 2.0f * pi_f;
          pot += tp.x * ad * ad;
          coeff += 2.0f * tp.x * ad;
        } else {                              // AMBER with a non-integer periodicity: general path
          float sn, cs;
          sincosf(fmaf(tp.w, phi, -tp.z), &sn, &cs);
          pot += tp.x * (1.0f + cs);
          coeff -= tp.w * tp.x * sn;
        }
      }
    }
    if (ENERGY) Eacc += (double)pot;
    // OpenMM-style projection (forces.py:582-600)
    const float ff0 = -coeff * n23 * (rnA * rnA), ff3 = coeff * n23 * (rnB * rnB);
    const float rn23sq = rn23 * rn23;
    const float ff1 = dot3(r12, r23) * rn23sq, ff2 = dot3(r34, r23) * rn23sq;
    const float ax = ff0 * cA.x, ay = ff0 * cA.y, az = ff0 * cA.z;
    const float bx = ff3 * cB.x, by = ff3 * cB.y, bz = ff3 * cB.z;
    const float sx = fmaf(ff1, ax, -ff2 * bx), sy = fmaf(ff1, ay, -ff2 * by), sz = fmaf(ff1, az, -ff2 * bz);
    f0x = -ax; f0y = -ay; f0z = -az;
    f1x = ax + sx; f1y = ay + sy; f1z = az + sz;
    f2x = bx - sx; f2y = by - sy; f2z = bz - sz;
    f3x = -bx; f3y = -by; f3z = -bz;
  }
  if (fuse14) {
    const float4 prm = __ldg(p14 + item);
    if (prm.w != 0.0f) {
      const f3 d = sub3(p0, p3);
      const float sc = pair14<ENERGY>(d, prm, P, E);
      f0x = fmaf(sc, d.x, f0x); f0y = fmaf(sc, d.y, f0y); f0z = fmaf(sc, d.z, f0z);
      f3x = fmaf(-sc, d.x, f3x); f3y = fmaf(-sc, d.y, f3y); f3z = fmaf(-sc, d.z, f3z);
    }
  }
  emit<FIXED>(S, sl.x & 0xffff, q & 0xff, f0x, f0y, f0z, flag);
  emit<FIXED>(S, sl.x >> 16, (q >> 8) & 0xff, f1x, f1y, f1z, flag);
  emit<FIXED>(S, sl.y & 0xffff, (q >> 16) & 0xff, f2x, f2y, f2z, flag);
  emit<FIXED>(S, sl.y >> 16, q >> 24, f3x, f3y, f3z, flag);
}

// 1-4 pairs that no dihedral item owns (only possible with hand-made parameter objects)
template <bool ENERGY, bool FIXED>
__device__ __forceinline__ void p14_task(const Smem& S, const DevTopo& T, const ForceParams& P, int item, Energies& E,
                                         uint32_t& flag) {
  if (item >= T.n14) return;
  const uint32_t q = __ldg(T.p14_idx + item), sl = __ldg(T.p14_slot + item);
  const float4 prm = __ldg(T.p14_prm + item);
  const f3 d = sub3(S.pos[q & 0xff], S.pos[(q >> 8) & 0xff]);
  const float sc = pair14<ENERGY>(d, prm, P, E);
  emit<FIXED>(S, sl & 0xffff, q & 0xff, sc * d.x, sc * d.y, sc * d.z, flag);
  emit<FIXED>(S, sl >> 16, (q >> 8) & 0xff, -sc * d.x, -sc * d.y, -sc * d.z, flag);
}

// ---- all tasks of this warp for one force evaluation; positions must be in S.pos ------------------
// The warp's static task list is sorted by kind on the host (T.kend[warp][k] = end of the k-th run), so every kind
// is a tight loop around its own body: no switch, and the next task word is fetched while the current task runs
// (the fetch -> decode -> branch chain of a switch-per-task loop cost 33 issue slots and ~150 cycles of exposed
// latency per task, profiles/r01_v11).
template <int MODE, bool ENERGY, bool FIXED>
__device__ __forceinline__ void run_tasks(const Smem& S, const DevTopo& T, const ForceParams& P, int warp, int lane,
                                          Energies& E, uint32_t& flag) {
  const uint32_t* tk = S.tasks + warp * kMaxTasks;
  int t = 0;
  uint32_t next = tk[0];
#define B200MD_TASK_LOOP(K, BODY)                                                     \
  {                                                                                   \
    const int end = T.kend[warp][K];                                                  \
    _Pragma("unroll 1") for (; t < end;) {                                            \
      const uint32_t task = next;                                                     \
      next = tk[++t];                                                                 \
      const uint32_t chunk = (task >> 8) & 0xffff, passes = task >> 24;               \
      const int item = chunk * 32 + lane;                                             \
      (void)passes; (void)item;                                                       \
      BODY;                                                                           \
    }                                                                                 \
  }
  B200MD_TASK_LOOP(0, (tile_task<MODE, ENERGY>(S, P, chunk, passes, warp, lane, E)))
  B200MD_TASK_LOOP(1, (tile_diag_task<MODE, ENERGY>(S, P, chunk & 0xff, (chunk >> 8) & 0xff, passes, warp, lane, E)))
  B200MD_TASK_LOOP(2, (torsion_task<ENERGY, FIXED>(S, P, T.dih_on != 0, T.p14_on != 0, item, T.ndih, T.dih_idx, T.dih_terms,
                                                   T.dih_term_prm, T.dih_slot, T.dih_p14, E.e[B200MD_E_DIHEDRALS], E, flag)))
  B200MD_TASK_LOOP(3, (torsion_task<ENERGY, FIXED>(S, P, true, false, item, T.nimp, T.imp_idx, T.imp_terms, T.imp_term_prm,
                                                   T.imp_slot, nullptr, E.e[B200MD_E_IMPROPERS], E, flag)))
  B200MD_TASK_LOOP(4, (angle_task<ENERGY, FIXED>(S, T, item, E, flag)))
  B200MD_TASK_LOOP(5, (bond_task<ENERGY, FIXED>(S, T, P, item, E, flag)))
  B200MD_TASK_LOOP(6, (p14_task<ENERGY, FIXED>(S, T, P, item, E, flag)))
#undef B200MD_TASK_LOOP
}

// ---- SMD end-to-end bias (potcalc.py:75-89; force = -dE/dr): one lane, during the task phase -------
__device__ __forceinline__ void smd_task(const Smem& S, const ForceParams& P, float center) {
  // one lane per replica: FP64 throughout, (c - d)^2 cancels when the restraint is nearly satisfied
  const float4 p1 = S.pos[P.smd_a1], p2 = S.pos[P.smd_a2];
  const double dx = (double)p2.x - (double)p1.x, dy = (double)p2.y - (double)p1.y, dz = (double)p2.z - (double)p1.z;
  const double dist = sqrt(dx * dx + dy * dy + dz * dz);
  const double dev = (double)center - dist;
  const double c = (double)P.smd_fk * dev / dist;  // force on atom2 = +fk (c - d) unit(r2 - r1)
  S.smd[0] = (float)(c * dx); S.smd[1] = (float)(c * dy); S.smd[2] = (float)(c * dz);
  S.smd[4] = (float)dist;
  S.smd_energy[0] = 0.5 * (double)P.smd_fk * dev * dev;
}

// ---- per-atom gather in a fixed order: j-side slot, 4 i-side partials, bonded slots, SMD ------------
template <bool FIXED>
__device__ __forceinline__ void gather_forces(const Smem& S, const ForceParams& P, int a, float& fx, float& fy,
                                              float& fz) {
  fx = S.fj[a]; fy = S.fj[kMaxAtoms + a]; fz = S.fj[2 * kMaxAtoms + a];
#pragma unroll
  for (int w = 0; w < kWarps; ++w) {
    fx += S.fi[(w * 3 + 0) * kMaxAtoms + a];
    fy += S.fi[(w * 3 + 1) * kMaxAtoms + a];
    fz += S.fi[(w * 3 + 2) * kMaxAtoms + a];
    S.fi[(w * 3 + 0) * kMaxAtoms + a] = 0.0f; S.fi[(w * 3 + 1) * kMaxAtoms + a] = 0.0f; S.fi[(w * 3 + 2) * kMaxAtoms + a] = 0.0f;  // re-arm
  }
  if (FIXED) {  // bonded forces arrived as fixed-point atomics; read and re-arm the accumulators
    fx += (float)S.fbi[a] * kFixedInv; fy += (float)S.fbi[kMaxAtoms + a] * kFixedInv; fz += (float)S.fbi[2 * kMaxAtoms + a] * kFixedInv;
    S.fbi[a] = 0; S.fbi[kMaxAtoms + a] = 0; S.fbi[2 * kMaxAtoms + a] = 0;
  } else {
    const int g1 = S.gstart[a + 1];
#pragma unroll 4
    for (int g = S.gstart[a]; g < g1; ++g) {
      const float* v = S.fb + 3 * g;
      fx += v[0]; fy += v[1]; fz += v[2];
    }
  }
  if (P.smd_on) {
    if (a == P.smd_a2) { fx += S.smd[0]; fy += S.smd[1]; fz += S.smd[2]; }
    if (a == P.smd_a1) { fx -= S.smd[0]; fy -= S.smd[1]; fz -= S.smd[2]; }
  }
}

// ---- block reduction of the FP64 accumulators (fixed order); thread 0 ends with the totals ---------
__device__ __forceinline__ void reduce_energies(const Smem& S, Energies& E, int warp, int lane) {
#pragma unroll
  for (int k = 0; k < B200MD_NUM_OBSERVABLES; ++k) {
    double v = E.e[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) S.ered[warp * B200MD_NUM_OBSERVABLES + k] = v;
  }
  __syncthreads();
  if (warp == 0 && lane == 0) {
#pragma unroll
    for (int k = 0; k < B200MD_NUM_OBSERVABLES; ++k)
      E.e[k] = (S.ered[k] + S.ered[B200MD_NUM_OBSERVABLES + k]) +
               (S.ered[2 * B200MD_NUM_OBSERVABLES + k] + S.ered[3 * B200MD_NUM_OBSERVABLES + k]);
  }
}

}  // namespace b200md
