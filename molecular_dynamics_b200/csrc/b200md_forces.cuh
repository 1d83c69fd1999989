// b200md_forces.cuh -- per-replica force evaluation (device functions shared by the single-shot and
// the fused persistent kernel).  Reference semantics: TorchMD_Fix/forces.py Forces.compute :82-344.
#pragma once
#include "b200md_kernels.cuh"

namespace b200md {

// Per-thread energy accumulators of the FP64 energy pass.
struct Energies {
  double e[kNumReduced];  // 8 energy slots + Ekin + end-to-end distance
  __device__ __forceinline__ void clear() {
#pragma unroll
    for (int k = 0; k < kNumReduced; ++k) e[k] = 0.0;
  }
};

// ---- one non-bonded pair (forces.py:258-316 with evaluate_LJ :379-413, evaluate_electrostatics
// :451-489).  MODE 0: no cutoff, plain Coulomb.  MODE 1: cutoff with optional reaction field and
// LJ switching.  Returns s = -(dE/dr)/r so that F_i += s * d, F_j -= s * d with d = r_i - r_j.
//
// The force path is written with explicit round-to-nearest intrinsics so that the results do not depend on
// compiler-chosen FMA contraction (every instantiation rounds identically).  (a12, nb6) = (12 A, -6 B).  Energies are not
// computed here: the fp32 rounding of dx, dy, dz limits a pair energy to ~1e-7 relative, and the lj / electrostatics
// sums cancel up to 1000:1 -- they come from the FP64 energy pass below.
// Two pairs of one i atom (j, j+1) in packed f32x2: one issue slot per two fp32 results.  Only rsqrt and the exclusion
// selects are per half.
struct PairConst {  // loop-invariant packed constants (MODE 1)
  u64 m2krf, rs, inv_sw, cm6, c15, cm10, one, nsdc;
};
template <int MODE>
__device__ __forceinline__ PairConst make_pair_const(const ForceParams& P) {
  PairConst C;
  if (MODE == 1) {
    C.m2krf = dup2(-2.0f * P.krf); C.rs = dup2(P.rs); C.inv_sw = dup2(P.inv_sw);
    C.cm6 = dup2(-6.0f); C.c15 = dup2(15.0f); C.cm10 = dup2(-10.0f); C.one = dup2(1.0f);
    C.nsdc = dup2(30.0f / 12.0f * P.inv_sw);  // -s'(t) = 30 t^2 (1 - t)^2 / (rc - rs), with the 1/12 of 12 E_lj folded in
  }
  return C;
}
template <int MODE>
__device__ __forceinline__ u64 pair_eval2(u64 dx, u64 dy, u64 dz, float kqi, u64 kq2, u64 qj, u64 a12, u64 nb6, bool e0,
                                          bool e1, const ForceParams& P, const PairConst& C) {
  const u64 r2 = fma2(dz, dz, fma2(dy, dy, mul2(dx, dx)));
  float r2a, r2b;
  upk(r2, r2a, r2b);
  float ia = fast_rsqrt(r2a), ib = fast_rsqrt(r2b);
  if (MODE == 1) { e0 = e0 || !(r2a <= P.rc2); e1 = e1 || !(r2b <= P.rc2); }
  ia = e0 ? 0.0f : ia; ib = e1 ? 0.0f : ib;
  const u64 rinv = pk(ia, ib);
  const u64 r2inv = mul2(rinv, rinv);
  const u64 r6 = mul2(mul2(r2inv, r2inv), r2inv);
  // An excluded / cut-off pair has rinv = 0: r2inv = 0 multiplies the whole coefficient below, so its (finite) reaction-field
  // term needs no select of its own
  const u64 qq = mul2(kq2, qj);
  const u64 glj1 = fma2(a12, r6, nb6);
  if (MODE == 0) {
    const u64 gc = mul2(qq, rinv);
    return mul2(fma2(glj1, r6, gc), r2inv);
  } else {
    const u64 gc = mul2(qq, fma2(C.m2krf, r2, rinv));  // -r * forces.py:478
    const u64 glj = mul2(glj1, r6);
    // forces.py:401-411 (pot * deriv / dist as written): F <- s F + E s' / r with t = (r - rs) / (rc - rs) clamped at 0 below
    // the switch distance (and for switch_dist = None, where rs = +inf makes it NaN): s = 1, s' = 0 there.
    //   s(t) = 1 + t^3 (-10 + 15 t - 6 t^2),   -s'(t) (rc - rs) = 30 t^2 (1 - t)^2 = 30 (t - t^2)^2,   12 E_lj = glj + nb6 r6
    const u64 r = mul2(r2, rinv);
    float ta, tb;
    upk(mul2(sub2(r, C.rs), C.inv_sw), ta, tb);
    const u64 t = pk(fmaxf(ta, 0.0f), fmaxf(tb, 0.0f));
    const u64 t2 = mul2(t, t);
    const u64 sv = fma2(mul2(t2, t), fma2(t, fma2(t, C.cm6, C.c15), C.cm10), C.one);
    const u64 q = sub2(t, t2);
    const u64 nsd = mul2(mul2(q, q), C.nsdc);
    const u64 elj12 = fma2(nb6, r6, glj);
    return mul2(add2(fma2(sv, glj, mul2(elj12, nsd)), gc), r2inv);
  }
}

// ---- TASK_TILE: j-chunk `chunk` (8 atoms) against i-blocks 0..passes-1 (32 lanes each) ----------
// Lane = i atom; the 8 j atoms arrive as 4 packed pairs (x[], y[], z[], q[] and the 12A / -6B rows are
// structure-of-arrays, so one LDS.128 delivers two f32x2 operands).  The pass loop is rolled (one body in the
// instruction cache); the i-side force of a pass is added to this warp's shared-memory partial.
template <int MODE, bool PBC, bool SKIP>
__device__ __forceinline__ void tile_task(const Smem& S, const ForceParams& P, int chunk, int passes, int warp,
                                          int lane) {
  const float4 nbox = PBC ? *(const float4*)S.box : make_float4(0.f, 0.f, 0.f, 0.f);
  const float4 ibox = PBC ? *(const float4*)(S.box + 4) : make_float4(0.f, 0.f, 0.f, 0.f);
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
    if (MODE == 1 && SKIP) {
      // Cutoff mode: a pass whose 32 x 8 pairs are ALL beyond the cutoff contributes exactly zero (pair_eval2 zeroes every
      // term of such a pair), so it is skipped on a warp vote -- bit-identical results.  The r^2 below are the expressions
      // pair_eval2 evaluates (same operands, same order): the compiler merges them, the test costs the compares and a vote.
      // For the extended deca-alanine conformers of the benchmark (end-to-end 31 A, cutoff 9 A) 6 of the 15 passes drop out.
      // (SKIP = false: the rigid-bond instantiation, where the test's live registers would spill.)
      bool in = false;
#pragma unroll
      for (int h = 0; h < kChunk / 4; ++h) {
        const ulonglong2 xj = *(const ulonglong2*)(sj + 4 * h), yj = *(const ulonglong2*)(sj + kMaxAtoms + 4 * h),
                         zj = *(const ulonglong2*)(sj + 2 * kMaxAtoms + 4 * h);
#pragma unroll
        for (int k = 0; k < 2; ++k) {
          u64 dx = sub2(xi, k ? xj.y : xj.x), dy = sub2(yi, k ? yj.y : yj.x), dz = sub2(zi, k ? zj.y : zj.x);
          if (PBC) { dx = wrap2(dx, nbox.x, ibox.x); dy = wrap2(dy, nbox.y, ibox.y); dz = wrap2(dz, nbox.z, ibox.z); }
          float ra, rb;
          upk(fma2(dz, dz, fma2(dy, dy, mul2(dx, dx))), ra, rb);
          in = in || ra <= P.rc2 || rb <= P.rc2;
        }
      }
      if (!__any_sync(0xffffffffu, in)) continue;
    }
    u64 ax = 0ull, ay = 0ull, az = 0ull;
#pragma unroll
    for (int h = 0; h < kChunk / 4; ++h) {
      const ulonglong2 xj = *(const ulonglong2*)(sj + 4 * h), yj = *(const ulonglong2*)(sj + kMaxAtoms + 4 * h),
                       zj = *(const ulonglong2*)(sj + 2 * kMaxAtoms + 4 * h), qj = *(const ulonglong2*)(sj + 3 * kMaxAtoms + 4 * h);
      const ulonglong2 aj = *(const ulonglong2*)(rowa + 16 * h), bj = *(const ulonglong2*)(rowb + 16 * h);
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const int s = 2 * h + k;
        u64 dx = sub2(xi, k ? xj.y : xj.x), dy = sub2(yi, k ? yj.y : yj.x), dz = sub2(zi, k ? zj.y : zj.x);
        if (PBC) { dx = wrap2(dx, nbox.x, ibox.x); dy = wrap2(dy, nbox.y, ibox.y); dz = wrap2(dz, nbox.z, ibox.z); }
        const u64 sc = pair_eval2<MODE>(dx, dy, dz, kqi, kq2, k ? qj.y : qj.x, k ? aj.y : aj.x, k ? bj.y : bj.x,
                                        (m8 >> (2 * s)) & 1u, (m8 >> (2 * s + 1)) & 1u, P, C);
        ax = fma2(sc, dx, ax); ay = fma2(sc, dy, ay); az = fma2(sc, dz, az);
        fjx[s] = fma2(sc, dx, fjx[s]); fjy[s] = fma2(sc, dy, fjy[s]); fjz[s] = fma2(sc, dz, fjz[s]);
      }
    }
    float* f = fiw + p * 32;
    f[0] = __fadd_rn(f[0], hsum2(ax)); f[kMaxAtoms] = __fadd_rn(f[kMaxAtoms], hsum2(ay)); f[2 * kMaxAtoms] = __fadd_rn(f[2 * kMaxAtoms], hsum2(az));
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
// r^2, hence identical s), so there is no partner force to hand back, no shuffle and
// no butterfly, and the whole step stays in packed FP32.  17 packed instructions per two pair evaluations against
// 26 scalar ones + 3 shuffles per pair of the rotation scheme it replaces (profiles/r01_v10: 46 issue slots per
// rotation step).  diagmask bit j: pair (lane, j) is not evaluated (self, exclusion, padding).
template <int MODE, bool PBC>
__device__ __forceinline__ void tile_diag_task(const Smem& S, const ForceParams& P, int block, int g0, int ng, int warp,
                                               int lane) {
  const float4 nbox = PBC ? *(const float4*)S.box : make_float4(0.f, 0.f, 0.f, 0.f);
  const float4 ibox = PBC ? *(const float4*)(S.box + 4) : make_float4(0.f, 0.f, 0.f, 0.f);
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
  u64 ax = 0ull, ay = 0ull, az = 0ull;
  // two groups (4 packed pair steps) per iteration: four independent dependency chains to interleave.  An odd `ng`
  // only occurs at the ragged end of a block, where the extra group is padding (all of its mask bits are set).
#pragma unroll 1
  for (int g2 = 0; g2 < ng; g2 += 2) {
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int g = g2 + u;
      const ulonglong2 xj = *(const ulonglong2*)(sj + 4 * g), yj = *(const ulonglong2*)(sj + kMaxAtoms + 4 * g),
                       zj = *(const ulonglong2*)(sj + 2 * kMaxAtoms + 4 * g), qj = *(const ulonglong2*)(sj + 3 * kMaxAtoms + 4 * g);
      const ulonglong2 aj = *(const ulonglong2*)(rowa + 16 * g), bj = *(const ulonglong2*)(rowb + 16 * g);
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        u64 dx = sub2(xi, k ? xj.y : xj.x), dy = sub2(yi, k ? yj.y : yj.x), dz = sub2(zi, k ? zj.y : zj.x);
        if (PBC) { dx = wrap2(dx, nbox.x, ibox.x); dy = wrap2(dy, nbox.y, ibox.y); dz = wrap2(dz, nbox.z, ibox.z); }
        const u64 sc = pair_eval2<MODE>(dx, dy, dz, kqi, kq2, k ? qj.y : qj.x, k ? aj.y : aj.x, k ? bj.y : bj.x,
                                        (dm >> (4 * u + 2 * k)) & 1u, (dm >> (4 * u + 2 * k + 1)) & 1u, P, C);
        ax = fma2(sc, dx, ax); ay = fma2(sc, dy, ay); az = fma2(sc, dz, az);
      }
    }
    dm >>= 8;
  }
  float* f = S.fi + warp * 3 * kMaxAtoms + i0 + lane;
  f[0] = __fadd_rn(f[0], hsum2(ax)); f[kMaxAtoms] = __fadd_rn(f[kMaxAtoms], hsum2(ay)); f[2 * kMaxAtoms] = __fadd_rn(f[2 * kMaxAtoms], hsum2(az));
}

// ---- bonded items: where a force vector goes ---------------------------------------------------------
// FIXED = false (single-shot Forces.compute): private atom-major 16-byte slots, summed per atom in a fixed
//   order after the barrier -- exact fp32 sums for arbitrarily large forces (generated / clashing frames).
// FIXED = true (fused integrator): 2^-17 kcal/mol/A fixed point, one shared-memory integer atomic per
//   component.  Integer addition is associative, so the sum is still order-independent and bit-reproducible,
//   and the slot buffer, its stores and the gather loop disappear.  The items scale their force coefficient by
//   kFixedScale themselves (one multiply per item instead of one per component); the per-atom gather raises the
//   range flag when an accumulator passes 2^29 (4 096 kcal/mol/A: the run has exploded long before).
constexpr float kFixedScale = 131072.0f, kFixedInv = 1.0f / 131072.0f;
constexpr int kFixedLimit = 1 << 29;

template <bool FIXED>
__device__ __forceinline__ void emit(const Smem& S, uint32_t slot, uint32_t atom, float x, float y, float z) {
  if (FIXED) {
    atomicAdd(S.fbi + atom, __float2int_rn(x));
    atomicAdd(S.fbi + kMaxAtoms + atom, __float2int_rn(y));
    atomicAdd(S.fbi + 2 * kMaxAtoms + atom, __float2int_rn(z));
  } else {
    S.fb[slot] = x; S.fb[S.fbs + slot] = y; S.fb[2 * S.fbs + slot] = z;
  }
}
// +v to (slot0, atom0) and -v to (slot1, atom1) from one conversion
template <bool FIXED>
__device__ __forceinline__ void emit_pm(const Smem& S, uint32_t slot0, uint32_t atom0, uint32_t slot1, uint32_t atom1,
                                        float x, float y, float z) {
  if (FIXED) {
    const int ix = __float2int_rn(x), iy = __float2int_rn(y), iz = __float2int_rn(z);
    atomicAdd(S.fbi + atom0, ix); atomicAdd(S.fbi + kMaxAtoms + atom0, iy); atomicAdd(S.fbi + 2 * kMaxAtoms + atom0, iz);
    atomicAdd(S.fbi + atom1, -ix); atomicAdd(S.fbi + kMaxAtoms + atom1, -iy); atomicAdd(S.fbi + 2 * kMaxAtoms + atom1, -iz);
  } else {
    S.fb[slot0] = x; S.fb[S.fbs + slot0] = y; S.fb[2 * S.fbs + slot0] = z;
    S.fb[slot1] = -x; S.fb[S.fbs + slot1] = -y; S.fb[2 * S.fbs + slot1] = -z;
  }
}

// ---- packed three-vectors: component pairs (item A, item B) ------------------------------------------
struct v3p { u64 x, y, z; };
__device__ __forceinline__ v3p sub3p(v3p a, v3p b) { return {sub2(a.x, b.x), sub2(a.y, b.y), sub2(a.z, b.z)}; }
// calculate_distances (forces.py:366-370): a - b, minimum image when the replica has a box
template <bool PBC>
__device__ __forceinline__ v3p diff3p(const Smem& S, v3p a, v3p b) {
  v3p d = sub3p(a, b);
  if (PBC) {
    const float4 nbox = *(const float4*)S.box, ibox = *(const float4*)(S.box + 4);
    d.x = wrap2(d.x, nbox.x, ibox.x); d.y = wrap2(d.y, nbox.y, ibox.y); d.z = wrap2(d.z, nbox.z, ibox.z);
  }
  return d;
}
__device__ __forceinline__ u64 dot3p(v3p a, v3p b) { return fma2(a.z, b.z, fma2(a.y, b.y, mul2(a.x, b.x))); }
__device__ __forceinline__ u64 neg2(u64 a) { return a ^ 0x8000000080000000ull; }
__device__ __forceinline__ v3p cross3p(v3p a, v3p b) {
  return {fma2(a.y, b.z, neg2(mul2(a.z, b.y))), fma2(a.z, b.x, neg2(mul2(a.x, b.z))), fma2(a.x, b.y, neg2(mul2(a.y, b.x)))};
}
__device__ __forceinline__ v3p scale3p(u64 s, v3p a) { return {mul2(s, a.x), mul2(s, a.y), mul2(s, a.z)}; }
__device__ __forceinline__ u64 rsqrt2(u64 v) { float a, b; upk(v, a, b); return pk(fast_rsqrt(a), fast_rsqrt(b)); }
// positions of atom `ia` (item A) and `ib` (item B) from the structure-of-arrays copy
__device__ __forceinline__ v3p load3p(const Smem& S, uint32_t ia, uint32_t ib) {
  return {pk(S.sx[ia], S.sx[ib]), pk(S.sx[kMaxAtoms + ia], S.sx[kMaxAtoms + ib]), pk(S.sx[2 * kMaxAtoms + ia], S.sx[2 * kMaxAtoms + ib])};
}
template <bool FIXED>
__device__ __forceinline__ void emit2(const Smem& S, uint32_t slot_a, uint32_t slot_b, uint32_t atom_a, uint32_t atom_b, v3p f) {
  float xa, xb, ya, yb, za, zb;
  upk(f.x, xa, xb); upk(f.y, ya, yb); upk(f.z, za, zb);
  emit<FIXED>(S, slot_a, atom_a, xa, ya, za);
  emit<FIXED>(S, slot_b, atom_b, xb, yb, zb);
}
template <bool FIXED>
__device__ __forceinline__ void emit2_pm(const Smem& S, uint32_t sa0, uint32_t sa1, uint32_t sb0, uint32_t sb1, uint32_t a0,
                                         uint32_t a1, uint32_t b0, uint32_t b1, v3p f) {
  float xa, xb, ya, yb, za, zb;
  upk(f.x, xa, xb); upk(f.y, ya, yb); upk(f.z, za, zb);
  emit_pm<FIXED>(S, sa0, a0, sa1, a1, xa, ya, za);
  emit_pm<FIXED>(S, sb0, b0, sb1, b1, xb, yb, zb);
}

// evaluate_bonds, forces.py:111-132, 492-501 -- two bonds per lane
template <bool FIXED, bool PBC>
__device__ __forceinline__ void bond_task(const Smem& S, const DevTopo& T, const ForceParams& P, int pair) {
  if (pair >= T.nbond2) return;
  const uint2 q = __ldg(T.bond2_idx + pair);
  const float4 prm = __ldg(T.bond2_prm + pair);  // kA, kB, r0A, r0B
  const uint4 sl = FIXED ? make_uint4(0u, 0u, 0u, 0u) : __ldg(T.bond2_slot + pair);
  const uint32_t a0 = q.x & 0xff, a1 = (q.x >> 8) & 0xff, b0 = q.y & 0xff, b1 = (q.y >> 8) & 0xff;
  const v3p d = diff3p<PBC>(S, load3p(S, a0, b0), load3p(S, a1, b1));
  const u64 r2 = dot3p(d, d);
  const u64 rinv = rsqrt2(r2), dist = mul2(r2, rinv);
  float da, db;
  upk(dist, da, db);
  const bool keep_a = !P.has_cutoff || da <= P.rc, keep_b = !P.has_cutoff || db <= P.rc;  // forces.py:117-125
  const float fs = FIXED ? -2.0f * kFixedScale : -2.0f;
  const u64 k2 = pk(keep_a ? fs * prm.x : 0.0f, keep_b ? fs * prm.y : 0.0f);
  const u64 c = mul2(mul2(k2, sub2(dist, pk(prm.z, prm.w))), rinv);  // -(dE/dr) / r (scaled)
  emit2_pm<FIXED>(S, sl.x & 0xffff, sl.x >> 16, sl.z & 0xffff, sl.z >> 16, a0, a1, b0, b1, scale3p(c, d));
}

// acos on [-1, 1] without branches: sqrt(1 - |x|) * P7(|x|) (Abramowitz & Stegun 4.4.46, |error| <= 2e-8 + fp32
// rounding), reflected for x < 0.  Used for the FORCE of the angle term; the energy pass takes acos in FP64.
__device__ __forceinline__ u64 acos2(u64 x) {
  float xa, xb;
  upk(x, xa, xb);
  const u64 t = pk(fabsf(xa), fabsf(xb));
  u64 p = dup2(-0.0012624911f);
  p = fma2(p, t, dup2(0.0066700901f)); p = fma2(p, t, dup2(-0.0170881256f)); p = fma2(p, t, dup2(0.0308918810f));
  p = fma2(p, t, dup2(-0.0501743046f)); p = fma2(p, t, dup2(0.0889789874f)); p = fma2(p, t, dup2(-0.2145988016f));
  p = fma2(p, t, dup2(1.5707963050f));
  float ua, ub;
  upk(sub2(dup2(1.0f), t), ua, ub);
  const u64 r = mul2(pk(fast_sqrt(ua), fast_sqrt(ub)), p);
  // x < 0: pi - r (pi split in two floats)
  const u64 sg = pk(xa < 0.0f ? -1.0f : 1.0f, xb < 0.0f ? -1.0f : 1.0f);
  const u64 hi = pk(xa < 0.0f ? 3.14159274101257324f : 0.0f, xb < 0.0f ? 3.14159274101257324f : 0.0f);
  const u64 lo = pk(xa < 0.0f ? -8.74227766e-8f : 0.0f, xb < 0.0f ? -8.74227766e-8f : 0.0f);
  return add2(fma2(sg, r, hi), lo);
}

// evaluate_angles, forces.py:134-145, 504-537 -- two angles per lane
template <bool FIXED, bool PBC>
__device__ __forceinline__ void angle_task(const Smem& S, const DevTopo& T, int pair) {
  if (pair >= T.nangle2) return;
  const uint2 q = __ldg(T.angle2_idx + pair);
  const float4 prm = __ldg(T.angle2_prm + pair);  // kA, kB, theta0A, theta0B
  const uint4 sl = FIXED ? make_uint4(0u, 0u, 0u, 0u) : __ldg(T.angle2_slot + pair);
  const uint32_t a0 = q.x & 0xff, a1 = (q.x >> 8) & 0xff, a2 = (q.x >> 16) & 0xff;
  const uint32_t b0 = q.y & 0xff, b1 = (q.y >> 8) & 0xff, b2 = (q.y >> 16) & 0xff;
  const v3p p1 = load3p(S, a1, b1);
  const v3p r21 = diff3p<PBC>(S, load3p(S, a0, b0), p1), r23 = diff3p<PBC>(S, load3p(S, a2, b2), p1);
  const u64 n21i = rsqrt2(dot3p(r21, r21)), n23i = rsqrt2(dot3p(r23, r23));
  float cta, ctb;
  upk(mul2(mul2(dot3p(r23, r21), n21i), n23i), cta, ctb);
  cta = fminf(fmaxf(cta, -1.0f), 1.0f); ctb = fminf(fmaxf(ctb, -1.0f), 1.0f);
  const u64 ct = pk(cta, ctb);
  const u64 dth = sub2(acos2(ct), pk(prm.z, prm.w));
  float sa, sb;
  upk(fma2(neg2(ct), ct, dup2(1.0f)), sa, sb);
  const float fs = FIXED ? -2.0f * kFixedScale : -2.0f;
  const u64 rs = pk(sa > 0.0f ? fs * fast_rsqrt(sa) : 0.0f, sb > 0.0f ? fs * fast_rsqrt(sb) : 0.0f);  // 0 where sin(theta) == 0
  const u64 coef = mul2(mul2(pk(prm.x, prm.y), dth), rs);
  const u64 c0 = mul2(coef, n21i), c2 = mul2(coef, n23i), c21 = mul2(ct, n21i), c23 = mul2(ct, n23i);
  const v3p u23 = scale3p(n23i, r23), u21 = scale3p(n21i, r21);
  const v3p f0 = {mul2(c0, fma2(c21, r21.x, neg2(u23.x))), mul2(c0, fma2(c21, r21.y, neg2(u23.y))), mul2(c0, fma2(c21, r21.z, neg2(u23.z)))};
  const v3p f2 = {mul2(c2, fma2(c23, r23.x, neg2(u21.x))), mul2(c2, fma2(c23, r23.y, neg2(u21.y))), mul2(c2, fma2(c23, r23.z, neg2(u21.z)))};
  const v3p f1 = {neg2(add2(f0.x, f2.x)), neg2(add2(f0.y, f2.y)), neg2(add2(f0.z, f2.z))};
  emit2<FIXED>(S, sl.x & 0xffff, sl.z & 0xffff, a0, b0, f0);
  emit2<FIXED>(S, sl.x >> 16, sl.z >> 16, a1, b1, f1);
  emit2<FIXED>(S, sl.y & 0xffff, sl.w & 0xffff, a2, b2, f2);
}

// Two LJ + Coulomb pairs with their own (12 A', -6 B', K qi qj') parameters: forces.py:176-226 (1-4 pairs are
// never cut, switched or reaction-field corrected).  Returns s with F_a += s d, F_b -= s d, d = r_a - r_b.
__device__ __forceinline__ u64 pair14_2(v3p d, u64 a12, u64 nb6, u64 qq, const ForceParams& P) {
  const u64 r2 = dot3p(d, d);
  u64 rinv = rsqrt2(r2);
  rinv = mul2(rinv, fma2(mul2(mul2(r2, rinv), dup2(-0.5f)), rinv, dup2(1.5f)));  // Newton step: the value is raised to r^-12
  const u64 r2inv = mul2(rinv, rinv), r6 = mul2(mul2(r2inv, r2inv), r2inv);
  a12 = mul2(a12, dup2(P.lj_on)); nb6 = mul2(nb6, dup2(P.lj_on)); qq = mul2(qq, dup2(P.el_on));
  const u64 ec = mul2(qq, rinv);
  const u64 glj = mul2(fma2(a12, r6, nb6), r6);
  return mul2(add2(glj, ec), r2inv);
}

// Proper torsions with integer periodicity, two per lane, as a Fourier series in phi (forces.py:147-174, 540-602):
//   E = K0 + sum_n a_n cos(n phi) + b_n sin(n phi),  a_n = sum k cos(phi0), b_n = sum k sin(phi0) over the terms of
// periodicity n, so any number of terms costs the same and there is no branch: cos / sin(n phi) come from
// (cos phi + i sin phi)^n, phi = -atan2(sinPhi, cosPhi) is never formed.  `nmax` (largest periodicity of the chunk)
// is warp-uniform.  The 1-4 pair on the end atoms (forces.py:176-226) is fused in (zero parameters if there is none).
template <bool FIXED, bool PBC>
__device__ __forceinline__ void dihedral_task(const Smem& S, const DevTopo& T, const ForceParams& P, int pair, int nmax) {
  if (pair >= T.ndih2) return;
  const uint2 q = __ldg(T.dih2_idx + pair);
  const uint4 sl = FIXED ? make_uint4(0u, 0u, 0u, 0u) : __ldg(T.dih2_slot + pair);
  const uint32_t a0 = q.x & 0xff, a1 = (q.x >> 8) & 0xff, a2 = (q.x >> 16) & 0xff, a3 = q.x >> 24;
  const uint32_t b0 = q.y & 0xff, b1 = (q.y >> 8) & 0xff, b2 = (q.y >> 16) & 0xff, b3 = q.y >> 24;
  const v3p p0 = load3p(S, a0, b0), p1 = load3p(S, a1, b1), p2 = load3p(S, a2, b2), p3 = load3p(S, a3, b3);
  const u64 fs = dup2(FIXED ? kFixedScale : 1.0f);
  v3p f0 = {0ull, 0ull, 0ull}, f1 = f0, f2 = f0, f3 = f0;
  if (T.dih_on) {
    const v3p r12 = diff3p<PBC>(S, p0, p1), r23 = diff3p<PBC>(S, p1, p2), r34 = diff3p<PBC>(S, p2, p3);
    const v3p cA = cross3p(r12, r23), cB = cross3p(r23, r34);
    const u64 n23sq = dot3p(r23, r23);
    const u64 rnA = rsqrt2(dot3p(cA, cA)), rnB = rsqrt2(dot3p(cB, cB)), rn23 = rsqrt2(n23sq), n23 = mul2(n23sq, rn23);
    // cosPhi = A.B / (|A||B|);  sinPhi = (C.B)/(|C||B|) with C = r23 x A  =>  |r23| (r12.B) / (|A||B|)
    const u64 rab = mul2(rnA, rnB);
    const u64 cosPhi = mul2(dot3p(cA, cB), rab), sinPhi = mul2(mul2(n23, dot3p(r12, cB)), rab);
    const u64 hinv = rsqrt2(fma2(cosPhi, cosPhi, mul2(sinPhi, sinPhi)));
    const u64 c = mul2(cosPhi, hinv), s = neg2(mul2(sinPhi, hinv));  // cos / sin of phi = -atan2(sinPhi, cosPhi)
    u64 cn = c, sn = s, coeff = 0ull;
    const float4* cf = T.dih2_coef + pair;
#pragma unroll 1
    for (int n = 1; n <= nmax; ++n) {
      const float4 k4 = __ldg(cf);  // n a_n (A, B), n b_n (A, B)
      cf += T.dih2_stride;
      const u64 ca = pk(k4.x, k4.y), cb = pk(k4.z, k4.w);
      coeff = fma2(cb, cn, fma2(neg2(ca), sn, coeff));  // dE/dphi += n (b_n cos(n phi) - a_n sin(n phi))
      const u64 t0 = fma2(cn, c, neg2(mul2(sn, s)));
      sn = fma2(cn, s, mul2(sn, c));
      cn = t0;
    }
    // OpenMM-style projection (forces.py:582-600), coefficient pre-scaled for the fixed-point accumulators
    coeff = mul2(mul2(coeff, fs), n23);
    const u64 g0 = mul2(coeff, mul2(rnA, rnA)), g3 = mul2(coeff, mul2(rnB, rnB));  // f0 = g0 cA, f3 = -g3 cB
    const u64 rn23sq = mul2(rn23, rn23);
    const u64 ff1 = mul2(dot3p(r12, r23), rn23sq), ff2 = mul2(dot3p(r34, r23), rn23sq);
    f0 = scale3p(g0, cA);
    const v3p b = scale3p(g3, cB);  // = -f3
    // s = ff1 a - ff2 b with a = -f0:  f1 = a + s, f2 = b - s
    const v3p sv = {neg2(fma2(ff1, f0.x, mul2(ff2, b.x))), neg2(fma2(ff1, f0.y, mul2(ff2, b.y))), neg2(fma2(ff1, f0.z, mul2(ff2, b.z)))};
    f1 = {sub2(sv.x, f0.x), sub2(sv.y, f0.y), sub2(sv.z, f0.z)};
    f2 = {sub2(b.x, sv.x), sub2(b.y, sv.y), sub2(b.z, sv.z)};
    f3 = {neg2(b.x), neg2(b.y), neg2(b.z)};
  }
  if (T.p14_on) {
    const float4 ab = __ldg(T.dih2_p14ab + pair);
    const float2 qq = __ldg(T.dih2_p14q + pair);
    const v3p d = diff3p<PBC>(S, p0, p3);
    const u64 sc = mul2(pair14_2(d, pk(ab.x, ab.y), pk(ab.z, ab.w), pk(qq.x, qq.y), P), fs);
    f0 = {fma2(sc, d.x, f0.x), fma2(sc, d.y, f0.y), fma2(sc, d.z, f0.z)};
    f3 = {fma2(neg2(sc), d.x, f3.x), fma2(neg2(sc), d.y, f3.y), fma2(neg2(sc), d.z, f3.z)};
  }
  emit2<FIXED>(S, sl.x & 0xffff, sl.z & 0xffff, a0, b0, f0);
  emit2<FIXED>(S, sl.x >> 16, sl.z >> 16, a1, b1, f1);
  emit2<FIXED>(S, sl.y & 0xffff, sl.w & 0xffff, a2, b2, f2);
  emit2<FIXED>(S, sl.y >> 16, sl.w >> 16, a3, b3, f3);
}

// 1-4 pairs that no packed dihedral owns (their dihedral has k = 0 everywhere, or is a general torsion)
template <bool FIXED, bool PBC>
__device__ __forceinline__ void p14_task(const Smem& S, const DevTopo& T, const ForceParams& P, int pair) {
  if (pair >= T.n14_2) return;
  const uint2 q = __ldg(T.p142_idx + pair);
  const uint4 sl = FIXED ? make_uint4(0u, 0u, 0u, 0u) : __ldg(T.p142_slot + pair);
  const float4 ab = __ldg(T.p142_ab + pair);
  const float2 qq = __ldg(T.p142_q + pair);
  const uint32_t a0 = q.x & 0xff, a1 = (q.x >> 8) & 0xff, b0 = q.y & 0xff, b1 = (q.y >> 8) & 0xff;
  const v3p d = diff3p<PBC>(S, load3p(S, a0, b0), load3p(S, a1, b1));
  const u64 sc = mul2(pair14_2(d, pk(ab.x, ab.y), pk(ab.z, ab.w), pk(qq.x, qq.y), P), dup2(FIXED ? kFixedScale : 1.0f));
  emit2_pm<FIXED>(S, sl.x & 0xffff, sl.x >> 16, sl.z & 0xffff, sl.z >> 16, a0, a1, b0, b1, scale3p(sc, d));
}

// General torsion, one item per lane: evaluate_torsion, forces.py:147-174 / 228-255, 540-602, for what the packed
// series cannot express -- impropers (either branch), CHARMM harmonic slices, non-integer or > 6 periodicities.
// Items [0, nspec) are special dihedrals (energy slot "dihedrals"), the rest impropers.
template <bool FIXED, bool PBC>
__device__ __forceinline__ void torsion_task(const Smem& S, const DevTopo& T, int item) {
  if (item < T.gen_begin || item >= T.gen_end) return;
  const uint32_t q = __ldg(T.gen_idx + item);
  const uint2 sl = FIXED ? make_uint2(0u, 0u) : __ldg(T.gen_slot + item);
  const float4 p0 = S.pos[q & 0xff], p1 = S.pos[(q >> 8) & 0xff], p2 = S.pos[(q >> 16) & 0xff], p3 = S.pos[q >> 24];
  f3 r12 = sub3(p0, p1), r23 = sub3(p1, p2), r34 = sub3(p2, p3);
  if (PBC) {
    const float4 nb = *(const float4*)S.box, ib = *(const float4*)(S.box + 4);
    r12 = {wrap1(r12.x, nb.x, ib.x), wrap1(r12.y, nb.y, ib.y), wrap1(r12.z, nb.z, ib.z)};
    r23 = {wrap1(r23.x, nb.x, ib.x), wrap1(r23.y, nb.y, ib.y), wrap1(r23.z, nb.z, ib.z)};
    r34 = {wrap1(r34.x, nb.x, ib.x), wrap1(r34.y, nb.y, ib.y), wrap1(r34.z, nb.z, ib.z)};
  }
  const f3 cA = cross3(r12, r23), cB = cross3(r23, r34);
  const float nA2 = dot3(cA, cA), nB2 = dot3(cB, cB), n23sq = dot3(r23, r23);
  const float rnA = fast_rsqrt(nA2), rnB = fast_rsqrt(nB2), rn23 = fast_rsqrt(n23sq), n23 = n23sq * rn23;
  // cosPhi = A.B / (|A||B|);  sinPhi = (C.B)/(|C||B|) with C = r23 x A  =>  |r23| (r12.B) / (|A||B|)
  const float rab = rnA * rnB;
  const float cosPhi = dot3(cA, cB) * rab, sinPhi = n23 * dot3(r12, cB) * rab;
  const float hinv = fast_rsqrt(fmaf(cosPhi, cosPhi, sinPhi * sinPhi));
  const float c = cosPhi * hinv, s = -sinPhi * hinv;  // cos / sin of phi = -atan2(sinPhi, cosPhi)
  const uint2 tr = __ldg(T.gen_terms + item);
  float coeff = 0.0f;
  for (uint32_t t = tr.x; t < tr.x + tr.y; ++t) {
    const float4 tp = __ldg(T.gen_term_prm + t);  // k0, per | marker, c0, s0
    if (tp.y > 0.0f) {                            // AMBER, integer periodicity (forces.py:565-569)
      float cn = c, sn = s;
      for (float k = 1.5f; k < tp.y; k += 1.0f) {
        const float t0 = fmaf(cn, c, -sn * s);
        sn = fmaf(cn, s, sn * c);
        cn = t0;
      }
      coeff -= tp.y * tp.x * fmaf(sn, tp.z, -cn * tp.w);     // -n k sin(n phi - phi0)
    } else {
      const float phi = atan2f(s, c);
      if (tp.y == kTermHarmonic) {          // CHARMM harmonic (forces.py:570-576)
        float ad = phi - tp.z;
        const float pi_f = 3.14159265358979323846f;
        if (ad < -pi_f) ad += 2.0f * pi_f; else if (ad > pi_f) ad -= 2.0f * pi_f;
        coeff += 2.0f * tp.x * ad;
      } else {                              // AMBER with a non-integer periodicity: general path
        coeff -= tp.w * tp.x * sinf(fmaf(tp.w, phi, -tp.z));  // E = k (1 + cos(per phi - phi0))
      }
    }
  }
  if (FIXED) coeff *= kFixedScale;
  // OpenMM-style projection (forces.py:582-600)
  const float ff0 = -coeff * n23 * (rnA * rnA), ff3 = coeff * n23 * (rnB * rnB);
  const float rn23sq = rn23 * rn23;
  const float ff1 = dot3(r12, r23) * rn23sq, ff2 = dot3(r34, r23) * rn23sq;
  const float ax = ff0 * cA.x, ay = ff0 * cA.y, az = ff0 * cA.z;
  const float bx = ff3 * cB.x, by = ff3 * cB.y, bz = ff3 * cB.z;
  const float sx = fmaf(ff1, ax, -ff2 * bx), sy = fmaf(ff1, ay, -ff2 * by), sz = fmaf(ff1, az, -ff2 * bz);
  emit<FIXED>(S, sl.x & 0xffff, q & 0xff, -ax, -ay, -az);
  emit<FIXED>(S, sl.x >> 16, (q >> 8) & 0xff, ax + sx, ay + sy, az + sz);
  emit<FIXED>(S, sl.y & 0xffff, (q >> 16) & 0xff, bx - sx, by - sy, bz - sz);
  emit<FIXED>(S, sl.y >> 16, q >> 24, -bx, -by, -bz);
}

// =====================================================================================================
// FP64 energy pass: every term of Forces.compute (forces.py:82-344) evaluated in double from the fp32 positions,
// the arithmetic of the reference run in torch.double on the same fp32-rounded inputs (the contract is
// 1e-6 relative PER TERM; lj / electrostatics cancel up to 1000:1, so fp32 pair terms cannot meet it).  Forces never
// come from here.  Item lists in topology order, one item per thread, strided by the CTA; per-thread FP64
// accumulators are block-reduced in a fixed order afterwards (reduce_energies).  Positions are read from S.pd
// (x[128], y[128], z[128], q[128] in double, written by the caller from its registers).
// =====================================================================================================
constexpr double kElecFactorD = 332.06371307417066;  // forces.py:373-376 with the reference's scipy constants

// 1 / sqrt(x) in double for x of ordinary magnitude (squared distances, squared norms of cross products): the hardware
// seed (MUFU.RSQ64H through rsqrt.approx.ftz.f64, ~2^-20 relative) and ONE third-order correction
// y (1 + e / 2 + 3 e^2 / 8), e = 1 - x y^2, which leaves e^3 ~ 1e-18: 5 FP64 instructions on one dependency chain against
// the ~15 of the library's rsqrt() with its special-case handling.  The energy pass is bound by the FP64 pipe
// (64 DFMA / clock / SM), so this chain length is what a pair costs.
__device__ __forceinline__ double drsqrt(double x) {
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  const double e = fma(-x * y, y, 1.0);
  return fma(y, e * fma(0.375, e, 0.5), y);
}

// Shared-memory double load by 32-bit shared address: inside the (not inlined) energy pass the compiler cannot prove
// that S.pd points to shared memory and would emit generic LD instead of LDS.
__device__ __forceinline__ double lds_f64(uint32_t saddr, int idx) {
  double v;
  asm("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(saddr + 8u * (uint32_t)idx));
  return v;
}
struct EPairConst {  // loop invariants of the pair loop, in registers
  uint32_t pd;       // shared address of S.pd
  double ta, tb, rs, inv_sw, rc2, krf, ncrf, bx, by, bz;
};

struct dv3 { double x, y, z; };
// calculate_distances (forces.py:366-370) in double; `pbc` is uniform over the CTA (one replica), so a run-time flag costs
// a predictable branch instead of a second copy of the FP64 code
__device__ __forceinline__ dv3 ediff(const EPairConst& C, bool pbc, int i, int j) {
  dv3 d = {lds_f64(C.pd, i) - lds_f64(C.pd, j), lds_f64(C.pd, kMaxAtoms + i) - lds_f64(C.pd, kMaxAtoms + j),
           lds_f64(C.pd, 2 * kMaxAtoms + i) - lds_f64(C.pd, 2 * kMaxAtoms + j)};
  if (pbc) {  // wrap_dist, forces.py:358-363; a zero box entry leaves its axis alone
    if (C.bx != 0.0) d.x -= C.bx * rint(d.x / C.bx);
    if (C.by != 0.0) d.y -= C.by * rint(d.y / C.by);
    if (C.bz != 0.0) d.z -= C.bz * rint(d.z / C.bz);
  }
  return d;
}
__device__ __forceinline__ double edot(dv3 a, dv3 b) { return fma(a.z, b.z, fma(a.y, b.y, a.x * b.x)); }
__device__ __forceinline__ dv3 ecross(dv3 a, dv3 b) {
  return {fma(a.y, b.z, -(a.z * b.y)), fma(a.z, b.x, -(a.x * b.z)), fma(a.x, b.y, -(a.y * b.x))};
}

// evaluate_torsion (forces.py:540-576), energy only.  Integer-periodicity AMBER terms never form phi: cos / sin(n phi)
// come from (cos phi + i sin phi)^n, cos / sin(phi0) from the host in FP64.
__device__ __forceinline__ double etorsion(const EPairConst& C, const DevTopo& T, bool pbc, int d) {
  const uint32_t q = __ldg(T.et_idx + d);
  const int a0 = q & 0xff, a1 = (q >> 8) & 0xff, a2 = (q >> 16) & 0xff, a3 = q >> 24;
  const dv3 r12 = ediff(C, pbc, a0, a1), r23 = ediff(C, pbc, a1, a2), r34 = ediff(C, pbc, a2, a3);
  const dv3 cA = ecross(r12, r23), cB = ecross(r23, r34);
  // forces.py:548-551: cosPhi = A.B / (|A||B|), sinPhi = C.B / (|C||B|) with C = r23 x A.  r23 is perpendicular to A, so
  // |C| = |r23||A| and C.B = |r23|^2 (r12.B): sinPhi = |r23| (r12.B) / (|A||B|) -- the same numbers to 1e-16 without the
  // third cross product, its norm and a renormalisation
  const double n23sq = edot(r23, r23), iab = drsqrt(edot(cA, cA)) * drsqrt(edot(cB, cB));
  const double cosPhi = edot(cA, cB) * iab, sinPhi = n23sq * drsqrt(n23sq) * edot(r12, cB) * iab;
  const double c1 = cosPhi, s1 = -sinPhi;  // cos / sin of phi = -atan2(sinPhi, cosPhi)
  double phi = 0.0, pot = 0.0;
  bool have_phi = false;
  const int t1 = __ldg(T.et_start + d + 1);
  for (int t = __ldg(T.et_start + d); t < t1; ++t) {
    const float4 tp = __ldg(T.et_term + t);
    const double k0 = tp.x;
    if (tp.w == 2.0f) {  // k (1 + cos(n phi - phi0)), forces.py:565-567
      const double2 cs0 = __ldg(T.et_cs + t);
      double cn = c1, sn = s1;
      for (int k = 1; k < (int)tp.z; ++k) {
        const double cc = fma(cn, c1, -(sn * s1));
        sn = fma(sn, c1, cn * s1);
        cn = cc;
      }
      pot += k0 * (1.0 + fma(cn, cs0.x, sn * cs0.y));
      continue;
    }
    if (!have_phi) { phi = -atan2(sinPhi, cosPhi); have_phi = true; }
    if (tp.w == 0.0f) {
      pot += k0 * (1.0 + cos((double)tp.z * phi - (double)tp.y));
    } else {  // CHARMM harmonic, forces.py:570-574
      double ad = phi - (double)tp.y;
      if (ad < -M_PI) ad += 2.0 * M_PI; else if (ad > M_PI) ad -= 2.0 * M_PI;
      pot += k0 * ad * ad;
    }
  }
  return pot;
}

// One non-bonded pair of Forces.ava_idx in double (forces.py:379-413, 451-489): (electrostatics, lj-family) energies.
// Straight-line (selects, no branches): the caller unrolls four of these so that their dependency chains overlap.
// S.pd carries q sqrt(K): the charge product below is K qi qj.  LJ = true: the plain "lj" term (A r^-12 - B r^-6) without
// the two table scalings the repulsion / combination cases need.
template <int MODE, bool PBC, bool LJ>
__device__ __forceinline__ void epair(const EPairConst& C, const double2* __restrict__ tab, uint32_t w, bool live, double& eel,
                                      double& elj) {
  const int i = w & 0xff, j = (w >> 8) & 0xff;
  double dx = lds_f64(C.pd, i) - lds_f64(C.pd, j), dy = lds_f64(C.pd, kMaxAtoms + i) - lds_f64(C.pd, kMaxAtoms + j),
         dz = lds_f64(C.pd, 2 * kMaxAtoms + i) - lds_f64(C.pd, 2 * kMaxAtoms + j);
  if (PBC) {  // wrap_dist, forces.py:358-363; a zero box entry leaves its axis alone
    if (C.bx != 0.0) dx -= C.bx * rint(dx / C.bx);
    if (C.by != 0.0) dy -= C.by * rint(dy / C.by);
    if (C.bz != 0.0) dz -= C.bz * rint(dz / C.bz);
  }
  const double r2 = fma(dz, dz, fma(dy, dy, dx * dx));
  const double2 ab = __ldg(tab + (w >> 16));
  const double rinv = drsqrt(r2), r2inv = rinv * rinv, r6 = r2inv * r2inv * r2inv;
  double qq = lds_f64(C.pd, 3 * kMaxAtoms + i) * lds_f64(C.pd, 3 * kMaxAtoms + j);
  double lj = (LJ ? fma(ab.x, r6, -ab.y) : fma(C.ta * ab.x, r6, C.tb * ab.y)) * r6;  // lj: A r^-12 - B r^-6; repulsion: A r^-12; repulsioncg: B r^-6
  qq = live ? qq : 0.0;  // a padding slot re-reads a valid pair with weight 0
  lj = live ? lj : 0.0;
  if (MODE == 1) {
    const double r = r2 * rinv;
    const double t = fmax((r - C.rs) * C.inv_sw, 0.0);  // forces.py:401-411; rs = +inf (t = NaN -> 0) without a switch
    lj *= fma(t * t * t, fma(t, fma(t, -6.0, 15.0), -10.0), 1.0);
    const bool in = r2 <= C.rc2;  // _filter_by_cutoff, forces.py:75-80
    eel += in ? qq * (rinv + fma(C.krf, r2, C.ncrf)) : 0.0;  // forces.py:476 (krf = crf = 0 without rfa)
    elj += in ? lj : 0.0;
  } else {
    eel = fma(qq, rinv, eel);
    elj += lj;
  }
}

// Forces.ava_idx (forces.py:258-316); the row-major list keeps i warp-uniform and j consecutive.  Four pairs per thread
// and iteration with their own accumulators: the FP64 chain of a pair is serial (~25 dependent instructions), and four
// warps per scheduler do not hide that on their own.
template <int MODE, bool PBC, bool LJ>
__device__ __forceinline__ void epair_loop(const EPairConst& C, const DevTopo& T, int tid, double& eel, double& elj) {
  double el[4] = {0.0, 0.0, 0.0, 0.0}, lj[4] = {0.0, 0.0, 0.0, 0.0};
  const uint32_t* __restrict__ list = T.e_pair;
  const double2* __restrict__ tab = T.e_ab;
  const int n = T.ne_pair;
  for (int k0 = tid; k0 < n; k0 += 4 * kThreads) {
    uint32_t w[4];
    bool live[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int k = k0 + u * kThreads;
      live[u] = k < n;
      w[u] = __ldg(list + (live[u] ? k : k0));
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) epair<MODE, PBC, LJ>(C, tab, w[u], live[u], el[u], lj[u]);
  }
  eel = (el[0] + el[1]) + (el[2] + el[3]);
  elj = (lj[0] + lj[1]) + (lj[2] + lj[3]);
}

template <int MODE>
__device__ __forceinline__ void energy_pass_impl(const Smem& S, const DevTopo& T, const ForceParams& P, bool pbc, int tid, Energies& E) {
  double eel = 0.0, elj = 0.0, ebond = 0.0, eang = 0.0, edih = 0.0, eimp = 0.0;
  EPairConst C;
  C.pd = (uint32_t)__cvta_generic_to_shared(S.pd);
  C.ta = (double)P.tab_a; C.tb = (double)P.tab_b; C.rs = P.rs_d; C.inv_sw = P.inv_sw_d; C.rc2 = P.rc_d * P.rc_d;
  C.krf = P.krf_d; C.ncrf = -P.crf_d;
  C.bx = -(double)S.box[0]; C.by = -(double)S.box[1]; C.bz = -(double)S.box[2];
  if (P.terms & (B200MD_T_ELECTROSTATICS | B200MD_T_LJ | B200MD_T_REPULSION | B200MD_T_REPULSIONCG)) {
    const bool plain_lj = P.tab_a == 1.0f && P.tab_b == -1.0f;
    if (pbc) epair_loop<MODE, true, false>(C, T, tid, eel, elj);  // periodic replicas: one general instantiation
    else if (plain_lj) epair_loop<MODE, false, true>(C, T, tid, eel, elj);
    else epair_loop<MODE, false, false>(C, T, tid, eel, elj);
  }
  // ---- 1-4 pairs (forces.py:176-226): never cut, switched or reaction-field corrected
  if (T.p14_on) {
    const double ljon = (double)P.lj_on, elon = (double)P.el_on;
    for (int k = tid; k < T.ne_14; k += kThreads) {
      const uint32_t w = __ldg(T.e14_idx + k);
      const dv3 d = ediff(C, pbc, w & 0xff, (w >> 8) & 0xff);
      const double r2 = edot(d, d), rinv = drsqrt(r2), r2inv = rinv * rinv, r6 = r2inv * r2inv * r2inv;
      elj += ljon * fma(__ldg(T.e14_prm + 3 * k), r6, -__ldg(T.e14_prm + 3 * k + 1)) * r6;
      eel += elon * __ldg(T.e14_prm + 3 * k + 2) * rinv;
    }
  }
  // ---- bonds (forces.py:111-132, 492-501)
  if (T.nbond2 > 0) {
    for (int k = tid; k < T.ne_bond; k += kThreads) {
      const uint32_t w = __ldg(T.eb_idx + k);
      const float2 kp = __ldg(T.eb_prm + k);
      const dv3 d = ediff(C, pbc, w & 0xff, (w >> 8) & 0xff);
      const double r2 = edot(d, d), dist = r2 * drsqrt(r2);
      const double x = dist - (double)kp.y;
      const bool keep = !P.has_cutoff || dist <= P.rc_d;
      ebond += keep ? (double)kp.x * x * x : 0.0;
    }
  }
  // ---- angles (forces.py:134-145, 504-518)
  if (T.nangle2 > 0) {
    for (int k = tid; k < T.ne_angle; k += kThreads) {
      const uint32_t w = __ldg(T.ea_idx + k);
      const float2 kp = __ldg(T.ea_prm + k);
      const int a1 = (w >> 8) & 0xff;
      const dv3 r21 = ediff(C, pbc, w & 0xff, a1), r23 = ediff(C, pbc, (w >> 16) & 0xff, a1);
      double ct = edot(r23, r21) * drsqrt(edot(r21, r21)) * drsqrt(edot(r23, r23));
      ct = ct < -1.0 ? -1.0 : (ct > 1.0 ? 1.0 : ct);
      const double dth = acos(ct) - (double)kp.y;
      eang += (double)kp.x * dth * dth;
    }
  }
  // ---- proper torsions and impropers (forces.py:147-174, 228-255)
  {
    const int k0 = T.dih_on ? 0 : T.ne_dih, k1 = T.gen_end > T.nspec ? T.ne_dih + T.ne_imp : T.ne_dih;  // enabled kinds
    for (int k = k0 + tid; k < k1; k += kThreads) {
      const double e = etorsion(C, T, pbc, k);
      if (k < T.ne_dih) edih += e; else eimp += e;
    }
  }
  E.e[B200MD_E_ELECTROSTATICS] += eel; E.e[B200MD_E_LJ] += elj; E.e[B200MD_E_BONDS] += ebond;
  E.e[B200MD_E_ANGLES] += eang; E.e[B200MD_E_DIHEDRALS] += edih; E.e[B200MD_E_IMPROPERS] += eimp;
}

// Inside the fused integrator the pass runs on report steps only: a call keeps its code (FP64 transcendentals) out of
// the hot instruction stream.  energy_kernel (b200md.cu) inlines energy_pass_impl instead.
template <int MODE>
__device__ __noinline__ void energy_pass(const Smem& S, const DevTopo& T, const ForceParams& P, bool pbc, int tid, Energies& E) {
  energy_pass_impl<MODE>(S, T, P, pbc, tid, E);
}

// ---- all tasks of this warp for one force evaluation; positions must be in S.pos ------------------
// The warp's static task list is sorted by kind on the host (T.kend[warp][k] = end of the k-th run), so every kind
// is a tight loop around its own body: no switch, and the next task word is fetched while the current task runs
// (the fetch -> decode -> branch chain of a switch-per-task loop cost 33 issue slots and ~150 cycles of exposed
// latency per task, profiles/r01_v11).
template <int MODE, bool FIXED, bool PBC, bool SKIP = true>
__device__ __forceinline__ void run_tasks(const Smem& S, const DevTopo& T, const ForceParams& P, int warp, int lane) {
  const uint32_t* tk = S.tasks + warp * kMaxTasks;
  const uint2 kends = T.kend8[warp];  // the 7 run ends as bytes: one constant load instead of one per kind
  int t = 0;
  uint32_t next = tk[0];
#define B200MD_TASK_LOOP(K, BODY)                                                     \
  {                                                                                   \
    const int end = (int)(((K) < 4 ? kends.x >> (8 * ((K) & 3)) : kends.y >> (8 * ((K) & 3))) & 0xffu);  \
    _Pragma("unroll 1") for (; t < end;) {                                            \
      const uint32_t task = next;                                                     \
      next = tk[++t];                                                                 \
      const uint32_t chunk = (task >> 8) & 0xffff, passes = task >> 24;               \
      const int item = chunk * 32 + lane;                                             \
      (void)passes; (void)item;                                                       \
      BODY;                                                                           \
    }                                                                                 \
  }
  B200MD_TASK_LOOP(0, (tile_task<MODE, PBC, SKIP>(S, P, chunk, passes, warp, lane)))
  B200MD_TASK_LOOP(1, (tile_diag_task<MODE, PBC>(S, P, chunk & 0xff, (chunk >> 8) & 0xff, passes, warp, lane)))
  B200MD_TASK_LOOP(2, (dihedral_task<FIXED, PBC>(S, T, P, item, (int)passes)))
  B200MD_TASK_LOOP(3, (torsion_task<FIXED, PBC>(S, T, item)))
  B200MD_TASK_LOOP(4, (angle_task<FIXED, PBC>(S, T, item)))
  B200MD_TASK_LOOP(5, (bond_task<FIXED, PBC>(S, T, P, item)))
  B200MD_TASK_LOOP(6, (p14_task<FIXED, PBC>(S, T, P, item)))
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
  S.smd_energy[1] += (double)P.smd_fk * dev * (double)P.smd_rate;  // work of the moving restraint, dW = dE/dc dc
}

// ---- per-atom gather in a fixed order: j-side slot, 4 i-side partials, bonded slots, SMD ------------
template <bool FIXED>
__device__ __forceinline__ void gather_forces(const Smem& S, const ForceParams& P, int a, float& fx, float& fy,
                                              float& fz, uint32_t& flag) {
  fx = S.fj[a]; fy = S.fj[kMaxAtoms + a]; fz = S.fj[2 * kMaxAtoms + a];
#pragma unroll
  for (int w = 0; w < kWarps; ++w) {
    fx += S.fi[(w * 3 + 0) * kMaxAtoms + a];
    fy += S.fi[(w * 3 + 1) * kMaxAtoms + a];
    fz += S.fi[(w * 3 + 2) * kMaxAtoms + a];
    S.fi[(w * 3 + 0) * kMaxAtoms + a] = 0.0f; S.fi[(w * 3 + 1) * kMaxAtoms + a] = 0.0f; S.fi[(w * 3 + 2) * kMaxAtoms + a] = 0.0f;  // re-arm
  }
  if (FIXED) {  // bonded forces arrived as fixed-point atomics; read and re-arm the accumulators
    const int ix = S.fbi[a], iy = S.fbi[kMaxAtoms + a], iz = S.fbi[2 * kMaxAtoms + a];
    if (max(max(abs(ix), abs(iy)), abs(iz)) > kFixedLimit) flag |= B200MD_FLAG_FORCE_RANGE;
    fx += (float)ix * kFixedInv; fy += (float)iy * kFixedInv; fz += (float)iz * kFixedInv;
    S.fbi[a] = 0; S.fbi[kMaxAtoms + a] = 0; S.fbi[2 * kMaxAtoms + a] = 0;
  } else {
    const int g1 = S.gstart[a + 1];  // ranges are multiples of 4 slots, planes 128-byte aligned: one LDS.128 per plane and quad
#pragma unroll 2
    for (int g = S.gstart[a]; g < g1; g += 4) {
      const float4 vx = *(const float4*)(S.fb + g), vy = *(const float4*)(S.fb + S.fbs + g), vz = *(const float4*)(S.fb + 2 * S.fbs + g);
      fx += (vx.x + vx.y) + (vx.z + vx.w); fy += (vy.x + vy.y) + (vy.z + vy.w); fz += (vz.x + vz.y) + (vz.z + vz.w);
    }
  }
  if (P.smd_on) {
    if (a == P.smd_a2) { fx += S.smd[0]; fy += S.smd[1]; fz += S.smd[2]; }
    if (a == P.smd_a1) { fx -= S.smd[0]; fy -= S.smd[1]; fz -= S.smd[2]; }
  }
}

// ---- rigid X-H bonds: SHAKE on the positions, RATTLE on the velocities ------------------------------
// One lane per hydrogen group (heavy atom p with its n <= 3 hydrogens): the groups do not share atoms, so the lanes
// are independent.  The group's n constraints are solved TOGETHER: Newton iterations on the n multipliers with the
// exact n x n Jacobian for the positions (the matrix form of SHAKE; quadratic convergence, so a fixed 3 iterations
// take a 2 fs step's ~1e-2 relative violation below fp32 resolution), one linear solve for the velocities (RATTLE is
// linear in its multipliers).  No data-dependent trip counts.  The step is issue-bound, so what counts is warp
// instructions, not lanes: the host sorts the groups by size and packs them into the lanes of the first warps
// (ALA10: 51 groups -> 2 warps instead of the 4 a lane-per-parent-atom mapping would run), and a warp whose groups
// all have one hydrogen runs the K = 1 instantiation (a scalar Newton iteration) instead of the 3 x 3 one.
constexpr int kShakeNewtonIters = 3;
constexpr float kShakeTol = 4.0e-6f;  // accepted |d^2 - r^2| / d^2 after the last iteration (flag beyond it)

// x = A^-1 b by cofactors (A is strongly diagonally dominant here: m_H << m_parent); K = 1: a division
template <int K>
__device__ __forceinline__ void solve_k(const float (&A)[K][K], const float (&b)[K], float (&x)[K]) {
  if constexpr (K == 1) {
    x[0] = b[0] * fast_rcp(A[0][0]);
  } else {
    const float c00 = A[1][1] * A[2][2] - A[1][2] * A[2][1], c01 = A[1][2] * A[2][0] - A[1][0] * A[2][2],
                c02 = A[1][0] * A[2][1] - A[1][1] * A[2][0];
    const float inv = fast_rcp(A[0][0] * c00 + A[0][1] * c01 + A[0][2] * c02);
    const float c10 = A[0][2] * A[2][1] - A[0][1] * A[2][2], c11 = A[0][0] * A[2][2] - A[0][2] * A[2][0],
                c12 = A[0][1] * A[2][0] - A[0][0] * A[2][1];
    const float c20 = A[0][1] * A[1][2] - A[0][2] * A[1][1], c21 = A[0][2] * A[1][0] - A[0][0] * A[1][2],
                c22 = A[0][0] * A[1][1] - A[0][1] * A[1][0];
    x[0] = (c00 * b[0] + c10 * b[1] + c20 * b[2]) * inv;
    x[1] = (c01 * b[0] + c11 * b[1] + c21 * b[2]) * inv;
    x[2] = (c02 * b[0] + c12 * b[1] + c22 * b[2]) * inv;
  }
}

// In:  S.pos = x(t) (constraints hold), S.sx = unconstrained x(t + dt).  Out: S.vel[a].xyz = displacement that
// restores the constraints of atom a's group (the slots were zeroed by their owners).  Displacements, not
// positions, travel through shared memory: the velocity correction d / dt keeps its relative precision.
// With multipliers l_k along the old bond vectors s_k = x_p(t) - x_hk(t):
//   d_p = w_p sum_k l_k s_k,  d_hk = -w_hk l_k s_k,  r_k(l) = u_k + d_p - d_hk,  sigma_k = |r_k|^2 - d_k^2 = 0.
template <int K>
__device__ __forceinline__ bool shake_group(const Smem& S, const DevTopo& T, int gi) {
  const int4 g = __ldg(T.cgroup + gi);  // (parent | n << 8, h0, h1, h2)
  const float4 d2v = __ldg(T.cdist2 + gi), wv = __ldg(T.cweight + gi);
  const int p = g.x & 0xff, n = g.x >> 8;
  const int h[3] = {g.y, g.z, g.w};
  const float d2[3] = {d2v.x, d2v.y, d2v.z};
  const float wp = wv.x, whv[3] = {wv.y, wv.z, wv.w};
  const float4 po = S.pos[p];
  const float pnx = S.sx[p], pny = S.sx[kMaxAtoms + p], pnz = S.sx[2 * kMaxAtoms + p];
  float wh[K], sx[K], sy[K], sz[K], ux[K], uy[K], uz[K], lam[K];
#pragma unroll
  for (int k = 0; k < K; ++k) {
    const int a = k < n ? h[k] : p;  // an unused row degenerates to s = u = 0 and is given the identity below
    const float4 ho = S.pos[a];
    wh[k] = whv[k];
    sx[k] = po.x - ho.x; sy[k] = po.y - ho.y; sz[k] = po.z - ho.z;
    ux[k] = pnx - S.sx[a]; uy[k] = pny - S.sx[kMaxAtoms + a]; uz[k] = pnz - S.sx[2 * kMaxAtoms + a];
    lam[k] = 0.0f;
  }
  float dpx = 0.0f, dpy = 0.0f, dpz = 0.0f;
#pragma unroll
  for (int it = 0; it < kShakeNewtonIters; ++it) {
    float A[K][K], b[K], rx[K], ry[K], rz[K], dl[K];
#pragma unroll
    for (int k = 0; k < K; ++k) {
      rx[k] = fmaf(wh[k] * lam[k], sx[k], ux[k] + dpx);
      ry[k] = fmaf(wh[k] * lam[k], sy[k], uy[k] + dpy);
      rz[k] = fmaf(wh[k] * lam[k], sz[k], uz[k] + dpz);
      b[k] = k < n ? d2[k] - (rx[k] * rx[k] + ry[k] * ry[k] + rz[k] * rz[k]) : 0.0f;
    }
#pragma unroll
    for (int k = 0; k < K; ++k)
#pragma unroll
      for (int l = 0; l < K; ++l) {
        const float sr = rx[k] * sx[l] + ry[k] * sy[l] + rz[k] * sz[l];
        A[k][l] = 2.0f * sr * (k == l ? wp + wh[k] : wp);
        if (k >= n || l >= n) A[k][l] = k == l ? 1.0f : 0.0f;
      }
    solve_k<K>(A, b, dl);
    dpx = 0.0f; dpy = 0.0f; dpz = 0.0f;
#pragma unroll
    for (int k = 0; k < K; ++k) {
      lam[k] += dl[k];
      dpx = fmaf(wp * lam[k], sx[k], dpx); dpy = fmaf(wp * lam[k], sy[k], dpy); dpz = fmaf(wp * lam[k], sz[k], dpz);
    }
  }
  bool ok = true;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    if (k < n) {
      const float hx = -wh[k] * lam[k] * sx[k], hy = -wh[k] * lam[k] * sy[k], hz = -wh[k] * lam[k] * sz[k];
      const float rx = ux[k] + (dpx - hx), ry = uy[k] + (dpy - hy), rz = uz[k] + (dpz - hz);
      ok = ok && fabsf(d2[k] - (rx * rx + ry * ry + rz * rz)) <= kShakeTol * d2[k];
      S.vel[h[k]] = make_float4(hx, hy, hz, 0.0f);
    }
  }
  S.vel[p] = make_float4(dpx, dpy, dpz, 0.0f);
  return ok;
}

// In: S.pos = x(t + dt) (constraints hold), S.vel = v after the second half kick.  Out: S.vel without the
// components along the rigid bonds:  v_p -= w_p sum_l m_l r_l,  v_hk += w_hk m_k r_k  with
//   sum_l m_l (w_p r_k.r_l + delta_kl w_hk r_k.r_k) = r_k.(v_p - v_hk).
template <int K>
__device__ __forceinline__ void rattle_group(const Smem& S, const DevTopo& T, int gi) {
  const int4 g = __ldg(T.cgroup + gi);
  const float4 wv = __ldg(T.cweight + gi);
  const int p = g.x & 0xff, n = g.x >> 8;
  const int h[3] = {g.y, g.z, g.w};
  const float wp = wv.x, whv[3] = {wv.y, wv.z, wv.w};
  const float4 pp = S.pos[p];
  float4 vp = S.vel[p];
  float wh[K], rx[K], ry[K], rz[K], vx[K], vy[K], vz[K], A[K][K], b[K], mu[K];
#pragma unroll
  for (int k = 0; k < K; ++k) {
    const int a = k < n ? h[k] : p;
    const float4 hp = S.pos[a], hv = S.vel[a];
    wh[k] = whv[k];
    rx[k] = pp.x - hp.x; ry[k] = pp.y - hp.y; rz[k] = pp.z - hp.z;
    vx[k] = hv.x; vy[k] = hv.y; vz[k] = hv.z;
    b[k] = rx[k] * (vp.x - vx[k]) + ry[k] * (vp.y - vy[k]) + rz[k] * (vp.z - vz[k]);
  }
#pragma unroll
  for (int k = 0; k < K; ++k)
#pragma unroll
    for (int l = 0; l < K; ++l) {
      const float rr = rx[k] * rx[l] + ry[k] * ry[l] + rz[k] * rz[l];
      A[k][l] = rr * (k == l ? wp + wh[k] : wp);
      if (k >= n || l >= n) A[k][l] = k == l ? 1.0f : 0.0f;
    }
  solve_k<K>(A, b, mu);
#pragma unroll
  for (int k = 0; k < K; ++k) {
    if (k < n) {
      vp.x = fmaf(-wp * mu[k], rx[k], vp.x); vp.y = fmaf(-wp * mu[k], ry[k], vp.y); vp.z = fmaf(-wp * mu[k], rz[k], vp.z);
      S.vel[h[k]] = make_float4(fmaf(wh[k] * mu[k], rx[k], vx[k]), fmaf(wh[k] * mu[k], ry[k], vy[k]),
                                fmaf(wh[k] * mu[k], rz[k], vz[k]), 0.0f);
    }
  }
  S.vel[p] = vp;
}

// ---- block reduction of the first NOBS FP64 accumulators (fixed order); thread 0 ends with the totals -------
template <int NOBS>
__device__ __forceinline__ void reduce_energies(const Smem& S, Energies& E, int warp, int lane) {
#pragma unroll
  for (int k = 0; k < NOBS; ++k) {
    double v = E.e[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) S.ered[warp * kNumReduced + k] = v;
  }
  __syncthreads();
  if (warp == 0 && lane == 0) {
#pragma unroll
    for (int k = 0; k < NOBS; ++k)
      E.e[k] = (S.ered[k] + S.ered[kNumReduced + k]) +
               (S.ered[2 * kNumReduced + k] + S.ered[3 * kNumReduced + k]);
  }
}

}  // namespace b200md
