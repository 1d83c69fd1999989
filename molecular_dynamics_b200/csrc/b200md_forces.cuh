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
// trajectory does not depend on how often energies are reported.  `ab` holds (12 A, 6 B).
// Energies (report steps only) use one Newton step on rsqrt: the MUFU error (2^-22) would otherwise be
// amplified 12x in r^-12 and miss the 1e-6 energy tolerance.
template <int MODE, bool ENERGY>
__device__ __forceinline__ float pair_eval(float dx, float dy, float dz, float kqi, float qj, float2 ab, bool excl,
                                           const ForceParams& P, float& pel, float& plj) {
  const float r2 = __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmul_rn(dx, dx)));
  float rinv = fast_rsqrt(r2);
  if (MODE == 1) excl = excl || !(r2 <= P.rc2);
  rinv = excl ? 0.0f : rinv;
  const float r2inv = __fmul_rn(rinv, rinv);
  const float r6 = __fmul_rn(__fmul_rn(r2inv, r2inv), r2inv);
  const float qq = (MODE == 1 && excl) ? 0.0f : __fmul_rn(kqi, qj);
  float glj = __fmul_rn(__fmaf_rn(ab.x, r6, -ab.y), r6);  // 12 A r^-12 - 6 B r^-6 = -r dE/dr
  float gc;
  float corr = 1.0f, elj = 0.0f;
  if (ENERGY || MODE == 1) {
    // Newton correction factor for rinv: rinv_exact ~= rinv * corr
    corr = __fmaf_rn(__fmul_rn(r2, r2inv), -0.5f, 1.5f);
    const float c2 = corr * corr;
    const float r6e = r6 * (c2 * c2 * c2);
    elj = (ab.x * r6e * (1.0f / 12.0f) - ab.y * (1.0f / 6.0f)) * r6e;  // A r^-12 - B r^-6
  }
  if (MODE == 0) {
    gc = __fmul_rn(qq, rinv);  // K qi qj / r = -r dE/dr
    if (ENERGY) pel += gc * corr;
  } else {
    gc = __fmul_rn(qq, __fmaf_rn(-2.0f * P.krf, r2, rinv));  // -r * forces.py:478
    const float rinv_e = rinv * corr;
    const float r = __fmul_rn(r2, rinv_e);
    if (r > P.rs) {  // forces.py:401-411 (note pot*deriv/dist as written)
      const float t = __fmul_rn(r - P.rs, P.inv_sw);
      const float t2 = __fmul_rn(t, t);
      const float sv = __fmaf_rn(__fmul_rn(t2, t), __fmaf_rn(t, __fmaf_rn(t, -6.0f, 15.0f), -10.0f), 1.0f);
      const float sd = __fmul_rn(__fmul_rn(t2, __fmaf_rn(t, __fmaf_rn(t, -30.0f, 60.0f), -30.0f)), P.inv_sw);
      glj = __fmaf_rn(sv, glj, -__fmul_rn(elj, sd));
      elj *= sv;
    }
    if (ENERGY) pel += qq * (rinv_e + fmaf(P.krf, r2, -P.crf));  // forces.py:476
  }
  if (ENERGY) plj += elj;
  return __fmul_rn(__fadd_rn(glj, gc), r2inv);
}

// ---- TASK_TILE: j-chunk `chunk` (8 atoms) against i-blocks 0..passes-1 (32 lanes each) ----------
template <int MODE, bool ENERGY>
__device__ __forceinline__ void tile_task(const Smem& S, const ForceParams& P, int chunk, int passes, int lane,
                                          float (&fi)[4][3], Energies& E) {
  float fj[kChunk][3];
#pragma unroll
  for (int s = 0; s < kChunk; ++s) fj[s][0] = fj[s][1] = fj[s][2] = 0.0f;
  const float4* pj = S.pos + chunk * kChunk;
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    if (p < passes) {
      const int i = p * 32 + lane;
      const float4 pi = S.pos[i];
      const float kqi = __fmul_rn(pi.w, kElecFactor);
      const float2* row = (const float2*)((const char*)S.ab + S.ab_row_off[i]) + chunk * kChunk;
      const uint32_t m8 = S.mask8[chunk * kMaxAtoms + i];
      float pel = 0.0f, plj = 0.0f, ax = 0.0f, ay = 0.0f, az = 0.0f;
#pragma unroll
      for (int s = 0; s < kChunk; ++s) {
        const float4 q = pj[s];
        const float dx = __fsub_rn(pi.x, q.x), dy = __fsub_rn(pi.y, q.y), dz = __fsub_rn(pi.z, q.z);
        const float sc = pair_eval<MODE, ENERGY>(dx, dy, dz, kqi, q.w, row[s], (m8 >> s) & 1u, P, pel, plj);
        ax = __fmaf_rn(sc, dx, ax); ay = __fmaf_rn(sc, dy, ay); az = __fmaf_rn(sc, dz, az);
        fj[s][0] = __fmaf_rn(sc, dx, fj[s][0]); fj[s][1] = __fmaf_rn(sc, dy, fj[s][1]); fj[s][2] = __fmaf_rn(sc, dz, fj[s][2]);
      }
      fi[p][0] = __fadd_rn(fi[p][0], ax); fi[p][1] = __fadd_rn(fi[p][1], ay); fi[p][2] = __fadd_rn(fi[p][2], az);
      if (ENERGY) { E.e[B200MD_E_ELECTROSTATICS] += (double)pel; E.e[B200MD_E_LJ] += (double)plj; }
    }
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
      v4[k][c] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
#pragma unroll
  for (int k = 0; k < 2; ++k)
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float send = b3 ? v4[k][c] : v4[k + 2][c];
      const float keep = b3 ? v4[k + 2][c] : v4[k][c];
      v2[k][c] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float send = b2 ? v2[0][c] : v2[1][c];
    const float keep = b2 ? v2[1][c] : v2[0][c];
    float v = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    v1[c] = v;
  }
  if ((lane & 3) == 0) {
    const int j = chunk * kChunk + (lane >> 2);
    S.fj[0 * kMaxAtoms + j] = -v1[0];
    S.fj[1 * kMaxAtoms + j] = -v1[1];
    S.fj[2 * kMaxAtoms + j] = -v1[2];
  }
}

// ---- bonded slots -----------------------------------------------------------------------------
__device__ __forceinline__ void put_slot(const Smem& S, const DevTopo& T, int type, int pos, int item, float x,
                                         float y, float z) {
  const int k = T.slot_base[type] + pos * T.slot_pad[type] + item;
  S.fb[k] = x;
  S.fb[T.nslots + k] = y;
  S.fb[2 * T.nslots + k] = z;
}

// evaluate_bonds, forces.py:111-132, 492-501
template <bool ENERGY>
__device__ __forceinline__ void bond_task(const Smem& S, const DevTopo& T, const ForceParams& P, int item, Energies& E) {
  if (item >= T.nbonds) return;
  const uint32_t q = __ldg(T.bond_idx + item);
  const float2 prm = __ldg(T.bond_prm + item);
  const float4 pa = S.pos[q & 0xff], pb = S.pos[(q >> 8) & 0xff];
  const f3 d = sub3(pa, pb);
  const float r2 = dot3(d, d);
  const float rinv = rsqrtf(r2), dist = r2 * rinv;
  const bool keep = !P.has_cutoff || dist <= P.rc;
  const float x = dist - prm.y;
  const float c = keep ? 2.0f * prm.x * x * rinv : 0.0f;  // dE/dr / r
  put_slot(S, T, TASK_BOND, 0, item, -c * d.x, -c * d.y, -c * d.z);
  put_slot(S, T, TASK_BOND, 1, item, c * d.x, c * d.y, c * d.z);
  if (ENERGY && keep) {
    // k (r - r0)^2 cancels ~40:1 in (r - r0): the energy (report steps only) takes r in FP64
    const double ddx = (double)pa.x - (double)pb.x, ddy = (double)pa.y - (double)pb.y, ddz = (double)pa.z - (double)pb.z;
    const double xd = sqrt(ddx * ddx + ddy * ddy + ddz * ddz) - (double)prm.y;
    E.e[B200MD_E_BONDS] += (double)prm.x * xd * xd;
  }
}

// evaluate_angles, forces.py:134-145, 504-537
template <bool ENERGY>
__device__ __forceinline__ void angle_task(const Smem& S, const DevTopo& T, int item, Energies& E) {
  if (item >= T.nangles) return;
  const uint32_t q = __ldg(T.angle_idx + item);
  const float2 prm = __ldg(T.angle_prm + item);
  const float4 p0 = S.pos[q & 0xff], p1 = S.pos[(q >> 8) & 0xff], p2 = S.pos[(q >> 16) & 0xff];
  const f3 r21 = sub3(p0, p1), r23 = sub3(p2, p1);
  const float n21i = rsqrtf(dot3(r21, r21)), n23i = rsqrtf(dot3(r23, r23));
  float ct = dot3(r23, r21) * n21i * n23i;
  ct = fminf(fmaxf(ct, -1.0f), 1.0f);
  const float th = acosf(ct), dth = th - prm.y;
  if (ENERGY) E.e[B200MD_E_ANGLES] += (double)(prm.x * dth * dth);
  const float st = sqrtf(fmaf(-ct, ct, 1.0f));
  const float coef = (st != 0.0f) ? -2.0f * prm.x * dth / st : 0.0f;
  const float a0 = coef * n21i, a2 = coef * n23i;
  const float c21 = ct * n21i, c23 = ct * n23i;
  const float f0x = a0 * (c21 * r21.x - r23.x * n23i), f0y = a0 * (c21 * r21.y - r23.y * n23i),
              f0z = a0 * (c21 * r21.z - r23.z * n23i);
  const float f2x = a2 * (c23 * r23.x - r21.x * n21i), f2y = a2 * (c23 * r23.y - r21.y * n21i),
              f2z = a2 * (c23 * r23.z - r21.z * n21i);
  put_slot(S, T, TASK_ANGLE, 0, item, f0x, f0y, f0z);
  put_slot(S, T, TASK_ANGLE, 1, item, -(f0x + f2x), -(f0y + f2y), -(f0z + f2z));
  put_slot(S, T, TASK_ANGLE, 2, item, f2x, f2y, f2z);
}

// evaluate_torsion, forces.py:147-174 / 228-255, 540-602 (dihedrals and impropers share it)
template <bool ENERGY>
__device__ __forceinline__ void torsion_task(const Smem& S, const DevTopo& T, int type, int item, int count,
                                             const uint32_t* idx, const uint2* terms, const float4* term_prm,
                                             double& Eacc) {
  if (item >= count) return;
  const uint32_t q = __ldg(idx + item);
  const float4 p0 = S.pos[q & 0xff], p1 = S.pos[(q >> 8) & 0xff], p2 = S.pos[(q >> 16) & 0xff], p3 = S.pos[q >> 24];
  const f3 r12 = sub3(p0, p1), r23 = sub3(p1, p2), r34 = sub3(p2, p3);
  const f3 cA = cross3(r12, r23), cB = cross3(r23, r34), cC = cross3(r23, cA);
  const float nA2 = dot3(cA, cA), nB2 = dot3(cB, cB), nC2 = dot3(cC, cC);
  const float rnA = rsqrtf(nA2), rnB = rsqrtf(nB2), rnC = rsqrtf(nC2);
  const float cosPhi = dot3(cA, cB) * rnB * rnA, sinPhi = dot3(cC, cB) * rnB * rnC;
  const float phi = -atan2f(sinPhi, cosPhi);
  const uint2 tr = __ldg(terms + item);
  float pot = 0.0f, coeff = 0.0f;
  for (uint32_t t = tr.x; t < tr.x + tr.y; ++t) {
    const float4 tp = __ldg(term_prm + t);  // k0, phi0, per, harmonic
    if (tp.w == 0.0f) {                     // AMBER, forces.py:565-569
      float sn, cs;
      sincosf(fmaf(tp.z, phi, -tp.y), &sn, &cs);
      pot += tp.x * (1.0f + cs);
      coeff -= tp.z * tp.x * sn;
    } else {                                // CHARMM harmonic, forces.py:570-576
      float ad = phi - tp.y;
      const float pi_f = 3.14159265358979323846f;
      if (ad < -pi_f) ad += 2.0f * pi_f; else if (ad > pi_f) ad -= 2.0f * pi_f;
      pot += tp.x * ad * ad;
      coeff += 2.0f * tp.x * ad;
    }
  }
  if (ENERGY) Eacc += (double)pot;
  const float nd2sq = dot3(r23, r23), nd2 = sqrtf(nd2sq), inv_nd2sq = 1.0f / nd2sq;
  const float ff0 = -coeff * nd2 / nA2, ff3 = coeff * nd2 / nB2;
  const float ff1 = dot3(r12, r23) * inv_nd2sq, ff2 = dot3(r34, r23) * inv_nd2sq;
  const f3 f0 = {ff0 * cA.x, ff0 * cA.y, ff0 * cA.z}, f3v = {ff3 * cB.x, ff3 * cB.y, ff3 * cB.z};
  const f3 s = {ff1 * f0.x - ff2 * f3v.x, ff1 * f0.y - ff2 * f3v.y, ff1 * f0.z - ff2 * f3v.z};
  put_slot(S, T, type, 0, item, -f0.x, -f0.y, -f0.z);
  put_slot(S, T, type, 1, item, f0.x + s.x, f0.y + s.y, f0.z + s.z);
  put_slot(S, T, type, 2, item, f3v.x - s.x, f3v.y - s.y, f3v.z - s.z);
  put_slot(S, T, type, 3, item, -f3v.x, -f3v.y, -f3v.z);
}

// 1-4 pairs, forces.py:176-226: own A/B scaled by scnb, Coulomb scaled by scee; never cut/switched.
template <bool ENERGY>
__device__ __forceinline__ void p14_task(const Smem& S, const DevTopo& T, const ForceParams& P, int item, Energies& E) {
  if (item >= T.n14) return;
  const uint32_t q = __ldg(T.p14_idx + item);
  const float4 prm = __ldg(T.p14_prm + item);
  const float4 pa = S.pos[q & 0xff], pb = S.pos[(q >> 8) & 0xff];
  const f3 d = sub3(pa, pb);
  const float rinv = rsqrt_nr(dot3(d, d));
  const float r2inv = rinv * rinv, r6 = r2inv * r2inv * r2inv;
  const float a = prm.x * P.lj_on, b = prm.y * P.lj_on, qq = prm.z * P.el_on;
  const float u = a * r6, t1 = u - b;
  const float elj = t1 * r6, ec = qq * rinv;
  const float sc = (6.0f * ((u + t1) * r6) + ec) * r2inv;
  put_slot(S, T, TASK_P14, 0, item, sc * d.x, sc * d.y, sc * d.z);
  put_slot(S, T, TASK_P14, 1, item, -sc * d.x, -sc * d.y, -sc * d.z);
  if (ENERGY) { E.e[B200MD_E_LJ] += (double)elj; E.e[B200MD_E_ELECTROSTATICS] += (double)ec; }
}

// ---- all tasks of this warp for one force evaluation; positions must be in S.pos ------------------
template <int MODE, bool ENERGY>
__device__ __forceinline__ void run_tasks(const Smem& S, const DevTopo& T, const ForceParams& P, int warp, int lane,
                                          Energies& E) {
  float fi[4][3];
#pragma unroll
  for (int p = 0; p < 4; ++p) fi[p][0] = fi[p][1] = fi[p][2] = 0.0f;
  const int nt = T.ntasks[warp];
  for (int t = 0; t < nt; ++t) {
    const uint32_t task = S.tasks[warp * kMaxTasks + t];
    const uint32_t kind = task & 0xff, chunk = (task >> 8) & 0xffff, passes = task >> 24;
    const int item = chunk * 32 + lane;
    switch (kind) {
      case TASK_TILE: tile_task<MODE, ENERGY>(S, P, chunk, passes, lane, fi, E); break;
      case TASK_BOND: bond_task<ENERGY>(S, T, P, item, E); break;
      case TASK_ANGLE: angle_task<ENERGY>(S, T, item, E); break;
      case TASK_DIHEDRAL:
        torsion_task<ENERGY>(S, T, TASK_DIHEDRAL, item, T.ndih, T.dih_idx, T.dih_terms, T.dih_term_prm, E.e[B200MD_E_DIHEDRALS]);
        break;
      case TASK_IMPROPER:
        torsion_task<ENERGY>(S, T, TASK_IMPROPER, item, T.nimp, T.imp_idx, T.imp_terms, T.imp_term_prm, E.e[B200MD_E_IMPROPERS]);
        break;
      case TASK_P14: p14_task<ENERGY>(S, T, P, item, E); break;
      default: break;
    }
  }
#pragma unroll
  for (int p = 0; p < 4; ++p)
#pragma unroll
    for (int c = 0; c < 3; ++c) S.fi[(warp * 3 + c) * kMaxAtoms + p * 32 + lane] = fi[p][c];
}

// ---- SMD end-to-end bias (potcalc.py:75-89; force = -dE/dr): one lane, during the task phase -------
__device__ __forceinline__ void smd_task(const Smem& S, const ForceParams& P, float center) {
  const float4 p1 = S.pos[P.smd_a1], p2 = S.pos[P.smd_a2];
  const f3 d = sub3(p2, p1);
  const float r2 = dot3(d, d), rinv = rsqrt_nr(r2), dist = r2 * rinv;
  const float dev = center - dist;
  const float c = P.smd_fk * dev * rinv;  // force on atom2 = +fk (c - d) unit(r2 - r1)
  S.smd[0] = c * d.x; S.smd[1] = c * d.y; S.smd[2] = c * d.z;
  S.smd[3] = 0.5f * P.smd_fk * dev * dev;
  S.smd[4] = dist;
}

// ---- per-atom gather in a fixed order: 4 i-side partials, the j-side slot, bonded slots, SMD --------
__device__ __forceinline__ void gather_forces(const Smem& S, const DevTopo& T, const ForceParams& P, int a,
                                              float& fx, float& fy, float& fz) {
  fx = S.fj[a]; fy = S.fj[kMaxAtoms + a]; fz = S.fj[2 * kMaxAtoms + a];
#pragma unroll
  for (int w = 0; w < kWarps; ++w) {
    fx += S.fi[(w * 3 + 0) * kMaxAtoms + a];
    fy += S.fi[(w * 3 + 1) * kMaxAtoms + a];
    fz += S.fi[(w * 3 + 2) * kMaxAtoms + a];
  }
  const int g0 = S.gstart[a], g1 = S.gstart[a + 1];
  for (int g = g0; g < g1; ++g) {
    const int k = S.glist[g];
    fx += S.fb[k]; fy += S.fb[T.nslots + k]; fz += S.fb[2 * T.nslots + k];
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
