// b200md_large.cuh -- solutes above 128 atoms (SURVEY 8f N4 "larger solutes": the adk-1.0.prmtop system named at
// V_Calculations/Test-10_diALA/potcalc.py:114).  Same reference semantics as the one-CTA-per-replica kernels
// (TorchMD_Fix/forces.py Forces.compute :82-344, upstream Integrator.step), but a replica no longer fits on chip:
//
//   large_bonded_kernel  one thread per bonded item (bonds, angles, proper torsions with all their terms, 1-4 pairs,
//                        impropers, the SMD restraint) in FP64; force vectors go to a per-atom 2^-30 fixed-point
//                        int64 accumulator in HBM by integer atomics (associative: the sum does not depend on the
//                        order the atomics land in, so results stay bit-reproducible);
//   large_pairs_kernel   all-pairs LJ + Coulomb, FULL shell: CTA = one 32-atom i-block of one replica, lane = i atom,
//                        the 4 warps split the j tiles of 32 atoms (staged per warp in shared memory, read as
//                        broadcasts), exclusions are one 32-bit mask word per (i atom, j tile); no j-side force, no
//                        atomics; the warps' partial sums meet in shared memory in a fixed order and the force row is
//                        written once, coalesced (with the bonded accumulator added when the bonded kernel ran before it
//                        on the same stream; by default that kernel runs beside this one on a second stream and
//                        large_combine_kernel adds the accumulator afterwards: same expression, same bits);
//   large_bbox_kernel    cutoff mode: bounding box of every 32-atom tile, so a warp can drop j tiles beyond the cutoff;
//   large_finish_kernel  per-replica energies = fixed-order sum of the per-CTA FP64 partials;
//   large_vv1 / large_vv2 / large_observe   the velocity-Verlet halves, the Langevin kick (same Philox stream as the
//                        fused kernel: keyed by seed, global replica, atom, global step) and the observable record.
// Pair FORCES are the packed pair_eval2<MODE> of b200md_forces.cuh (two j atoms per FFMA2 issue slot, FP32).  Pair
// ENERGIES (ENERGY instantiation) are evaluated in FP64 from the fp32 positions, half shell (tiles above the diagonal
// and j > i inside the diagonal tile): the lj / electrostatics sums cancel too far for fp32 pair terms to meet the
// 1e-6-per-term contract.
#pragma once
#include "b200md_forces.cuh"

namespace b200md {

constexpr int kLThreads = 128;  // pair kernel: 4 warps on one i-block
constexpr int kLBondedThreads = 128;
constexpr double kLFixScale = 1073741824.0;  // 2^30 per kcal/mol/A: range +-8.6e9, resolution 9.3e-10
constexpr double kLFixInv = 1.0 / 1073741824.0;
constexpr double kLElec = 332.06371307417066;

struct LargeTopo {
  int natoms, ntypes, nwords;  // nwords = ceil(natoms / 32) = number of 32-atom blocks
  const int* atom_type;        // [32 nwords] (padding: 0)
  const float* charge;         // [32 nwords]
  const float* mass;           // [32 nwords] (padding: 1)
  const float* tab12a;         // [ntypes^2] 12 A
  const float* tab6b;          // [ntypes^2]  6 B
  const double2* tab64;        // [ntypes^2] (A, B) in double for the FP64 pair energies
  const uint32_t* mask;        // [32 nwords][nwords] bit j of word w: pair (i, 32 w + j) is NOT evaluated
  int nbonds, nangles, ndih, n14, nimp;
  const int2* bond_idx;   const float2* bond_prm;
  const int* angle_idx;   const float2* angle_prm;   // 3 ints per angle
  const int4* dih_idx;    const int* dih_start;  const float4* dih_term;   // terms of dihedral d: [dih_start[d], dih_start[d+1]); (k, phi0, per, kind)
  const double2* dih_cs;  const double2* imp_cs;                          // (cos phi0, sin phi0) per term; kind: 0 AMBER, 1 harmonic, 2 AMBER with integer per 1..6
  const int2* idx14;      const float4* prm14;                            // A, B, scnb, scee
  const int4* imp_idx;    const int* imp_start;  const float4* imp_term;
};

struct d3 { double x, y, z; };
__device__ __forceinline__ d3 dsub(d3 a, d3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ d3 dcross(d3 a, d3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
__device__ __forceinline__ double ddot(d3 a, d3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ d3 dscale(d3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ d3 lpos(const float* p, int i) { return {(double)p[3 * i], (double)p[3 * i + 1], (double)p[3 * i + 2]}; }
// wrap_dist (forces.py:358-363): every component, unless the whole box diagonal is zero
__device__ __forceinline__ d3 lwrap(d3 d, const double* b, bool vac) {
  if (vac) return d;
  d.x -= b[0] * rint(d.x / b[0]); d.y -= b[1] * rint(d.y / b[1]); d.z -= b[2] * rint(d.z / b[2]);
  return d;
}
__device__ __forceinline__ void ladd(long long* f, int i, d3 a, double s) {
  atomicAdd((unsigned long long*)(f + 3 * i), (unsigned long long)__double2ll_rn(s * a.x * kLFixScale));
  atomicAdd((unsigned long long*)(f + 3 * i + 1), (unsigned long long)__double2ll_rn(s * a.y * kLFixScale));
  atomicAdd((unsigned long long*)(f + 3 * i + 2), (unsigned long long)__double2ll_rn(s * a.z * kLFixScale));
}

// evaluate_torsion (forces.py:540-602) for one quadruple with its terms; returns the energy.
// Terms of the AMBER branch with an integer periodicity 1..6 (marker 2) never form phi: cos / sin(n phi) come from
// (cos phi + i sin phi)^n and cos / sin(phi0) from the host (FP64), which takes atan2, sin and cos out of the common case
// (the FP64 transcendentals were most of this kernel's instructions).  Other terms use phi = -atan2(sin, cos) as written.
__device__ __forceinline__ double ltorsion(const float* p, const double* bx, bool vac, int4 q, const float4* term,
                                           const double2* term_cs, int t0, int t1, long long* f) {
  const d3 r12 = lwrap(dsub(lpos(p, q.x), lpos(p, q.y)), bx, vac);
  const d3 r23 = lwrap(dsub(lpos(p, q.y), lpos(p, q.z)), bx, vac);
  const d3 r34 = lwrap(dsub(lpos(p, q.z), lpos(p, q.w)), bx, vac);
  const d3 cA = dcross(r12, r23), cB = dcross(r23, r34), cC = dcross(r23, cA);
  const double nA = sqrt(ddot(cA, cA)), nB = sqrt(ddot(cB, cB)), nC = sqrt(ddot(cC, cC));
  const d3 ncB = dscale(cB, 1.0 / nB);
  const double cosPhi = ddot(cA, ncB) / nA, sinPhi = ddot(cC, ncB) / nC;  // forces.py:548-551; phi = -atan2(sinPhi, cosPhi)
  const double hinv = 1.0 / sqrt(cosPhi * cosPhi + sinPhi * sinPhi);
  const double c1 = cosPhi * hinv, s1 = -sinPhi * hinv;  // cos phi, sin phi
  double phi = 0.0;
  bool have_phi = false;
  double pot = 0.0, coeff = 0.0;
  for (int t = t0; t < t1; ++t) {
    const float4 tp = __ldg(term + t);
    const double k0 = tp.x, phi0 = tp.y, per = tp.z;
    if (tp.w == 2.0f) {  // AMBER branch, integer periodicity: k (1 + cos(n phi - phi0)), forces.py:565-569
      const double2 cs0 = term_cs[t];
      const int n = (int)tp.z;
      double cn = c1, sn = s1;
      for (int k = 1; k < n; ++k) {
        const double cc = cn * c1 - sn * s1;
        sn = sn * c1 + cn * s1;
        cn = cc;
      }
      pot += k0 * (1.0 + (cn * cs0.x + sn * cs0.y));
      coeff += -per * k0 * (sn * cs0.x - cn * cs0.y);
      continue;
    }
    if (!have_phi) { phi = -atan2(sinPhi, cosPhi); have_phi = true; }
    if (tp.w == 0.0f) {  // AMBER branch, any periodicity
      const double ad = per * phi - phi0;
      pot += k0 * (1.0 + cos(ad));
      coeff += -per * k0 * sin(ad);
    } else {  // CHARMM harmonic branch, forces.py:570-576
      double ad = phi - phi0;
      if (ad < -M_PI) ad += 2.0 * M_PI; else if (ad > M_PI) ad -= 2.0 * M_PI;
      pot += k0 * ad * ad;
      coeff += 2.0 * k0 * ad;
    }
  }
  const double nd2 = sqrt(ddot(r23, r23)), n2d2 = nd2 * nd2;  // OpenMM projection, forces.py:582-600
  const double ff0 = (-coeff * nd2) / (nA * nA), ff1 = ddot(r12, r23) / n2d2, ff2 = ddot(r34, r23) / n2d2;
  const double ff3 = (coeff * nd2) / (nB * nB);
  const d3 f0 = dscale(cA, ff0), f3 = dscale(cB, ff3);
  const d3 s = dsub(dscale(f0, ff1), dscale(f3, ff2));
  ladd(f, q.x, f0, -1.0);
  ladd(f, q.y, {f0.x + s.x, f0.y + s.y, f0.z + s.z}, 1.0);
  ladd(f, q.z, {f3.x - s.x, f3.y - s.y, f3.z - s.z}, 1.0);
  ladd(f, q.w, f3, -1.0);
  return pot;
}

// FP64 block reduction of `n` values per thread in a fixed order; result valid in thread 0
template <int NV, int THREADS>
__device__ __forceinline__ void lreduce(double (&e)[NV], double* sred /* [THREADS / 32][NV] */, int tid) {
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    double v = e[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    e[k] = v;
  }
  if ((tid & 31) == 0) {
#pragma unroll
    for (int k = 0; k < NV; ++k) sred[(tid >> 5) * NV + k] = e[k];
  }
  __syncthreads();
  if (tid == 0) {
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      double v = 0.0;
      for (int w = 0; w < THREADS / 32; ++w) v += sred[w * NV + k];
      e[k] = v;
    }
  }
  __syncthreads();
}

// grid (nblk, min(R, 65535)); partial[(r * npart + blockIdx.x) * 8 + slot]
__global__ void __launch_bounds__(kLBondedThreads)
large_bonded_kernel(const LargeTopo T, const ForceParams P, const long long R, const float* __restrict__ pos,
                    const float* __restrict__ box, long long* __restrict__ fbond, double* __restrict__ partial,
                    const int npart, const float* __restrict__ smd_centers, const double smd_center, double* __restrict__ smd_rec) {
  __shared__ double sred[(kLBondedThreads / 32) * B200MD_NUM_ENERGIES];
  const int tid = threadIdx.x;
  const int N = T.natoms;
  const uint32_t terms = P.terms;
  const int nb = (terms & B200MD_T_BONDS) ? T.nbonds : 0;
  const int na = (terms & B200MD_T_ANGLES) ? T.nangles : 0;
  const int nd = (terms & B200MD_T_DIHEDRALS) ? T.ndih : 0;
  const bool lj14 = terms & B200MD_T_LJ, el14 = terms & B200MD_T_ELECTROSTATICS;
  const int n4 = ((terms & B200MD_T_14) && (lj14 || el14)) ? T.n14 : 0;
  const int ni = (terms & B200MD_T_IMPROPERS) ? T.nimp : 0;
  const int e1 = nb, e2 = e1 + na, e3 = e2 + nd, e4 = e3 + n4, e5 = e4 + ni;
  for (long long r = blockIdx.y; r < R; r += gridDim.y) {
    const float* p = pos + r * N * 3;
    long long* f = fbond + r * N * 3;
    double bx[3] = {0.0, 0.0, 0.0};
    if (box) { bx[0] = __ldg(box + r * 9); bx[1] = __ldg(box + r * 9 + 4); bx[2] = __ldg(box + r * 9 + 8); }
    const bool vac = bx[0] == 0.0 && bx[1] == 0.0 && bx[2] == 0.0;
    double e[B200MD_NUM_ENERGIES];
#pragma unroll
    for (int k = 0; k < B200MD_NUM_ENERGIES; ++k) e[k] = 0.0;
    for (int g = blockIdx.x * kLBondedThreads + tid; g < e5; g += gridDim.x * kLBondedThreads) {
      if (g < e1) {  // forces.py:111-132, 492-501
        const int2 ij = __ldg(T.bond_idx + g);
        const float2 kp = __ldg(T.bond_prm + g);
        const d3 d = lwrap(dsub(lpos(p, ij.x), lpos(p, ij.y)), bx, vac);
        const double dist = sqrt(ddot(d, d));
        if (P.has_cutoff && !(dist <= (double)P.rc)) continue;
        const double x = dist - (double)kp.y;
        e[B200MD_E_BONDS] += (double)kp.x * x * x;
        const double c = 2.0 * (double)kp.x * x / dist;
        ladd(f, ij.x, d, -c);
        ladd(f, ij.y, d, c);
      } else if (g < e2) {  // forces.py:134-145, 504-537
        const int a = g - e1;
        const int q0 = __ldg(T.angle_idx + 3 * a), q1 = __ldg(T.angle_idx + 3 * a + 1), q2 = __ldg(T.angle_idx + 3 * a + 2);
        const float2 kp = __ldg(T.angle_prm + a);
        const d3 r21 = lwrap(dsub(lpos(p, q0), lpos(p, q1)), bx, vac);
        const d3 r23 = lwrap(dsub(lpos(p, q2), lpos(p, q1)), bx, vac);
        const double n23i = 1.0 / sqrt(ddot(r23, r23)), n21i = 1.0 / sqrt(ddot(r21, r21));
        double ct = ddot(r23, r21) * n21i * n23i;
        ct = ct < -1.0 ? -1.0 : (ct > 1.0 ? 1.0 : ct);
        const double dth = acos(ct) - (double)kp.y;
        e[B200MD_E_ANGLES] += (double)kp.x * dth * dth;
        const double st = sqrt(1.0 - ct * ct);
        const double coef = st != 0.0 ? -2.0 * (double)kp.x * dth / st : 0.0;
        const d3 f0 = dscale(dsub(dscale(r21, ct * n21i), dscale(r23, n23i)), coef * n21i);
        const d3 f2 = dscale(dsub(dscale(r23, ct * n23i), dscale(r21, n21i)), coef * n23i);
        ladd(f, q0, f0, 1.0);
        ladd(f, q2, f2, 1.0);
        ladd(f, q1, {f0.x + f2.x, f0.y + f2.y, f0.z + f2.z}, -1.0);
      } else if (g < e3) {  // forces.py:147-174
        const int d = g - e2;
        e[B200MD_E_DIHEDRALS] += ltorsion(p, bx, vac, __ldg(T.dih_idx + d), T.dih_term, T.dih_cs, __ldg(T.dih_start + d), __ldg(T.dih_start + d + 1), f);
      } else if (g < e4) {  // forces.py:176-226: never cut, switched or reaction-field corrected
        const int k = g - e3;
        const int2 ij = __ldg(T.idx14 + k);
        const float4 q = __ldg(T.prm14 + k);
        const d3 d = lwrap(dsub(lpos(p, ij.x), lpos(p, ij.y)), bx, vac);
        const double r2 = ddot(d, d), rinv2 = 1.0 / r2, rinv6 = rinv2 * rinv2 * rinv2, rinv1 = sqrt(rinv2);
        double c = 0.0;  // -(dE/dr) / r
        if (lj14) {
          e[B200MD_E_LJ] += ((double)q.x * rinv6 * rinv6 - (double)q.y * rinv6) / (double)q.z;
          c += (12.0 * (double)q.x * rinv6 * rinv6 - 6.0 * (double)q.y * rinv6) * rinv2 / (double)q.z;
        }
        if (el14) {
          const double pot = kLElec * (double)__ldg(T.charge + ij.x) * (double)__ldg(T.charge + ij.y) * rinv1 / (double)q.w;
          e[B200MD_E_ELECTROSTATICS] += pot;
          c += pot * rinv2;
        }
        ladd(f, ij.x, d, c);
        ladd(f, ij.y, d, -c);
      } else {  // forces.py:228-255
        const int d = g - e4;
        e[B200MD_E_IMPROPERS] += ltorsion(p, bx, vac, __ldg(T.imp_idx + d), T.imp_term, T.imp_cs, __ldg(T.imp_start + d), __ldg(T.imp_start + d + 1), f);
      }
    }
    if (P.smd_on && blockIdx.x == 0 && tid == 0) {  // potcalc.py:75-89, not wrapped
      const double center = smd_centers ? (double)__ldg(smd_centers + r) : smd_center;
      const d3 d = dsub(lpos(p, P.smd_a2), lpos(p, P.smd_a1));
      const double dist = sqrt(ddot(d, d)), dev = center - dist;
      e[B200MD_E_EXTERNAL] += 0.5 * (double)P.smd_fk * dev * dev;
      const double c = (double)P.smd_fk * dev / dist;
      ladd(f, P.smd_a2, d, c);
      ladd(f, P.smd_a1, d, -c);
      if (smd_rec) {  // [R][2]: end-to-end distance of this step, work since the first step of the call
        smd_rec[2 * r] = dist;
        smd_rec[2 * r + 1] += (double)P.smd_fk * dev * (double)P.smd_rate;
      }
    }
    lreduce<B200MD_NUM_ENERGIES, kLBondedThreads>(e, sred, tid);
    if (tid == 0) {
      double* o = partial + (r * npart + blockIdx.x) * B200MD_NUM_ENERGIES;
#pragma unroll
      for (int k = 0; k < B200MD_NUM_ENERGIES; ++k) o[k] = e[k];
    }
  }
}

// Axis-aligned bounding box of every 32-atom tile of every replica (cutoff mode, vacuum replicas): lets a warp of the
// pair kernel drop a j tile none of its i atoms can reach.  One warp per tile; bbox[(r * nwords + tile) * 8 + {0..2, 4..6}] = lo, hi.
__global__ void __launch_bounds__(128)
large_bbox_kernel(const LargeTopo T, const long long R, const float* __restrict__ pos, float* __restrict__ bbox) {
  const int lane = threadIdx.x & 31, tile = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (tile >= T.nwords) return;
  const int j = 32 * tile + lane;
  for (long long r = blockIdx.y; r < R; r += gridDim.y) {
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    if (j < T.natoms) {
#pragma unroll
      for (int c = 0; c < 3; ++c) lo[c] = hi[c] = __ldg(pos + (r * T.natoms + j) * 3 + c);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        lo[c] = fminf(lo[c], __shfl_xor_sync(0xffffffffu, lo[c], o));
        hi[c] = fmaxf(hi[c], __shfl_xor_sync(0xffffffffu, hi[c], o));
      }
    if (lane < 3) {
      bbox[(r * T.nwords + tile) * 8 + lane] = lane == 0 ? lo[0] : (lane == 1 ? lo[1] : lo[2]);
      bbox[(r * T.nwords + tile) * 8 + 4 + lane] = lane == 0 ? hi[0] : (lane == 1 ? hi[1] : hi[2]);
    }
  }
}

template <int OFF>
__device__ __forceinline__ float lds_f32(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1+%2];" : "=f"(v) : "r"(addr), "n"(OFF));
  return v;
}
// Row stride of the shared-memory type-pair table in entries: odd, so that the (12 A, -6 B) entries of one column sit
// in distinct banks for up to 16 row types (lanes differ in their row = type of i, the column = type of j is uniform)
__host__ __device__ inline int large_tab_stride(int ntypes) { return ntypes | 1; }

// grid (nwords, min(R, 65535)).  Shared memory: the two type-pair tables (2 ntypes^2 floats, dynamic) + statics.
// The j tile of a warp is staged structure-of-arrays (x[32], y[32], z[32], q[32], type[32]) so that one broadcast LDS.64
// delivers the operands of TWO j atoms as a packed f32x2: the pair arithmetic is pair_eval2 (FFMA2 / FMUL2 / FADD2, one
// issue slot per two pairs), like the on-chip tiles.
template <int MODE, bool ENERGY, bool PBC>
__device__ __forceinline__ void large_pairs_body(const LargeTopo& T, const ForceParams& P, const float* __restrict__ p,
                                                 const uint32_t tbase, float* sj /* [4][32] */, int* stj, double* sjd /* [4][32] */,
                                                 const float3 nbox, const float3 ibox, int ib, int warp, int lane, float xi,
                                                 float yi, float zi, float kqi, const float* __restrict__ bb, float& fx,
                                                 float& fy, float& fz, double& eel, double& elj) {
  const int N = T.natoms;
  const uint32_t* mrow = T.mask + (size_t)(32 * ib + lane) * T.nwords;
  const PairConst C = make_pair_const<MODE>(P);
  const u64 xi2 = dup2(xi), yi2 = dup2(yi), zi2 = dup2(zi), kq2 = dup2(kqi);
  u64 ax = 0ull, ay = 0ull, az = 0ull;
  // the j atoms, their table offsets and the mask word of the NEXT tile are fetched while the current one is evaluated
  float jx = 0.f, jy = 0.f, jz = 0.f, jq = 0.f;
  int jtype = 0;
  uint32_t mn = 0xffffffffu;
  // cutoff mode, vacuum: the tile's bounding box against this lane's i atom.  Component-wise |xi - box| <= |xi - xj| for
  // every atom of the tile and fp32 subtraction / fma are monotonic, so d2 <= r2 of every pair of the lane: a tile that all
  // lanes place beyond the cutoff holds only pairs the cutoff test of pair_eval2 would zero -- results are bit-identical
  const bool boxtest = MODE == 1 && !PBC && bb != nullptr;
  float4 blo = make_float4(0.f, 0.f, 0.f, 0.f), bhi = blo;
  const bool live_i = 32 * ib + lane < N;
  auto fetch = [&](int jt) {
    if (boxtest) { blo = __ldg((const float4*)(bb + 8 * jt)); bhi = __ldg((const float4*)(bb + 8 * jt + 4)); }
    const int j = 32 * jt + lane;
    jx = jy = jz = jq = 0.f;
    if (j < N) { jx = __ldg(p + 3 * j); jy = __ldg(p + 3 * j + 1); jz = __ldg(p + 3 * j + 2); jq = __ldg(T.charge + j) * P.el_on; }
    jtype = __ldg(T.atom_type + j) * 8;  // byte offset of the (12 A, -6 B) entry in this lane's table row
    mn = __ldg(mrow + jt);
  };
  if (warp < T.nwords) fetch(warp);
  for (int jt = warp; jt < T.nwords; jt += kLThreads / 32) {
    const uint32_t m = mn;
    bool far = false;
    if (boxtest) {
      const float ex = fmaxf(fmaxf(__fsub_rn(blo.x, xi), __fsub_rn(xi, bhi.x)), 0.0f), ey = fmaxf(fmaxf(__fsub_rn(blo.y, yi), __fsub_rn(yi, bhi.y)), 0.0f);
      const float ez = fmaxf(fmaxf(__fsub_rn(blo.z, zi), __fsub_rn(zi, bhi.z)), 0.0f);
      far = !live_i || !(__fmaf_rn(ez, ez, __fmaf_rn(ey, ey, __fmul_rn(ex, ex))) <= P.rc2);
    }
    __syncwarp();
    sj[lane] = jx; sj[32 + lane] = jy; sj[64 + lane] = jz; sj[96 + lane] = jq;
    stj[lane] = jtype;
    if (ENERGY && jt >= ib) { sjd[lane] = (double)jx; sjd[32 + lane] = (double)jy; sjd[64 + lane] = (double)jz; sjd[96 + lane] = (double)jq; }
    __syncwarp();
    if (jt + kLThreads / 32 < T.nwords) fetch(jt + kLThreads / 32);
    if (__all_sync(0xffffffffu, far || m == 0xffffffffu)) continue;
#pragma unroll 4
    for (int jj = 0; jj < 32; jj += 2) {
      const u64 xj = *(const u64*)(sj + jj), yj = *(const u64*)(sj + 32 + jj), zj = *(const u64*)(sj + 64 + jj);
      const u64 qj = *(const u64*)(sj + 96 + jj);
      const int2 tj = *(const int2*)(stj + jj);
      const u64 a12 = pk(lds_f32<0>(tbase + tj.x), lds_f32<0>(tbase + tj.y)), nb6 = pk(lds_f32<4>(tbase + tj.x), lds_f32<4>(tbase + tj.y));
      u64 dx = sub2(xi2, xj), dy = sub2(yi2, yj), dz = sub2(zi2, zj);
      if (PBC) { dx = wrap2(dx, nbox.x, ibox.x); dy = wrap2(dy, nbox.y, ibox.y); dz = wrap2(dz, nbox.z, ibox.z); }
      const u64 sc = pair_eval2<MODE>(dx, dy, dz, kqi, kq2, qj, a12, nb6, (m >> jj) & 1u, (m >> (jj + 1)) & 1u, P, C);
      ax = fma2(sc, dx, ax); ay = fma2(sc, dy, ay); az = fma2(sc, dz, az);
    }
    if (ENERGY && jt >= ib) {  // FP64 energies of the unordered pairs this tile owns (forces.py:379-413, 451-489 in double)
      const double xd = (double)xi, yd = (double)yi, zd = (double)zi;
      const double kq = kLElec * (double)(__ldg(T.charge + 32 * ib + lane) * P.el_on);
      const double rc2 = P.rc_d * P.rc_d, ta = (double)P.tab_a, tb = (double)P.tab_b;
      const double2* trow = T.tab64 + (size_t)__ldg(T.atom_type + 32 * ib + lane) * T.ntypes;
      const int jfirst = jt == ib ? lane + 1 : 0;  // diagonal tile: j > i
      for (int jj = 0; jj < 32; ++jj) {
        if (jj < jfirst || ((m >> jj) & 1u)) continue;
        double dx = xd - sjd[jj], dy = yd - sjd[32 + jj], dz = zd - sjd[64 + jj];
        if (PBC) {
          const double bx = -(double)nbox.x, by = -(double)nbox.y, bz = -(double)nbox.z;
          if (bx != 0.0) dx -= bx * rint(dx / bx);
          if (by != 0.0) dy -= by * rint(dy / by);
          if (bz != 0.0) dz -= bz * rint(dz / bz);
        }
        const double r2 = dx * dx + dy * dy + dz * dz;
        if (MODE == 1 && !(r2 <= rc2)) continue;
        const double rinv = rsqrt(r2), r2inv = rinv * rinv, r6 = r2inv * r2inv * r2inv;
        const double2 ab = __ldg(trow + (stj[jj] >> 3));
        const double qq = kq * sjd[96 + jj];
        double lj = (ta * ab.x * r6 + tb * ab.y) * r6;
        if (MODE == 1) {
          const double r = r2 * rinv;
          if (r > P.rs_d) {
            const double t = (r - P.rs_d) * P.inv_sw_d;
            lj *= 1.0 + t * t * t * (-10.0 + t * (15.0 - t * 6.0));
          }
          eel += qq * (rinv + P.krf_d * r2 - P.crf_d);
        } else {
          eel += qq * rinv;
        }
        elj += lj;
      }
    }
  }
  fx = hsum2(ax); fy = hsum2(ay); fz = hsum2(az);
}

template <int MODE, bool ENERGY>
__global__ void __launch_bounds__(kLThreads)
large_pairs_kernel(const LargeTopo T, const ForceParams P, const long long R, const float* __restrict__ pos,
                   const float* __restrict__ box, const long long* __restrict__ fbond, float* __restrict__ forces,
                   double* __restrict__ partial, const int npart, const int part0, uint32_t* __restrict__ status,
                   const float* __restrict__ bbox) {
  extern __shared__ __align__(16) char lsmem[];
  float* sAB = (float*)lsmem;  // [ntypes][stride] entries (12 A tab_a, 6 B tab_b)
  const int tstride = large_tab_stride(T.ntypes);
  __shared__ __align__(16) float sj[kLThreads / 32][4 * 32];
  __shared__ __align__(16) int stj[kLThreads / 32][32];
  __shared__ double sjd[ENERGY ? kLThreads / 32 : 1][ENERGY ? 4 * 32 : 1];
  __shared__ float sf[kLThreads / 32][3][32];
  __shared__ double sred[(kLThreads / 32) * 2];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int N = T.natoms, ib = blockIdx.x;
  for (int k = tid; k < T.ntypes * T.ntypes; k += kLThreads) {
    const int t = k / T.ntypes, u = k - t * T.ntypes;
    sAB[2 * (t * tstride + u)] = __ldg(T.tab12a + k) * P.tab_a;
    sAB[2 * (t * tstride + u) + 1] = __ldg(T.tab6b + k) * P.tab_b;
  }
  const int i = 32 * ib + lane;
  const bool live = i < N;
  const uint32_t tbase = (uint32_t)__cvta_generic_to_shared(sAB) + (uint32_t)(__ldg(T.atom_type + i) * tstride * 8);
  const float kqi = __fmul_rn(__ldg(T.charge + i) * P.el_on, kElecFactor);
  const bool nonbonded = (P.terms & (B200MD_T_ELECTROSTATICS | B200MD_T_LJ | B200MD_T_REPULSION | B200MD_T_REPULSIONCG)) != 0;
  __syncthreads();
  for (long long r = blockIdx.y; r < R; r += gridDim.y) {
    const float* p = pos + r * N * 3;
    float xi = 0.f, yi = 0.f, zi = 0.f;
    if (live) { xi = __ldg(p + 3 * i); yi = __ldg(p + 3 * i + 1); zi = __ldg(p + 3 * i + 2); }
    if (warp == 0 && live && status != nullptr && (xi != xi || yi != yi || zi != zi)) atomicOr(status + r, B200MD_FLAG_NAN_POS);
    const float b0 = box ? __ldg(box + r * 9) : 0.0f, b1 = box ? __ldg(box + r * 9 + 4) : 0.0f, b2 = box ? __ldg(box + r * 9 + 8) : 0.0f;
    const bool pbc = b0 != 0.0f || b1 != 0.0f || b2 != 0.0f;  // forces.py:359
    const float3 nbox = make_float3(-b0, -b1, -b2);
    const float3 ibox = make_float3(b0 != 0.0f ? 1.0f / b0 : 0.0f, b1 != 0.0f ? 1.0f / b1 : 0.0f, b2 != 0.0f ? 1.0f / b2 : 0.0f);
    float fx = 0.f, fy = 0.f, fz = 0.f;
    double e[2] = {0.0, 0.0};
    if (nonbonded) {
      double* sd = sjd[ENERGY ? warp : 0];
      if (pbc) large_pairs_body<MODE, ENERGY, true>(T, P, p, tbase, sj[warp], stj[warp], sd, nbox, ibox, ib, warp, lane, xi, yi, zi, kqi, nullptr, fx, fy, fz, e[0], e[1]);
      else large_pairs_body<MODE, ENERGY, false>(T, P, p, tbase, sj[warp], stj[warp], sd, nbox, ibox, ib, warp, lane, xi, yi, zi, kqi,
                                                 bbox ? bbox + r * T.nwords * 8 : nullptr, fx, fy, fz, e[0], e[1]);
    }
    sf[warp][0][lane] = fx; sf[warp][1][lane] = fy; sf[warp][2][lane] = fz;
    __syncthreads();
    if (tid < 96) {  // thread (c, lane): fixed-order sum over the warps + the bonded accumulator, one coalesced row
      const int a = tid / 3, c = tid - 3 * a, ia = 32 * ib + a;
      if (ia < N) {
        const float s = __fadd_rn(__fadd_rn(sf[0][c][a], sf[1][c][a]), __fadd_rn(sf[2][c][a], sf[3][c][a]));
        // fbond == nullptr: the bonded kernel runs concurrently on a second stream, large_combine_kernel adds it afterwards
        const double fb = fbond != nullptr ? (double)fbond[(r * N + ia) * 3 + c] * kLFixInv : 0.0;
        forces[(r * N + ia) * 3 + c] = fbond != nullptr ? (float)((double)s + fb) : s;
      }
    }
    if (ENERGY) {
      lreduce<2, kLThreads>(e, sred, tid);
      if (tid == 0) {
        double* o = partial + (r * npart + part0 + ib) * B200MD_NUM_ENERGIES;
        o[B200MD_E_ELECTROSTATICS] = e[0]; o[B200MD_E_LJ] = e[1];
#pragma unroll
        for (int k = 2; k < B200MD_NUM_ENERGIES; ++k) o[k] = 0.0;
      }
    } else {
      __syncthreads();
    }
  }
}

// forces = pair row (fp32) + bonded accumulator: the same expression, hence the same bits, as the fused epilogue above
__global__ void large_combine_kernel(const long long n, const long long* __restrict__ fbond, float* __restrict__ forces) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (g < n) forces[g] = (float)((double)forces[g] + (double)fbond[g] * kLFixInv);
}

// energies[r][k] = sum over the npart per-CTA partials, in index order
__global__ void large_finish_kernel(const long long R, const double* __restrict__ partial, const int npart,
                                    double* __restrict__ energies) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= R * B200MD_NUM_ENERGIES) return;
  const long long r = g / B200MD_NUM_ENERGIES;
  const int k = (int)(g - r * B200MD_NUM_ENERGIES);
  double s = 0.0;
  for (int q = 0; q < npart; ++q) s += partial[(r * npart + q) * B200MD_NUM_ENERGIES + k];
  energies[g] = s;
}

// _first_VV of upstream torchmd.integrator: vel += 0.5 dt f / m ; pos += dt vel (one thread per atom)
__global__ void large_vv1_kernel(const LargeTopo T, const float dt, const long long R, float* __restrict__ pos,
                                 float* __restrict__ vel, const float* __restrict__ forces) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= R * T.natoms) return;
  const int a = (int)(g % T.natoms);
  const float hdt_m = 0.5f * dt * (1.0f / __ldg(T.mass + a));
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float v = fmaf(hdt_m, forces[3 * g + c], vel[3 * g + c]);
    vel[3 * g + c] = v;
    pos[3 * g + c] = fmaf(dt, v, pos[3 * g + c]);
  }
}

// Langevin kick (vel += -gamma vel dt + vcoeff N(0,1)) and _second_VV; with `ekin` the per-block kinetic-energy
// partials of a report step.  grid (ceil(N / 256), R): blocks do not straddle replicas.
__global__ void __launch_bounds__(256)
large_vv2_kernel(const LargeTopo T, const StepParams I, const unsigned long long gstep, const long long R,
                 float* __restrict__ vel, const float* __restrict__ forces, float* __restrict__ pos,
                 double* __restrict__ ekin, uint32_t* __restrict__ status) {
  __shared__ double sred[8];
  const int tid = threadIdx.x, a = blockIdx.x * 256 + tid;
  for (long long r = blockIdx.y; r < R; r += gridDim.y) {
    double e[1] = {0.0};
    if (a < T.natoms) {
      const long long g = r * T.natoms + a;
      const float mass = __ldg(T.mass + a), invm = 1.0f / mass, hdt_m = 0.5f * I.dt * invm;
      float v[3] = {vel[3 * g], vel[3 * g + 1], vel[3 * g + 2]};
      if (I.temperature > 0.0f) {
        const float vcoef = sqrtf(2.0f * I.gamma * invm * kBoltzman * I.temperature * I.dt), vdamp = I.gamma * I.dt;
        float n[3];
        normal3(I.seed, (unsigned long long)(I.replica0 + r), (uint32_t)a, gstep, n);
#pragma unroll
        for (int c = 0; c < 3; ++c) v[c] += fmaf(vcoef, n[c], -vdamp * v[c]);
      }
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        v[c] = fmaf(hdt_m, forces[3 * g + c], v[c]);
        vel[3 * g + c] = v[c];
      }
      e[0] = 0.5 * (double)mass * ((double)v[0] * v[0] + (double)v[1] * v[1] + (double)v[2] * v[2]);
      const float x = pos[3 * g], y = pos[3 * g + 1], z = pos[3 * g + 2];
      if (status != nullptr && (x != x || y != y || z != z)) atomicOr(status + r, B200MD_FLAG_NAN_POS);
    }
    if (ekin != nullptr) {
      lreduce<1, 256>(e, sred, tid);
      if (tid == 0) ekin[r * gridDim.x + blockIdx.x] = e[0];
    }
  }
}

// one observable record per replica: 8 energies, Ekin (fixed-order sum of the block partials), SMD distance and work
__global__ void large_observe_kernel(const long long R, const double* __restrict__ energies, const double* __restrict__ ekin,
                                     const int nek, const double* __restrict__ smd_rec, double* __restrict__ out) {
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= R) return;
  double* o = out + r * B200MD_NUM_OBSERVABLES;
  for (int k = 0; k < B200MD_NUM_ENERGIES; ++k) o[k] = energies[r * B200MD_NUM_ENERGIES + k];
  double s = 0.0;
  for (int q = 0; q < nek; ++q) s += ekin[r * nek + q];
  o[B200MD_OBS_EKIN] = s;
  o[B200MD_OBS_E2E_DIST] = smd_rec ? smd_rec[2 * r] : 0.0;
  o[B200MD_OBS_SMD_WORK] = smd_rec ? smd_rec[2 * r + 1] : 0.0;
}

}  // namespace b200md
