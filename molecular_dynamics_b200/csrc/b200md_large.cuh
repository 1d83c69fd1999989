// b200md_large.cuh -- solutes above 128 atoms (SURVEY 8f N4 "larger solutes": the adk-1.0.prmtop system named at
// V_Calculations/Test-10_diALA/potcalc.py:114).  Same reference semantics as the one-CTA-per-replica kernels
// (TorchMD_Fix/forces.py Forces.compute :82-344, upstream Integrator.step), but a replica no longer fits on chip:
//
//   large_pairs_kernel          LJ + Coulomb forces, HALF shell: CTA = one rectangle (BI >= BJ) of two blocks of <= 288
//                               atoms of one replica, both staged in shared memory; a warp holds a chunk of 8 j atoms as
//                               broadcast operands and meets it with 32-atom i tiles (lane = i atom) in packed FP32
//                               (pair_eval2 of b200md_forces.cuh); i-side force -> the warp's shared-memory row, j-side
//                               force -> 24 registers over the passes, transposed once per chunk by a shuffle butterfly.
//                               Cutoff mode: tile bounding boxes give each chunk its pass list by one ballot, a listed
//                               pass still drops out on a warp vote (bit-identical with and without the skips);
//   large_bonded_forces_kernel  bonds, angles, torsions with all their terms, 1-4 pairs, impropers, the SMD restraint in
//                               fp32: CTA = one window of 64 atoms, every item that touches the window is evaluated into
//                               private shared-memory slots (no atomics), thread a sums atom a's slots in a fixed order;
//   -> both write PARTIAL force rows fpart[replica][slot][atom][3] (pairs: slot = the partner block, bonded: the last
//      slot; every slot of every atom is written exactly once per call), so the two launches are independent and run
//      beside each other on two streams; the force row is the fixed-order sum of the slots, taken by the consumer:
//   large_sum_kernel            Forces.compute;
//   large_vv2_kernel            the integrator: Langevin kick + second half-kick of step s, first half-kick + drift of step
//                               s + 1 and the tile bounding boxes of the moved atoms in ONE launch between two force
//                               evaluations (large_vv1_kernel only opens a call and follows snapshot steps);
//   large_bonded_energy_kernel, large_pair_energy_kernel   report steps / Forces.compute with energies: every term in FP64
//                               from the fp32 positions, item- / tile-centric, no forces; they only read the positions
//                               and run on the second stream; large_finish_kernel sums the per-CTA partials in index order;
//   large_bbox_kernel, large_observe_kernel   tile boxes of a Forces.compute call; the observable record.
//   large_shake_kernel, large_rattle_kernel, large_ekin_kernel   rigid X-H bonds: one thread per hydrogen group (the group
//                               solvers of the on-chip path on registers), the kinetic energy of the constrained velocities;
//   large_steps_kernel          few replicas: runs of steps between two report steps as ONE cooperative launch.
// The Langevin noise is the same Philox stream as the fused kernel (keyed by seed, global replica, atom, global step).
// Energies are FP64 because the lj / electrostatics sums cancel too far for fp32 pair terms to meet the 1e-6-per-term
// contract; forces are fp32 like the on-chip kernels.  Results are bit-reproducible and do not depend on
// the replica count of a call or on how often energies are recorded (every summation order is fixed by the topology).
// Round 1 for comparison: full-shell 32 x 32 tiles (every pair evaluated from both atoms), one FP64 thread per bonded item
// with int64 global atomics, 4-6 launches per step: 397 / 246 us per Langevin step of 256 x 1 040 atoms (plain / cutoff
// 9 A); now 255 / 120 us (profiles/r02_large_solute.md).
#pragma once
#include <cooperative_groups.h>

#include "b200md_forces.cuh"

namespace b200md {

// Loads of data another CTA may have written earlier in the SAME launch (positions, tile boxes, partial rows in the
// persistent multi-step kernel): COH = true reads through L2 (ld.global.cg); the per-launch kernels keep the read-only path.
template <bool COH> __device__ __forceinline__ float ldm(const float* a) { return COH ? __ldcg(a) : __ldg(a); }
template <bool COH> __device__ __forceinline__ float4 ldm4(const float4* a) { return COH ? __ldcg(a) : __ldg(a); }
constexpr int kLThreads = 128;  // pair kernels: 4 warps per CTA
constexpr int kLBondedThreads = 128;
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
  // bonded forces by WINDOWS of kLWin atoms (large_bonded_forces_kernel): window w evaluates the entries
  // [win_start[w], win_start[w + 1]) = every item that touches one of its atoms (index into the concatenated list
  // [bonds | angles | dihedrals | 1-4 | impropers]); win_slot holds, per role of the item, the shared-memory slot of that
  // contribution (0xffff: the atom belongs to another window).  Slots are atom-major: atom a of the window owns
  // [gstart[a], gstart[a + 1]) - gstart[first atom of the window], in the fixed order bonds, angles, dihedrals, 1-4, impropers.
  const int* win_start;      // [nwin + 1]
  const int4* win_kend;      // [nwin] first entry of the window's angles, dihedrals, 1-4 pairs, impropers (entries are kind-sorted)
  const int4* win_atoms;     // [entries] atom indices of the item
  const float4* win_prm;     // [entries] bond / angle: (k, x0); torsions: (first term, end term) as int bits; 1-4: (A / scnb, B / scnb, q_i q_j / scee, -)
  const ushort4* win_slot;   // [entries]
  const int* gstart;         // [natoms + 1]
  const int* win_lo;         // [nwin] first atom the window's items reference
  const int* win_n;          // [nwin] how many atoms from win_lo they span, 0 when more than kLHalo (then read from HBM)
  int nwin, win_slots_max, halo_max;  // windows; largest slot count / staged span of a window (shared memory = 12 bytes x each)
  int bt, nblk;                       // pair rectangles: tiles per atom block (<= kLBT), number of blocks
  const int* phi_start;               // [natoms + 1] dihedral-angle backward: items of atom a are [phi_start[a], phi_start[a + 1])
  const int* phi_items;               // dihedral << 2 | role
  // rigid X-H bonds (SHAKE / RATTLE): groups of a heavy atom and its 1 .. 3 hydrogens
  int ngroups, ncons, cmax;           // groups, constraints (-1: a group the solvers cannot hold), largest group
  const int4* cgroup;                 // (parent | n << 24, h0, h1, h2)
  const float4* cdist2;               // squared bond lengths
  const float4* cweight;              // 1 / mass of (parent, h0, h1, h2)
};
constexpr int kLWin = 64, kLHalo = 320;

// Bonded items are written once, on a scalar type T: T = float evaluates the FORCES (what every step needs; the
// one-thread-per-item kernel was FP64 throughout in round 1 and, at 109 us per step of 256 x 1 040 atoms, the largest
// kernel of a cutoff-mode step), T = double evaluates the ENERGIES of report steps (FORCE = false, NoSink).  A report
// step runs both, so the forces -- and with them the trajectory -- do not depend on how often energies are recorded.
template <typename T> struct v3t { T x, y, z; };
template <typename T> __device__ __forceinline__ v3t<T> vsub(v3t<T> a, v3t<T> b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
template <typename T> __device__ __forceinline__ v3t<T> vcross(v3t<T> a, v3t<T> b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
template <typename T> __device__ __forceinline__ T vdot(v3t<T> a, v3t<T> b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
template <typename T> __device__ __forceinline__ v3t<T> vscale(v3t<T> a, T s) { return {a.x * s, a.y * s, a.z * s}; }
template <typename T> __device__ __forceinline__ v3t<T> vpos(const float* p, int i) { return {(T)p[3 * i], (T)p[3 * i + 1], (T)p[3 * i + 2]}; }
// Where an item reads its atoms: the replica's rows in HBM, or (bonded force windows) 16-byte rows staged in shared memory,
// addressed as shared so that one LDS.128 delivers an atom (a generic pointer that is sometimes shared costs an LD each).
struct GlobalPos {
  const float* p;
  template <typename T> __device__ __forceinline__ v3t<T> get(int i) const { return {(T)p[3 * i], (T)p[3 * i + 1], (T)p[3 * i + 2]}; }
};
struct SharedPos {
  uint32_t base;  // shared byte address of atom 0's row (the staged span starts at atom lo: base = rows - 16 lo)
  template <typename T> __device__ __forceinline__ v3t<T> get(int i) const {
    float x, y, z, w;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(x), "=f"(y), "=f"(z), "=f"(w) : "r"(base + 16u * (uint32_t)i));
    return {(T)x, (T)y, (T)z};
  }
};
__device__ __forceinline__ float trint(float x) { return rintf(x); }
__device__ __forceinline__ double trint(double x) { return rint(x); }
// 1 / x, 1 / sqrt(x): double for the energy path, the hardware approximations with one Newton step where the
// value is raised to a power in float (the force path, like the on-chip kernels: ~1 ulp, a tenth of the instructions of the
// IEEE sequences nvcc emits without fast-math)
__device__ __forceinline__ double trcp(double x) { return 1.0 / x; }
__device__ __forceinline__ double trsqrt(double x) { return drsqrt(x); }  // hardware seed + one third-order step: < 1 ulp of FP64
__device__ __forceinline__ float trcp(float x) { const float y = fast_rcp(x); return y * (2.0f - x * y); }
__device__ __forceinline__ float trsqrt(float x) { return rsqrt_nr(x); }
// wrap_dist (forces.py:358-363): every component, unless the whole box diagonal is zero
template <typename T> __device__ __forceinline__ v3t<T> vwrap(v3t<T> d, const T* b, bool vac) {
  if (vac) return d;
  d.x -= b[0] * trint(d.x / b[0]); d.y -= b[1] * trint(d.y / b[1]); d.z -= b[2] * trint(d.z / b[2]);
  return d;
}
// Where the force vectors of an item go.  Round 1 added each to a per-atom int64 fixed-point accumulator with three 64-bit
// GLOBAL atomics (~18 M per step of 256 x 1 040 atoms: the kernel sat at ~95 us whatever its arithmetic cost).  Measured
// alternatives: private slots in global memory + a gather kernel (5.8 M scattered 12-byte stores and their re-read:
// 63 + 41 us); an atom-centric kernel that re-evaluates an item for every atom it touches (lanes of a warp sit at different
// kinds: 304 us).  Kept: slots in SHARED memory, one CTA per window of kLWin atoms -- the CTA evaluates every item that
// touches the window (items across a window edge are evaluated by both windows), writes the vectors of ITS atoms to their
// private slots, and sums each atom's slots in a fixed order: no atomic, no global scratch, exact fp32, any magnitude.
struct WindowSink {
  float4* slots;  // shared memory, one 16-byte slot per contribution (one STS.128 / LDS.128 each)
  ushort4 sl;    // slot of each role, 0xffff = not in this window
};
struct NoSink {};
template <typename T> __device__ __forceinline__ void lput(WindowSink& f, int role, v3t<T> a, T s) {
  const unsigned sidx = role == 0 ? f.sl.x : (role == 1 ? f.sl.y : (role == 2 ? f.sl.z : f.sl.w));
  if (sidx != 0xffffu) f.slots[sidx] = make_float4((float)(s * a.x), (float)(s * a.y), (float)(s * a.z), 0.0f);
}
template <typename T> __device__ __forceinline__ void lput(NoSink&, int, v3t<T>, T) {}

// evaluate_torsion (forces.py:540-602) for one quadruple with its terms; returns the energy, adds the forces when FORCE.
// Terms of the AMBER branch with an integer periodicity 1..6 (marker 2) never form phi: cos / sin(n phi) come from
// (cos phi + i sin phi)^n and cos / sin(phi0) from the host (FP64).  Other terms use phi = -atan2(sin, cos) as written.
template <typename T, bool FORCE, typename SINK, typename POS>
__device__ __forceinline__ T ltorsion(const POS& p, const T* bx, bool vac, int4 q, const float4* term,
                                      const double2* term_cs, int t0, int t1, SINK& f) {
  typedef v3t<T> V;
  const V r12 = vwrap(vsub(p.template get<T>(q.x), p.template get<T>(q.y)), bx, vac);
  const V r23 = vwrap(vsub(p.template get<T>(q.y), p.template get<T>(q.z)), bx, vac);
  const V r34 = vwrap(vsub(p.template get<T>(q.z), p.template get<T>(q.w)), bx, vac);
  const V cA = vcross(r12, r23), cB = vcross(r23, r34), cC = vcross(r23, cA);
  const T inA = trsqrt(vdot(cA, cA)), inB = trsqrt(vdot(cB, cB)), inC = trsqrt(vdot(cC, cC));
  const V ncB = vscale(cB, inB);
  const T cosPhi = vdot(cA, ncB) * inA, sinPhi = vdot(cC, ncB) * inC;  // forces.py:548-551; phi = -atan2(sinPhi, cosPhi)
  const T hinv = trsqrt(cosPhi * cosPhi + sinPhi * sinPhi);
  const T c1 = cosPhi * hinv, s1 = -sinPhi * hinv;  // cos phi, sin phi
  T phi = 0;
  bool have_phi = false;
  T pot = 0, coeff = 0;
  for (int t = t0; t < t1; ++t) {
    const float4 tp = __ldg(term + t);
    const T k0 = tp.x, phi0 = tp.y, per = tp.z;
    if (tp.w == 2.0f) {  // AMBER branch, integer periodicity: k (1 + cos(n phi - phi0)), forces.py:565-569
      const double2 cs0 = term_cs[t];
      const int n = (int)tp.z;
      T cn = c1, sn = s1;
      for (int k = 1; k < n; ++k) {
        const T cc = cn * c1 - sn * s1;
        sn = sn * c1 + cn * s1;
        cn = cc;
      }
      pot += k0 * ((T)1 + (cn * (T)cs0.x + sn * (T)cs0.y));
      coeff += -per * k0 * (sn * (T)cs0.x - cn * (T)cs0.y);
      continue;
    }
    if (!have_phi) { phi = -atan2(sinPhi, cosPhi); have_phi = true; }
    if (tp.w == 0.0f) {  // AMBER branch, any periodicity
      const T ad = per * phi - phi0;
      pot += k0 * ((T)1 + cos(ad));
      coeff += -per * k0 * sin(ad);
    } else {  // CHARMM harmonic branch, forces.py:570-576
      T ad = phi - phi0;
      if (ad < (T)-M_PI) ad += (T)(2.0 * M_PI); else if (ad > (T)M_PI) ad -= (T)(2.0 * M_PI);
      pot += k0 * ad * ad;
      coeff += (T)2 * k0 * ad;
    }
  }
  if (FORCE) {
    const T n23sq = vdot(r23, r23), in23 = trsqrt(n23sq), nd2 = n23sq * in23, in2d2 = in23 * in23;  // OpenMM projection, forces.py:582-600
    const T ff0 = (-coeff * nd2) * (inA * inA), ff1 = vdot(r12, r23) * in2d2, ff2 = vdot(r34, r23) * in2d2;
    const T ff3 = (coeff * nd2) * (inB * inB);
    const V f0 = vscale(cA, ff0), f3 = vscale(cB, ff3);
    const V s = vsub(vscale(f0, ff1), vscale(f3, ff2));
    lput<T>(f, 0, f0, (T)-1);
    lput<T>(f, 1, {f0.x + s.x, f0.y + s.y, f0.z + s.z}, (T)1);
    lput<T>(f, 2, {f3.x - s.x, f3.y - s.y, f3.z - s.z}, (T)1);
    lput<T>(f, 3, f3, (T)-1);
  }
  return pot;
}

// One bonded item g of the concatenated list [bonds | angles | dihedrals | 1-4 pairs | impropers] (EVERY kind present: the
// callers skip the kinds a call has switched off): adds its energy to the slot it is booked under, hands its force vectors
// to the sink when FORCE.
struct LItemRanges { int e1, e2, e3, e4, e5; bool lj14, el14; uint32_t on; /* bit k: kind k is evaluated */ };
__device__ __forceinline__ LItemRanges litem_ranges(const LargeTopo& T, uint32_t terms) {
  LItemRanges G;
  G.e1 = T.nbonds; G.e2 = G.e1 + T.nangles; G.e3 = G.e2 + T.ndih; G.e4 = G.e3 + T.n14; G.e5 = G.e4 + T.nimp;
  G.lj14 = terms & B200MD_T_LJ; G.el14 = terms & B200MD_T_ELECTROSTATICS;
  G.on = ((terms & B200MD_T_BONDS) ? 1u : 0u) | ((terms & B200MD_T_ANGLES) ? 2u : 0u) | ((terms & B200MD_T_DIHEDRALS) ? 4u : 0u) |
         (((terms & B200MD_T_14) && (G.lj14 || G.el14)) ? 8u : 0u) | ((terms & B200MD_T_IMPROPERS) ? 16u : 0u);
  return G;
}
__device__ __forceinline__ int litem_kind(const LItemRanges& G, int g) { return (g >= G.e1) + (g >= G.e2) + (g >= G.e3) + (g >= G.e4); }
// The arithmetic of one item of kind `kind` on atoms q with parameters prm (bond / angle: (k, x0); 1-4: (A, B, scnb, scee)
// and qq14 = q_i q_j; torsions: terms [t0, t1) of the kind's term list).
template <typename T, bool FORCE, typename SINK, typename POS>
__device__ __forceinline__ void litem_eval(const LargeTopo& T_, const ForceParams& P, const LItemRanges& G, const int kind,
                                           const int4 q, const float4 prm, const T qq14, const int t0, const int t1,
                                           const POS& p, const T* bx, bool vac, SINK& f, T* e /* [8] */) {
  typedef v3t<T> V;
  if (kind == 0) {  // forces.py:111-132, 492-501
    const V d = vwrap(vsub(p.template get<T>(q.x), p.template get<T>(q.y)), bx, vac);
    const T r2b = vdot(d, d), idist = trsqrt(r2b), dist = r2b * idist;
    if (P.has_cutoff && !(dist <= (T)P.rc)) {  // forces.py:117-125: dropped (its slots must still read as zero)
      if (FORCE) { const V zero = {(T)0, (T)0, (T)0}; lput<T>(f, 0, zero, (T)0); lput<T>(f, 1, zero, (T)0); }
      return;
    }
    const T x = dist - (T)prm.y;
    e[B200MD_E_BONDS] += (T)prm.x * x * x;
    if (FORCE) {
      const T c = (T)2 * (T)prm.x * x * idist;
      lput<T>(f, 0, d, -c);
      lput<T>(f, 1, d, c);
    }
  } else if (kind == 1) {  // forces.py:134-145, 504-537
    const V r21 = vwrap(vsub(p.template get<T>(q.x), p.template get<T>(q.y)), bx, vac);
    const V r23 = vwrap(vsub(p.template get<T>(q.z), p.template get<T>(q.y)), bx, vac);
    const T n23i = trsqrt(vdot(r23, r23)), n21i = trsqrt(vdot(r21, r21));
    T ct = vdot(r23, r21) * n21i * n23i;
    ct = ct < (T)-1 ? (T)-1 : (ct > (T)1 ? (T)1 : ct);
    const T dth = acos(ct) - (T)prm.y;
    e[B200MD_E_ANGLES] += (T)prm.x * dth * dth;
    if (FORCE) {
      const T st2 = (T)1 - ct * ct;
      const T coef = st2 > (T)0 ? (T)-2 * (T)prm.x * dth * trsqrt(st2) : (T)0;  // 0 where sin(theta) == 0 (forces.py:521-523)
      const V f0 = vscale(vsub(vscale(r21, ct * n21i), vscale(r23, n23i)), coef * n21i);
      const V f2 = vscale(vsub(vscale(r23, ct * n23i), vscale(r21, n21i)), coef * n23i);
      lput<T>(f, 0, f0, (T)1);
      lput<T>(f, 2, f2, (T)1);
      lput<T>(f, 1, {f0.x + f2.x, f0.y + f2.y, f0.z + f2.z}, (T)-1);
    }
  } else if (kind == 2) {  // forces.py:147-174
    e[B200MD_E_DIHEDRALS] += ltorsion<T, FORCE, SINK, POS>(p, bx, vac, q, T_.dih_term, T_.dih_cs, t0, t1, f);
  } else if (kind == 3) {  // forces.py:176-226: never cut, switched or reaction-field corrected
    const V d = vwrap(vsub(p.template get<T>(q.x), p.template get<T>(q.y)), bx, vac);
    const T r2 = vdot(d, d), rinv1 = trsqrt(r2), rinv2 = rinv1 * rinv1, rinv6 = rinv2 * rinv2 * rinv2;
    T c = 0;  // -(dE/dr) / r
    if (G.lj14) {
      const T iscnb = trcp((T)prm.z);
      e[B200MD_E_LJ] += ((T)prm.x * rinv6 * rinv6 - (T)prm.y * rinv6) * iscnb;
      c += ((T)12 * (T)prm.x * rinv6 * rinv6 - (T)6 * (T)prm.y * rinv6) * rinv2 * iscnb;
    }
    if (G.el14) {
      const T pot = (T)kLElec * qq14 * rinv1 * trcp((T)prm.w);
      e[B200MD_E_ELECTROSTATICS] += pot;
      c += pot * rinv2;
    }
    if (FORCE) {
      lput<T>(f, 0, d, c);
      lput<T>(f, 1, d, -c);
    }
  } else {  // forces.py:228-255
    e[B200MD_E_IMPROPERS] += ltorsion<T, FORCE, SINK, POS>(p, bx, vac, q, T_.imp_term, T_.imp_cs, t0, t1, f);
  }
}

// Item g of the concatenated list [bonds | angles | dihedrals | 1-4 pairs | impropers]: its indices and parameters from the
// per-kind arrays (the FP64 energy kernel; the force kernel reads the same data inline from its window lists).
template <typename T, bool FORCE, typename SINK>
__device__ __forceinline__ void litem(const LargeTopo& T_, const ForceParams& P, const LItemRanges& G, int g, const float* pg,
                                      const T* bx, bool vac, SINK& f, T* e /* [8] */) {
  const GlobalPos p = {pg};
  int4 q = make_int4(0, 0, 0, 0);
  float4 prm = make_float4(0.f, 0.f, 0.f, 0.f);
  T qq14 = 0;
  int t0 = 0, t1 = 0;
  const int kind = litem_kind(G, g);
  if (kind == 0) {
    const int2 ij = __ldg(T_.bond_idx + g);
    const float2 kp = __ldg(T_.bond_prm + g);
    q.x = ij.x; q.y = ij.y; prm.x = kp.x; prm.y = kp.y;
  } else if (kind == 1) {
    const int a = g - G.e1;
    q.x = __ldg(T_.angle_idx + 3 * a); q.y = __ldg(T_.angle_idx + 3 * a + 1); q.z = __ldg(T_.angle_idx + 3 * a + 2);
    const float2 kp = __ldg(T_.angle_prm + a);
    prm.x = kp.x; prm.y = kp.y;
  } else if (kind == 2) {
    const int d = g - G.e2;
    q = __ldg(T_.dih_idx + d); t0 = __ldg(T_.dih_start + d); t1 = __ldg(T_.dih_start + d + 1);
  } else if (kind == 3) {
    const int k = g - G.e3;
    const int2 ij = __ldg(T_.idx14 + k);
    q.x = ij.x; q.y = ij.y; prm = __ldg(T_.prm14 + k);
    qq14 = (T)__ldg(T_.charge + ij.x) * (T)__ldg(T_.charge + ij.y);
  } else {
    const int d = g - G.e4;
    q = __ldg(T_.imp_idx + d); t0 = __ldg(T_.imp_start + d); t1 = __ldg(T_.imp_start + d + 1);
  }
  litem_eval<T, FORCE, SINK, GlobalPos>(T_, P, G, kind, q, prm, qq14, t0, t1, p, bx, vac, f, e);
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

// Bonded ENERGIES (report steps / Forces.compute with energies), FP64, item-centric: one thread per item, no forces.
// grid (nblk, min(R, 65535)); partial[(r * npart + blockIdx.x) * 8 + slot]
__global__ void __launch_bounds__(kLBondedThreads)
large_bonded_energy_kernel(const LargeTopo T, const ForceParams P, const long long R, const float* __restrict__ pos,
                           const float* __restrict__ box, double* __restrict__ partial, const int npart,
                           const float* __restrict__ smd_centers, const double smd_center) {
  __shared__ double sred[(kLBondedThreads / 32) * B200MD_NUM_ENERGIES];
  const int tid = threadIdx.x;
  const int N = T.natoms;
  const LItemRanges G = litem_ranges(T, P.terms);
  NoSink none;
  for (long long r = blockIdx.y; r < R; r += gridDim.y) {
    const float* p = pos + r * N * 3;
    double bxd[3] = {0.0, 0.0, 0.0};
    if (box) { bxd[0] = __ldg(box + r * 9); bxd[1] = __ldg(box + r * 9 + 4); bxd[2] = __ldg(box + r * 9 + 8); }
    const bool vac = bxd[0] == 0.0 && bxd[1] == 0.0 && bxd[2] == 0.0;
    double e[B200MD_NUM_ENERGIES];
#pragma unroll
    for (int k = 0; k < B200MD_NUM_ENERGIES; ++k) e[k] = 0.0;
    for (int g = blockIdx.x * kLBondedThreads + tid; g < G.e5; g += gridDim.x * kLBondedThreads)
      if (((G.on >> litem_kind(G, g)) & 1u)) litem<double, false, NoSink>(T, P, G, g, p, bxd, vac, none, e);
    if (P.smd_on && blockIdx.x == 0 && tid == 0) {
      const double center = smd_centers ? (double)__ldg(smd_centers + r) : smd_center;
      const v3t<double> d = vsub(vpos<double>(p, P.smd_a2), vpos<double>(p, P.smd_a1));
      const double dev = center - sqrt(vdot(d, d));
      e[B200MD_E_EXTERNAL] += 0.5 * (double)P.smd_fk * dev * dev;
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

// ---- pair forces: HALF shell over rectangles of atom blocks ------------------------------------------------------------
// The 32-atom tiles are grouped into nblk blocks of bt <= kLBT tiles.  One CTA takes one rectangle (BI >= BJ) of one
// replica: both blocks are staged in shared memory (the i block as float4 rows, the j block structure-of-arrays); warp w
// holds, for every j tile of BJ in turn, the tile's w-th chunk of 8 atoms in broadcast operands and meets it with the
// 32-atom i tiles of BI ("passes", lane = i atom) exactly like the on-chip tile task of b200md_forces.cuh: packed
// pair_eval2, the i-side force of a pass added to the warp's private shared-memory row, the j-side force kept in 24
// registers over the passes and transposed once per chunk by a shuffle butterfly.  In a diagonal rectangle a chunk only
// meets i tiles ABOVE its own tile, and the 32 x 32 pairs inside a tile are evaluated full-shell by its lanes (no j side).
// Every unordered pair is evaluated once -- round 1 evaluated each from both atoms.
// Cutoff mode: lane l < bt tests tile box (BI, l) against the j tile's box (large_bbox_kernel; under periodic boundaries
// by the minimum image of the box centres, lgap), a ballot is the chunk's pass list; a listed pass still drops out on a warp vote when none of its 32 x 8 pairs is within the cutoff
// (the r^2 of the vote are the expressions pair_eval2 evaluates: bit-identical with and without the skips).
// The CTA writes the block's partial force rows fpart[r][slot][atom][3]: atoms of BI -> slot BJ, atoms of BJ -> slot BI
// (diagonal: both sides summed into slot BI), so atom a of block B finds slot k written by rectangle (max(B, k), min(B, k)):
// every slot of every atom is written exactly once per launch, no atomics, no zeroing, a fixed summation order.
constexpr int kLBT = 9;          // tiles per block at most
constexpr int kLB = 32 * kLBT;   // atoms per block at most
struct RectSmem {
  float4 ipos[kLB];  // x, y, z, charge * el_on * kElecFactor
  float jx[kLB], jy[kLB], jz[kLB], jq[kLB];  // jq: charge * el_on
  int irow[kLB];     // byte offset of the i atom's row in the type-pair table
  int jcol[kLB];     // byte offset of the j atom's column
  float fi[kLThreads / 32][3][kLB];
  float fj[3][kLB];
};

// Lower bound of |x_a - x_b| in one component for atoms a, b inside the intervals [alo, ahi], [blo, bhi].  Vacuum: the gap
// of the intervals, in the monotonic fp32 arithmetic of the pair loop (an exact bound).  Periodic (box length -nb, 1 / L =
// ib; 0, 0 for an open direction): the minimum-image distance of the interval centres less the two half widths -- any pair
// is x_a - x_b = (c_a - c_b) + (d_a - d_b) with |d| <= the half widths, so no image brings it closer; the callers leave a
// 0.1 % margin on the squared distance for the rounding of this expression (3e-6 A at 100 A against 4.5e-3 A of margin).
template <bool PBC>
__device__ __forceinline__ float lgap(const float alo, const float ahi, const float blo, const float bhi, const float nb,
                                      const float ib) {
  if (!PBC) return fmaxf(fmaxf(__fsub_rn(blo, ahi), __fsub_rn(alo, bhi)), 0.0f);
  float c = 0.5f * ((alo + ahi) - (blo + bhi));
  c = fmaf(nb, rintf(c * ib), c);
  return fmaxf(fabsf(c) - 0.5f * ((ahi - alo) + (bhi - blo)), 0.0f);
}

template <int MODE, bool PBC, bool COH>
__device__ __forceinline__ void large_rect(const LargeTopo& T, const ForceParams& P, const float* __restrict__ p,
                                           const float* __restrict__ bb, RectSmem& S, const uint32_t tab, const int BI,
                                           const int BJ, const float3 nbox, const float3 ibox, float* __restrict__ out_i,
                                           float* __restrict__ out_j, const int tid) {
  const int N = T.natoms, bt = T.bt, nw = T.nwords;
  const int warp = tid >> 5, lane = tid & 31;
  const bool diag = BI == BJ;
  const int tstride = large_tab_stride(T.ntypes);
  const int nia = min(32 * bt, max(0, 32 * nw - 32 * bt * BI)), nja = min(32 * bt, max(0, 32 * nw - 32 * bt * BJ));  // padded atoms of the blocks
  const bool boxtest = MODE == 1 && bb != nullptr;
  const float rc2_box = PBC ? 1.001f * P.rc2 : P.rc2;
  if (boxtest) {  // the whole rectangle beyond the cutoff (chains far apart): its rows are zero
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    const int blk = lane < 16 ? BI : BJ, tl = lane & 15, tile = blk * bt + tl;
    if (tl < bt && tile < nw) {
      const float4 a = ldm4<COH>((const float4*)(bb + 8 * tile)), b = ldm4<COH>((const float4*)(bb + 8 * tile + 4));
      lo[0] = a.x; lo[1] = a.y; lo[2] = a.z; hi[0] = b.x; hi[1] = b.y; hi[2] = b.z;
    }
#pragma unroll
    for (int o = 8; o > 0; o >>= 1)
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        lo[c] = fminf(lo[c], __shfl_xor_sync(0xffffffffu, lo[c], o));
        hi[c] = fmaxf(hi[c], __shfl_xor_sync(0xffffffffu, hi[c], o));
      }
    float g2 = 0.0f;
#pragma unroll
    for (int c = 0; c < 3; ++c) {  // lanes 0..15 hold BI's box, 16..31 BJ's
      const float olo = __shfl_xor_sync(0xffffffffu, lo[c], 16), ohi = __shfl_xor_sync(0xffffffffu, hi[c], 16);
      const float g = lgap<PBC>(lo[c], hi[c], olo, ohi, c == 0 ? nbox.x : (c == 1 ? nbox.y : nbox.z), c == 0 ? ibox.x : (c == 1 ? ibox.y : ibox.z));
      g2 = __fmaf_rn(g, g, g2);
    }
    if (!(g2 <= rc2_box)) {  // (uniform over the CTA: every warp computes the same boxes)
      for (int k = tid; k < 3 * nia; k += kLThreads) if (32 * bt * BI + k / 3 < N) out_i[k] = 0.0f;
      if (!diag) for (int k = tid; k < 3 * nja; k += kLThreads) if (32 * bt * BJ + k / 3 < N) out_j[k] = 0.0f;
      return;
    }
  }
  // stage the two blocks, zero the accumulators
  for (int t = tid; t < 32 * bt; t += kLThreads) {
    const int ai = 32 * bt * BI + t, aj = 32 * bt * BJ + t;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    int row = 0;
    if (ai < N) {
      v = make_float4(ldm<COH>(p + 3 * ai), ldm<COH>(p + 3 * ai + 1), ldm<COH>(p + 3 * ai + 2), __fmul_rn(__ldg(T.charge + ai) * P.el_on, kElecFactor));
      row = __ldg(T.atom_type + ai) * tstride * 8;
    }
    S.ipos[t] = v; S.irow[t] = row;
    float x = 0.f, y = 0.f, z = 0.f, q = 0.f;
    int col = 0;
    if (aj < N) { x = ldm<COH>(p + 3 * aj); y = ldm<COH>(p + 3 * aj + 1); z = ldm<COH>(p + 3 * aj + 2); q = __ldg(T.charge + aj) * P.el_on; col = __ldg(T.atom_type + aj) * 8; }
    S.jx[t] = x; S.jy[t] = y; S.jz[t] = z; S.jq[t] = q; S.jcol[t] = col;
  }
  for (int k = tid; k < (kLThreads / 32) * 3 * kLB; k += kLThreads) (&S.fi[0][0][0])[k] = 0.0f;
  for (int k = tid; k < 3 * kLB; k += kLThreads) (&S.fj[0][0])[k] = 0.0f;
  __syncthreads();
  const PairConst C = make_pair_const<MODE>(P);
  // ---- chunks of 8 j atoms against the i tiles ----
  for (int jtl = 0; jtl < bt; ++jtl) {
    const int jt = BJ * bt + jtl;
    if (jt >= nw) break;
    const int j0 = 32 * jtl + 8 * warp;  // this warp's chunk of the tile
    // pass list: i tiles of BI (above the j tile in a diagonal rectangle) that exist and, in cutoff mode, can reach the j tile
    bool need = lane < bt && BI * bt + lane < nw && (!diag || lane > jtl);
    if (boxtest && need) {
      const int it = BI * bt + lane;
      const float4 ilo = ldm4<COH>((const float4*)(bb + 8 * it)), ihi = ldm4<COH>((const float4*)(bb + 8 * it + 4));
      const float4 tlo = ldm4<COH>((const float4*)(bb + 8 * jt)), thi = ldm4<COH>((const float4*)(bb + 8 * jt + 4));
      const float gx = lgap<PBC>(ilo.x, ihi.x, tlo.x, thi.x, nbox.x, ibox.x), gy = lgap<PBC>(ilo.y, ihi.y, tlo.y, thi.y, nbox.y, ibox.y);
      const float gz = lgap<PBC>(ilo.z, ihi.z, tlo.z, thi.z, nbox.z, ibox.z);
      need = __fmaf_rn(gz, gz, __fmaf_rn(gy, gy, __fmul_rn(gx, gx))) <= rc2_box;
    }
    uint32_t todo = __ballot_sync(0xffffffffu, need);
    if (todo == 0u) continue;  // fj of the chunk stays zero
    int jc[8];
    {
      const int4 c0 = *(const int4*)(S.jcol + j0), c1 = *(const int4*)(S.jcol + j0 + 4);
      jc[0] = c0.x; jc[1] = c0.y; jc[2] = c0.z; jc[3] = c0.w; jc[4] = c1.x; jc[5] = c1.y; jc[6] = c1.z; jc[7] = c1.w;
    }
    u64 fjx[4], fjy[4], fjz[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) fjx[q] = fjy[q] = fjz[q] = 0ull;
    bool any = MODE == 0;  // cutoff mode: did any listed pass survive its vote?
    // the i atom, its table row and its mask word of the NEXT pass are fetched while the current one computes
    int ibl_n = __ffs(todo) - 1;
    float4 pi_n = S.ipos[32 * ibl_n + lane];
    int roff_n = S.irow[32 * ibl_n + lane];
    uint32_t mw_n = __ldg(T.mask + (size_t)(32 * (BI * bt + ibl_n) + lane) * nw + jt);
#pragma unroll 1
    while (todo) {
      asm volatile("" ::: "memory");  // re-read the 8 broadcast j atoms every pass instead of pinning 32 registers
      const int ibl = ibl_n;
      todo &= todo - 1;
      const float4 pi = pi_n;
      const uint32_t taddr = tab + (uint32_t)roff_n;
      const uint32_t m8 = (mw_n >> (8 * warp)) & 0xffu;
      if (todo) {
        ibl_n = __ffs(todo) - 1;
        pi_n = S.ipos[32 * ibl_n + lane];
        roff_n = S.irow[32 * ibl_n + lane];
        mw_n = __ldg(T.mask + (size_t)(32 * (BI * bt + ibl_n) + lane) * nw + jt);
      }
      const float kqi = pi.w;
      const u64 xi = dup2(pi.x), yi = dup2(pi.y), zi = dup2(pi.z), kq2 = dup2(kqi);
      if (MODE == 1) {
        bool in = false;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const ulonglong2 xj = *(const ulonglong2*)(S.jx + j0 + 4 * h), yj = *(const ulonglong2*)(S.jy + j0 + 4 * h), zj = *(const ulonglong2*)(S.jz + j0 + 4 * h);
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
        any = true;
      }
      u64 ax = 0ull, ay = 0ull, az = 0ull;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const ulonglong2 xj = *(const ulonglong2*)(S.jx + j0 + 4 * h), yj = *(const ulonglong2*)(S.jy + j0 + 4 * h), zj = *(const ulonglong2*)(S.jz + j0 + 4 * h);
        const ulonglong2 qj = *(const ulonglong2*)(S.jq + j0 + 4 * h);
#pragma unroll
        for (int k = 0; k < 2; ++k) {
          const int q = 2 * h + k;
          const u64 a12 = pk(lds_f32<0>(taddr + jc[2 * q]), lds_f32<0>(taddr + jc[2 * q + 1]));
          const u64 nb6 = pk(lds_f32<4>(taddr + jc[2 * q]), lds_f32<4>(taddr + jc[2 * q + 1]));
          u64 dx = sub2(xi, k ? xj.y : xj.x), dy = sub2(yi, k ? yj.y : yj.x), dz = sub2(zi, k ? zj.y : zj.x);
          if (PBC) { dx = wrap2(dx, nbox.x, ibox.x); dy = wrap2(dy, nbox.y, ibox.y); dz = wrap2(dz, nbox.z, ibox.z); }
          const u64 sc = pair_eval2<MODE>(dx, dy, dz, kqi, kq2, k ? qj.y : qj.x, a12, nb6, (m8 >> (2 * q)) & 1u, (m8 >> (2 * q + 1)) & 1u, P, C);
          ax = fma2(sc, dx, ax); ay = fma2(sc, dy, ay); az = fma2(sc, dz, az);
          fjx[q] = fma2(sc, dx, fjx[q]); fjy[q] = fma2(sc, dy, fjy[q]); fjz[q] = fma2(sc, dz, fjz[q]);
        }
      }
      float* f = &S.fi[warp][0][32 * ibl + lane];
      f[0] = __fadd_rn(f[0], hsum2(ax)); f[kLB] = __fadd_rn(f[kLB], hsum2(ay)); f[2 * kLB] = __fadd_rn(f[2 * kLB], hsum2(az));
    }
    if (!any) continue;  // every pass dropped out: fj of the chunk stays zero
    // transposing butterfly (as tile_task): 24 per-lane partials -> per-atom sums; after the xor-16/8/4 stages lane l holds
    // atom (l >> 2) & 7 of the chunk, xor-2/1 finish the sum.  Fixed order => deterministic.
    float fj[8][3];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      upk(fjx[q], fj[2 * q][0], fj[2 * q + 1][0]);
      upk(fjy[q], fj[2 * q][1], fj[2 * q + 1][1]);
      upk(fjz[q], fj[2 * q][2], fj[2 * q + 1][2]);
    }
    const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4;
    float v4[4][3], v2[2][3], v1[3];
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float send = b4 ? fj[k][c] : fj[k + 4][c], keep = b4 ? fj[k + 4][c] : fj[k][c];
        v4[k][c] = __fadd_rn(keep, __shfl_xor_sync(0xffffffffu, send, 16));
      }
#pragma unroll
    for (int k = 0; k < 2; ++k)
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float send = b3 ? v4[k][c] : v4[k + 2][c], keep = b3 ? v4[k + 2][c] : v4[k][c];
        v2[k][c] = __fadd_rn(keep, __shfl_xor_sync(0xffffffffu, send, 8));
      }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float send = b2 ? v2[0][c] : v2[1][c], keep = b2 ? v2[1][c] : v2[0][c];
      float v = __fadd_rn(keep, __shfl_xor_sync(0xffffffffu, send, 4));
      v = __fadd_rn(v, __shfl_xor_sync(0xffffffffu, v, 2));
      v1[c] = __fadd_rn(v, __shfl_xor_sync(0xffffffffu, v, 1));
    }
    if ((lane & 3) == 0) {
      const int j = j0 + (lane >> 2);
      S.fj[0][j] = -v1[0]; S.fj[1][j] = -v1[1]; S.fj[2][j] = -v1[2];
    }
  }
  // ---- pairs inside a tile (diagonal rectangles): lane = i atom meets the 32 atoms of its own tile, no j side ----
  if (diag) {
    for (int tl = warp; tl < bt; tl += kLThreads / 32) {
      const int tile = BI * bt + tl;
      if (tile >= nw) break;
      const float4 pi = S.ipos[32 * tl + lane];
      const uint32_t taddr = tab + (uint32_t)S.irow[32 * tl + lane];
      const uint32_t m = __ldg(T.mask + (size_t)(32 * tile + lane) * nw + tile);
      const float kqi = pi.w;
      const u64 xi = dup2(pi.x), yi = dup2(pi.y), zi = dup2(pi.z), kq2 = dup2(kqi);
      u64 ax = 0ull, ay = 0ull, az = 0ull;
#pragma unroll 4
      for (int jj = 0; jj < 32; jj += 2) {
        const float* sj = S.jx + 32 * tl + jj;
        const u64 xj = *(const u64*)sj, yj = *(const u64*)(sj + kLB), zj = *(const u64*)(sj + 2 * kLB), qj = *(const u64*)(sj + 3 * kLB);
        const int2 tj = *(const int2*)(S.jcol + 32 * tl + jj);
        const u64 a12 = pk(lds_f32<0>(taddr + tj.x), lds_f32<0>(taddr + tj.y)), nb6 = pk(lds_f32<4>(taddr + tj.x), lds_f32<4>(taddr + tj.y));
        u64 dx = sub2(xi, xj), dy = sub2(yi, yj), dz = sub2(zi, zj);
        if (PBC) { dx = wrap2(dx, nbox.x, ibox.x); dy = wrap2(dy, nbox.y, ibox.y); dz = wrap2(dz, nbox.z, ibox.z); }
        const u64 sc = pair_eval2<MODE>(dx, dy, dz, kqi, kq2, qj, a12, nb6, (m >> jj) & 1u, (m >> (jj + 1)) & 1u, P, C);
        ax = fma2(sc, dx, ax); ay = fma2(sc, dy, ay); az = fma2(sc, dz, az);
      }
      float* f = &S.fi[warp][0][32 * tl + lane];
      f[0] = __fadd_rn(f[0], hsum2(ax)); f[kLB] = __fadd_rn(f[kLB], hsum2(ay)); f[2 * kLB] = __fadd_rn(f[2 * kLB], hsum2(az));
    }
  }
  __syncthreads();
  for (int k = tid; k < 3 * nia; k += kLThreads) {
    const int t = k / 3, c = k - 3 * t;
    if (32 * bt * BI + t < N) {
      float v = __fadd_rn(__fadd_rn(S.fi[0][c][t], S.fi[1][c][t]), __fadd_rn(S.fi[2][c][t], S.fi[3][c][t]));
      if (diag) v = __fadd_rn(v, S.fj[c][t]);
      out_i[k] = v;
    }
  }
  if (!diag)
    for (int k = tid; k < 3 * nja; k += kLThreads) {
      const int t = k / 3, c = k - 3 * t;
      if (32 * bt * BJ + t < N) out_j[k] = S.fj[c][t];
    }
}

// Every force of one Forces.compute in ONE launch, as partial rows: grid (nrect + nwin, min(R, 65535)).
//   CTA x < nrect : pair rectangle x = BI (BI + 1) / 2 + BJ of one replica (large_rect above) -> slots 0 .. nblk - 1;
//   CTA x >= nrect: bonded window w = x - nrect of kLWin = 64 atoms (fp32): every item that touches the window is evaluated
//                   into private shared-memory slots (positions of the atoms those items reference are staged in shared
//                   memory first when they span at most kLHalo atoms -- true of any chain-ordered topology), then thread
//                   a < 64 sums atom a's slots in their fixed order, adds the SMD restraint and writes slot nblk.
// The two kinds of CTA share the SMs (the bonded ones wait on gathered loads, the pair ones issue packed FP32), and the
// force row of an atom is the fixed-order sum of its nblk + 1 slots, taken by the consumer (large_sum_kernel, or the
// integrator's half-kick kernel, which reads the rows anyway).
// Shared memory (dynamic): type-pair tables + RectSmem  or  bonded slots [win_slots_max][3] + staged positions [halo_max][3].
template <bool COH>
__device__ __forceinline__ float large_row_sum(const float* fpart, const long long r, const int nslot, const int slot0,
                                               const int N, const int a, const int c) {
  const float* q = fpart + ((size_t)r * nslot * N + a) * 3 + c;
  float f = COH ? __ldcg(q + (size_t)(nslot - 1) * N * 3) : q[(size_t)(nslot - 1) * N * 3];  // bonded
  for (int k = slot0; k < nslot - 1; ++k) f = __fadd_rn(f, COH ? __ldcg(q + (size_t)k * N * 3) : q[(size_t)k * N * 3]);
  return f;
}
__host__ __device__ inline int large_tab_bytes(const LargeTopo& T) { return (2 * T.ntypes * large_tab_stride(T.ntypes) * (int)sizeof(float) + 15) & ~15; }
inline int large_pairs_dyn_bytes(const LargeTopo& T) { return large_tab_bytes(T) + (int)sizeof(RectSmem); }
inline int large_bonded_dyn_bytes(const LargeTopo& T) { return (int)sizeof(float4) * (T.win_slots_max + T.halo_max); }
inline int large_steps_dyn_bytes(const LargeTopo& T) { return large_tab_bytes(T) + std::max((int)sizeof(RectSmem), large_bonded_dyn_bytes(T)); }

// The type-pair tables of a CTA, staged once per launch: [ntypes][stride] entries (12 A tab_a, 6 B tab_b)
__device__ __forceinline__ void large_stage_tables(const LargeTopo& T, const ForceParams& P, float* sAB, const int tid) {
  const int tstride = large_tab_stride(T.ntypes);
  for (int k = tid; k < T.ntypes * T.ntypes; k += kLThreads) {
    const int t = k / T.ntypes, u = k - t * T.ntypes;
    sAB[2 * (t * tstride + u)] = __ldg(T.tab12a + k) * P.tab_a;
    sAB[2 * (t * tstride + u) + 1] = __ldg(T.tab6b + k) * P.tab_b;
  }
}

// One pair rectangle (index x = BI (BI + 1) / 2 + BJ) of replica r.  lsmem: tables | RectSmem.
template <int MODE, bool COH>
__device__ __forceinline__ void large_pairs_unit(const LargeTopo& T, const ForceParams& P, const float* pos, const float* box,
                                                 float* fpart, const float* bbox, const long long r, const int x, char* lsmem,
                                                 const int tid) {
  const int N = T.natoms, nslot = T.nblk + 1;
  RectSmem& S = *(RectSmem*)(lsmem + large_tab_bytes(T));
  const uint32_t tab = (uint32_t)__cvta_generic_to_shared(lsmem);
  int BI = 0;
  while ((BI + 1) * (BI + 2) / 2 <= x) ++BI;
  const int BJ = x - BI * (BI + 1) / 2;
  const float* p = pos + r * N * 3;
  const float b0 = box ? __ldg(box + r * 9) : 0.0f, b1 = box ? __ldg(box + r * 9 + 4) : 0.0f, b2 = box ? __ldg(box + r * 9 + 8) : 0.0f;
  const bool pbc = b0 != 0.0f || b1 != 0.0f || b2 != 0.0f;  // forces.py:359
  const float3 nbox = make_float3(-b0, -b1, -b2);
  const float3 ibox = make_float3(b0 != 0.0f ? 1.0f / b0 : 0.0f, b1 != 0.0f ? 1.0f / b1 : 0.0f, b2 != 0.0f ? 1.0f / b2 : 0.0f);
  float* out_i = fpart + (((size_t)r * nslot + BJ) * N + 32 * T.bt * BI) * 3;
  float* out_j = fpart + (((size_t)r * nslot + BI) * N + 32 * T.bt * BJ) * 3;
  __syncthreads();  // the tables (first unit); the previous unit's rows have been read
  const float* bb = bbox ? bbox + r * T.nwords * 8 : nullptr;
  if (pbc) large_rect<MODE, true, COH>(T, P, p, bb, S, tab, BI, BJ, nbox, ibox, out_i, out_j, tid);
  else large_rect<MODE, false, COH>(T, P, p, bb, S, tab, BI, BJ, nbox, ibox, out_i, out_j, tid);
}

// grid (nrect, min(R, 65535))
template <int MODE>
__global__ void __launch_bounds__(kLThreads, 5)  // (4 and 6 CTAs per SM measured within 2 % of 5)
large_pairs_kernel(const LargeTopo T, const ForceParams P, const long long R, const float* __restrict__ pos,
                   const float* __restrict__ box, float* __restrict__ fpart, const float* __restrict__ bbox) {
  extern __shared__ __align__(16) char lsmem[];
  large_stage_tables(T, P, (float*)lsmem, threadIdx.x);
  for (long long r = blockIdx.y; r < R; r += gridDim.y)
    large_pairs_unit<MODE, false>(T, P, pos, box, fpart, bbox, r, (int)blockIdx.x, lsmem, threadIdx.x);
}

// The bonded window w of replica r (+ the SMD restraint, the NaN flag of its atoms) -> the last slot of the partial rows.
// wsmem: slots [win_slots_max] float4 | staged positions [halo_max] float4.
template <bool COH>
__device__ __forceinline__ void large_bonded_unit(const LargeTopo& T, const ForceParams& P, const float* pos, const float* box,
                                                  float* fpart, uint32_t* status, const float* smd_centers, const double smd_center,
                                                  double* smd_rec, const long long r, const int w, char* wsmem, const int tid) {
  const int N = T.natoms, nslot = T.nblk + 1;
  float4* lslots = (float4*)wsmem;
  float4* shalo = lslots + T.win_slots_max;
  const LItemRanges G = litem_ranges(T, P.terms);
  const int k0 = __ldg(T.win_start + w), k1 = __ldg(T.win_start + w + 1);
  const int4 ke = __ldg(T.win_kend + w);  // first entry of the angles, dihedrals, 1-4 pairs, impropers of the window
  const int hlo = __ldg(T.win_lo + w), hn = __ldg(T.win_n + w);
  const int a = kLWin * w + (tid >> 1), half = tid & 1;  // two threads gather an atom: its even / its odd slots
  const bool mine = a < N;
  const int g0 = __ldg(T.gstart + kLWin * w);
  const int s0 = mine ? __ldg(T.gstart + a) - g0 : 0, s1 = mine ? __ldg(T.gstart + a + 1) - g0 : 0;
  const float* p = pos + r * N * 3;
  const float b0 = box ? __ldg(box + r * 9) : 0.0f, b1 = box ? __ldg(box + r * 9 + 4) : 0.0f, b2 = box ? __ldg(box + r * 9 + 8) : 0.0f;
  const bool pbc = b0 != 0.0f || b1 != 0.0f || b2 != 0.0f;
  __syncthreads();  // the slots and staged positions of the previous unit are no longer read
  for (int k = tid; k < hn; k += kLThreads) shalo[k] = make_float4(ldm<COH>(p + 3 * (hlo + k)), ldm<COH>(p + 3 * (hlo + k) + 1), ldm<COH>(p + 3 * (hlo + k) + 2), 0.0f);
  __syncthreads();
  const float bxf[3] = {b0, b1, b2};
  auto items = [&](const auto& pe) {
    WindowSink sink;
    sink.slots = lslots;
    float ef[B200MD_NUM_ENERGIES] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};  // by-product of the float path, not read
    for (int k = k0 + tid; k < k1; k += kLThreads) {
      // everything the item needs arrives in three coalesced loads of the window's own lists (round 2 first read the item
      // number, then gathered its indices and its parameters from the per-kind arrays: three dependent loads deep)
      const int4 q = __ldg(T.win_atoms + k);
      const float4 prm = __ldg(T.win_prm + k);
      sink.sl = __ldg(T.win_slot + k);
      const int kind = (k >= ke.x) + (k >= ke.y) + (k >= ke.z) + (k >= ke.w);
      if (((G.on >> kind) & 1u)) {
        float4 pk4 = prm;
        float qq = 0.0f;
        int t0 = 0, t1 = 0;
        if (kind == 3) { qq = prm.z; pk4.z = 1.0f; pk4.w = 1.0f; }  // scnb, scee are folded into A, B, q_i q_j on the host
        else if (kind == 2 || kind == 4) { t0 = __float_as_int(prm.x); t1 = __float_as_int(prm.y); }
        litem_eval<float, true, WindowSink>(T, P, G, kind, q, pk4, qq, t0, t1, pe, bxf, !pbc, sink, ef);
      } else {  // a kind the call has switched off: its slots read as zero
        const v3t<float> zero = {0.0f, 0.0f, 0.0f};
#pragma unroll
        for (int c = 0; c < 4; ++c) lput<float>(sink, c, zero, 0.0f);
      }
    }
  };
  if (hn > 0) items(SharedPos{(uint32_t)__cvta_generic_to_shared(shalo) - 16u * (uint32_t)hlo});
  else items(GlobalPos{p});  // the window's items span more than kLHalo atoms: from HBM
  __syncthreads();
  float bx = 0.0f, by = 0.0f, bz = 0.0f;
  for (int sidx = s0 + half; sidx < s1; sidx += 2) { const float4 v = lslots[sidx]; bx += v.x; by += v.y; bz += v.z; }
  bx = __fadd_rn(bx, __shfl_xor_sync(0xffffffffu, bx, 1));  // (even slots in order) + (odd slots in order): a fixed order
  by = __fadd_rn(by, __shfl_xor_sync(0xffffffffu, by, 1));
  bz = __fadd_rn(bz, __shfl_xor_sync(0xffffffffu, bz, 1));
  if (mine && half == 0) {
    const float x = ldm<COH>(p + 3 * a), y = ldm<COH>(p + 3 * a + 1), z = ldm<COH>(p + 3 * a + 2);
    if (status != nullptr && (x != x || y != y || z != z)) atomicOr(status + r, B200MD_FLAG_NAN_POS);
    if (P.smd_on && (a == P.smd_a1 || a == P.smd_a2)) {  // potcalc.py:75-89, not wrapped; FP64: (c - d) cancels near the centre
      const double center = smd_centers ? (double)__ldg(smd_centers + r) : smd_center;
      const v3t<double> p1 = {(double)ldm<COH>(p + 3 * P.smd_a1), (double)ldm<COH>(p + 3 * P.smd_a1 + 1), (double)ldm<COH>(p + 3 * P.smd_a1 + 2)};
      const v3t<double> p2 = {(double)ldm<COH>(p + 3 * P.smd_a2), (double)ldm<COH>(p + 3 * P.smd_a2 + 1), (double)ldm<COH>(p + 3 * P.smd_a2 + 2)};
      const v3t<double> d = vsub(p2, p1);
      const double dist = sqrt(vdot(d, d)), dev = center - dist;
      const double c = (double)P.smd_fk * dev / dist;  // force on atom2 = +fk (c - d) unit(r2 - r1)
      const float sg = a == P.smd_a2 ? 1.0f : -1.0f;
      bx += sg * (float)(c * d.x); by += sg * (float)(c * d.y); bz += sg * (float)(c * d.z);
      if (smd_rec && a == P.smd_a2) {  // [R][2]: end-to-end distance of this step, work since the first step of the call
        smd_rec[2 * r] = dist;
        smd_rec[2 * r + 1] += (double)P.smd_fk * dev * (double)P.smd_rate;
      }
    }
    float* o = fpart + (((size_t)r * nslot + T.nblk) * N + a) * 3;
    o[0] = bx; o[1] = by; o[2] = bz;
  }
}

// grid (nwin, min(R, 65535))
__global__ void __launch_bounds__(kLBondedThreads, 8)
large_bonded_forces_kernel(const LargeTopo T, const ForceParams P, const long long R, const float* __restrict__ pos,
                           const float* __restrict__ box, float* __restrict__ fpart, uint32_t* __restrict__ status,
                           const float* __restrict__ smd_centers, const double smd_center, double* __restrict__ smd_rec) {
  extern __shared__ __align__(16) char lsmem[];
  for (long long r = blockIdx.y; r < R; r += gridDim.y)
    large_bonded_unit<false>(T, P, pos, box, fpart, status, smd_centers, smd_center, smd_rec, r, (int)blockIdx.x, lsmem, threadIdx.x);
}

// forces[r][a][c] = bonded slot + pair slots slot0 .. nblk - 1 in index order (Forces.compute; the integrator's half-kick
// kernel takes the same sum itself)
__global__ void large_sum_kernel(const LargeTopo T, const long long R, const float* __restrict__ fpart, const int slot0,
                                 float* __restrict__ forces) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= R * T.natoms * 3) return;
  const long long r = g / (T.natoms * 3);
  const int k = (int)(g - r * T.natoms * 3), a = k / 3, c = k - 3 * a;
  forces[g] = large_row_sum<false>(fpart, r, T.nblk + 1, slot0, T.natoms, a, c);
}

// FP64 energies of the nonbonded pairs (report steps / Forces.compute with energies; forces.py:379-413, 451-489 in double
// from the fp32 positions), half shell: CTA = 32-atom i tile, lane = i atom, the 4 warps split the j tiles jt >= ib
// (j > i inside the diagonal tile).  grid (nwords, min(R, 65535)); partial[(r * npart + part0 + ib) * 8 + {el, lj}].
// Only reads the positions: the host runs it beside the force launch on the second stream.
template <int MODE>
__global__ void __launch_bounds__(kLThreads)
large_pair_energy_kernel(const LargeTopo T, const ForceParams P, const long long R, const float* __restrict__ pos,
                         const float* __restrict__ box, double* __restrict__ partial, const int npart, const int part0,
                         const float* __restrict__ bbox) {
  __shared__ double sjd[kLThreads / 32][4 * 32];
  __shared__ int stj[kLThreads / 32][32];
  __shared__ double sred[(kLThreads / 32) * 2];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, ib = blockIdx.x;
  const int N = T.natoms, i = 32 * ib + lane;
  const bool live = i < N;
  const double kq = kLElec * (double)(__ldg(T.charge + i) * P.el_on);
  const double2* trow = T.tab64 + (size_t)__ldg(T.atom_type + i) * T.ntypes;
  const uint32_t* mrow = T.mask + (size_t)i * T.nwords;
  const double rc2 = P.rc_d * P.rc_d, ta = (double)P.tab_a, tb = (double)P.tab_b;
  for (long long r = blockIdx.y; r < R; r += gridDim.y) {
    const float* p = pos + r * N * 3;
    const double bx = box ? (double)__ldg(box + r * 9) : 0.0, by = box ? (double)__ldg(box + r * 9 + 4) : 0.0, bz = box ? (double)__ldg(box + r * 9 + 8) : 0.0;
    const bool pbc = bx != 0.0 || by != 0.0 || bz != 0.0;
    const float* bb = (MODE == 1 && bbox != nullptr) ? bbox + r * T.nwords * 8 : nullptr;
    double xd = 0.0, yd = 0.0, zd = 0.0;
    if (live) { xd = (double)__ldg(p + 3 * i); yd = (double)__ldg(p + 3 * i + 1); zd = (double)__ldg(p + 3 * i + 2); }
    double e[2] = {0.0, 0.0};
    for (int jt = ib + warp; jt < T.nwords; jt += kLThreads / 32) {
      if (bb != nullptr) {  // warp-uniform: the two tile boxes are further apart than the cutoff
        const float4 ilo = __ldg((const float4*)(bb + 8 * ib)), ihi = __ldg((const float4*)(bb + 8 * ib + 4));
        const float4 tlo = __ldg((const float4*)(bb + 8 * jt)), thi = __ldg((const float4*)(bb + 8 * jt + 4));
        float gx, gy, gz;
        if (pbc) {  // minimum image of the box centres (lgap); its own rounding is covered by the 0.1 % margin below
          gx = lgap<true>(ilo.x, ihi.x, tlo.x, thi.x, -(float)bx, bx != 0.0 ? (float)(1.0 / bx) : 0.0f);
          gy = lgap<true>(ilo.y, ihi.y, tlo.y, thi.y, -(float)by, by != 0.0 ? (float)(1.0 / by) : 0.0f);
          gz = lgap<true>(ilo.z, ihi.z, tlo.z, thi.z, -(float)bz, bz != 0.0 ? (float)(1.0 / bz) : 0.0f);
        } else {
          gx = lgap<false>(ilo.x, ihi.x, tlo.x, thi.x, 0.0f, 0.0f); gy = lgap<false>(ilo.y, ihi.y, tlo.y, thi.y, 0.0f, 0.0f);
          gz = lgap<false>(ilo.z, ihi.z, tlo.z, thi.z, 0.0f, 0.0f);
        }
        // (the gap components are fp32 lower bounds of the FP64 |d| below: the margin covers their rounding)
        if (__fmaf_rn(gz, gz, __fmaf_rn(gy, gy, __fmul_rn(gx, gx))) > (pbc ? 1.001f : 1.00001f) * P.rc2) continue;
      }
      const int j = 32 * jt + lane;
      double* sd = sjd[warp];
      __syncwarp();
      float jx = 0.f, jy = 0.f, jz = 0.f, jq = 0.f;
      if (j < N) { jx = __ldg(p + 3 * j); jy = __ldg(p + 3 * j + 1); jz = __ldg(p + 3 * j + 2); jq = __ldg(T.charge + j) * P.el_on; }
      sd[lane] = (double)jx; sd[32 + lane] = (double)jy; sd[64 + lane] = (double)jz; sd[96 + lane] = (double)jq;

      stj[warp][lane] = __ldg(T.atom_type + j);
      __syncwarp();
      const uint32_t m = __ldg(mrow + jt);
      const int jfirst = jt == ib ? lane + 1 : 0;  // diagonal tile: j > i
      for (int jj = 0; jj < 32; ++jj) {
        if (jj < jfirst || ((m >> jj) & 1u)) continue;
        double dx = xd - sd[jj], dy = yd - sd[32 + jj], dz = zd - sd[64 + jj];
        if (pbc) {
          if (bx != 0.0) dx -= bx * rint(dx / bx);
          if (by != 0.0) dy -= by * rint(dy / by);
          if (bz != 0.0) dz -= bz * rint(dz / bz);
        }
        const double r2 = dx * dx + dy * dy + dz * dz;
        if (MODE == 1 && !(r2 <= rc2)) continue;
        const double rinv = drsqrt(r2), r2inv = rinv * rinv, r6 = r2inv * r2inv * r2inv;
        const double2 ab = __ldg(trow + stj[warp][jj]);
        const double qq = kq * sd[96 + jj];
        double lj = (ta * ab.x * r6 + tb * ab.y) * r6;
        if (MODE == 1) {
          const double rr = r2 * rinv;
          if (rr > P.rs_d) {
            const double t = (rr - P.rs_d) * P.inv_sw_d;
            lj *= 1.0 + t * t * t * (-10.0 + t * (15.0 - t * 6.0));
          }
          e[0] += qq * (rinv + P.krf_d * r2 - P.crf_d);
        } else {
          e[0] += qq * rinv;
        }
        e[1] += lj;
      }
    }
    lreduce<2, kLThreads>(e, sred, tid);
    if (tid == 0) {
      double* o = partial + (r * npart + part0 + ib) * B200MD_NUM_ENERGIES;
      o[B200MD_E_ELECTROSTATICS] = e[0]; o[B200MD_E_LJ] = e[1];
#pragma unroll
      for (int k = 2; k < B200MD_NUM_ENERGIES; ++k) o[k] = 0.0;
    }
  }
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
// (pos_old: rigid bonds -- x(t) is kept for SHAKE, which works along the old bond vectors)
__global__ void large_vv1_kernel(const LargeTopo T, const float dt, const long long R, float* __restrict__ pos,
                                 float* __restrict__ vel, const float* __restrict__ forces, float* __restrict__ pos_old) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= R * T.natoms) return;
  const int a = (int)(g % T.natoms);
  const float hdt_m = 0.5f * dt * (1.0f / __ldg(T.mass + a));
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float v = fmaf(hdt_m, forces[3 * g + c], vel[3 * g + c]);
    vel[3 * g + c] = v;
    const float x = pos[3 * g + c];
    if (pos_old != nullptr) pos_old[3 * g + c] = x;
    pos[3 * g + c] = fmaf(dt, v, x);
  }
}

// Atom a of replica r between two force evaluations: the force row (fixed-order sum of the partial rows, stored for the
// caller), the Langevin kick (vel += -gamma vel dt + vcoeff N(0,1)), _second_VV of step `gstep` and -- fuse_next -- the
// _first_VV of the next step (same atom, same thread: vel += 0.5 dt f / m ; pos += dt vel, large_vv1_kernel's arithmetic
// expression by expression).  Returns the kinetic energy after _second_VV; lo = hi = the atom's (moved) position.
template <bool COH>
__device__ __forceinline__ double large_vv2_atom(const LargeTopo& T, const StepParams& I, const unsigned long long gstep,
                                                 const long long r, const int a, float* vel, float* forces, float* pos,
                                                 uint32_t* status, const int fuse_next, const float* fpart, const int slot0,
                                                 float (&lo)[3], float (&hi)[3]) {
  const long long g = r * T.natoms + a;
  const float mass = __ldg(T.mass + a), invm = 1.0f / mass, hdt_m = 0.5f * I.dt * invm;
  float v[3] = {vel[3 * g], vel[3 * g + 1], vel[3 * g + 2]};
  float f[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    f[c] = large_row_sum<COH>(fpart, r, T.nblk + 1, slot0, T.natoms, a, c);
    forces[3 * g + c] = f[c];
  }
  if (I.temperature > 0.0f) {
    const float vcoef = sqrtf(2.0f * I.gamma * invm * kBoltzman * I.temperature * I.dt), vdamp = I.gamma * I.dt;
    float n[3];
    normal3(I.seed, (unsigned long long)(I.replica0 + r), (uint32_t)a, gstep, n);
#pragma unroll
    for (int c = 0; c < 3; ++c) v[c] += fmaf(vcoef, n[c], -vdamp * v[c]);
  }
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    v[c] = fmaf(hdt_m, f[c], v[c]);
    vel[3 * g + c] = v[c];
  }
  const double ek = 0.5 * (double)mass * ((double)v[0] * v[0] + (double)v[1] * v[1] + (double)v[2] * v[2]);
  float x = pos[3 * g], y = pos[3 * g + 1], z = pos[3 * g + 2];
  if (status != nullptr && (x != x || y != y || z != z)) atomicOr(status + r, B200MD_FLAG_NAN_POS);
  if (fuse_next) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      v[c] = fmaf(hdt_m, f[c], v[c]);
      vel[3 * g + c] = v[c];
    }
    x = fmaf(I.dt, v[0], x); y = fmaf(I.dt, v[1], y); z = fmaf(I.dt, v[2], z);
    pos[3 * g] = x; pos[3 * g + 1] = y; pos[3 * g + 2] = z;
  }
  lo[0] = hi[0] = x; lo[1] = hi[1] = y; lo[2] = hi[2] = z;
  return ek;
}
// bounding box of the 32-atom tile held by this warp (lane = atom a, a warp-aligned), as large_bbox_kernel writes it
__device__ __forceinline__ void large_tile_box_store(const LargeTopo& T, float* bbox, const long long r, const int a, const int lane,
                                                     float (&lo)[3], float (&hi)[3]) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      lo[c] = fminf(lo[c], __shfl_xor_sync(0xffffffffu, lo[c], o));
      hi[c] = fmaxf(hi[c], __shfl_xor_sync(0xffffffffu, hi[c], o));
    }
  const int tile = a >> 5;
  if (lane < 3 && tile < T.nwords) {
    bbox[(r * T.nwords + tile) * 8 + lane] = lane == 0 ? lo[0] : (lane == 1 ? lo[1] : lo[2]);
    bbox[(r * T.nwords + tile) * 8 + 4 + lane] = lane == 0 ? hi[0] : (lane == 1 ? hi[1] : hi[2]);
  }
}

// The integrator's launch between two force evaluations.  grid (ceil(N / 256), R): blocks do not straddle replicas.  With
// `ekin` the per-block kinetic-energy partials of a report step; with fuse_next and `bbox` the bounding boxes of the moved
// tiles (a warp is a tile: blocks start at multiples of 256 atoms) -- one launch instead of three.  The host does not fuse
// on snapshot steps and on the last step of a call, where pos must stay x(t + dt).
__global__ void __launch_bounds__(256)
large_vv2_kernel(const LargeTopo T, const StepParams I, const unsigned long long gstep, const long long R,
                 float* __restrict__ vel, float* __restrict__ forces, float* __restrict__ pos,
                 double* __restrict__ ekin, uint32_t* __restrict__ status, const int fuse_next, float* __restrict__ bbox,
                 const float* __restrict__ fpart, const int slot0) {
  __shared__ double sred[8];
  const int tid = threadIdx.x, a = blockIdx.x * 256 + tid;
  for (long long r = blockIdx.y; r < R; r += gridDim.y) {
    double e[1] = {0.0};
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    if (a < T.natoms) e[0] = large_vv2_atom<false>(T, I, gstep, r, a, vel, forces, pos, status, fuse_next, fpart, slot0, lo, hi);
    if (fuse_next && bbox != nullptr) large_tile_box_store(T, bbox, r, a, tid & 31, lo, hi);
    if (ekin != nullptr) {
      lreduce<1, 256>(e, sred, tid);
      if (tid == 0) ekin[r * gridDim.x + blockIdx.x] = e[0];
    }
  }
}

// Steps [s_begin, s_end) of one Integrator.step call in ONE cooperative launch (none of them a report step): per step
//   phase F  the pair rectangles and bonded windows of all replicas, handed out CTA by CTA (the units of
//            large_pairs_kernel / large_bonded_forces_kernel, same code, same partial rows);
//   -- grid barrier --
//   phase V  large_vv2_kernel's atoms (second half-kick of the step, first half-kick + drift of the next, tile boxes);
//   -- grid barrier --
// For calls with few replicas of a large solute the per-step launches of the other path are latency, not work; here the
// state stays in L2 between the phases and a step costs its two grid barriers plus its longest unit (a Langevin step of
// ONE 3 328-atom replica, cutoff mode: 30.5 us as 3 launches + 2 stream joins, 27.9 us here -- the longest rectangle is
// most of it).  Results are bit-identical to the per-step launches (same unit code; positions, boxes and rows another
// CTA wrote are read through L2, ldm<true>).  With more replicas the per-step launches are faster (the two kinds of CTA
// in one kernel share one register allocation: 16 x 3 328 atoms 60 us against 49 us), so the host takes this path only
// while the units of a step fill at most half of the resident grid.
template <int MODE>
__global__ void __launch_bounds__(kLThreads, 4)
large_steps_kernel(const LargeTopo T, const ForceParams P, const StepParams I, const long long R, float* pos, float* vel,
                   float* forces, const float* __restrict__ box, float* fpart, float* bbox, uint32_t* status,
                   double* smd_rec, const int s_begin, const int s_end, const int snapshots_on, const int pairs_on) {
  extern __shared__ __align__(16) char lsmem[];
  cooperative_groups::grid_group grid = cooperative_groups::this_grid();
  const int tid = threadIdx.x;
  large_stage_tables(T, P, (float*)lsmem, tid);
  char* wsmem = lsmem + large_tab_bytes(T);  // the bonded windows use the RectSmem region
  const int nrect = pairs_on ? T.nblk * (T.nblk + 1) / 2 : 0, slot0 = pairs_on ? 0 : T.nblk;
  const long long npair = (long long)nrect * R, nunits = npair + (long long)T.nwin * R;
  const int chunks = (T.natoms + kLThreads - 1) / kLThreads;
  for (int s = s_begin; s < s_end; ++s) {
    const unsigned long long gstep = I.step0 + (unsigned long long)s;
    const double center = (double)P.smd_c0 + (double)P.smd_rate * (double)(gstep + 1);
    for (long long u = blockIdx.x; u < nunits; u += gridDim.x) {
      if (u < npair) large_pairs_unit<MODE, true>(T, P, pos, box, fpart, bbox, u / nrect, (int)(u % nrect), lsmem, tid);
      else large_bonded_unit<true>(T, P, pos, box, fpart, status, P.smd_centers, center, smd_rec, (u - npair) / T.nwin, (int)((u - npair) % T.nwin), wsmem, tid);
    }
    grid.sync();
    const bool snap = snapshots_on && I.snapshot_every > 0 && (s + 1) % I.snapshot_every == 0;
    const int fuse = (s + 1 < I.nsteps && !snap) ? 1 : 0;
    for (long long c = blockIdx.x; c < (long long)chunks * R; c += gridDim.x) {
      const long long r = c / chunks;
      const int a = (int)(c % chunks) * kLThreads + tid;
      float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
      if (a < T.natoms) large_vv2_atom<true>(T, I, gstep, r, a, vel, forces, pos, status, fuse, fpart, slot0, lo, hi);
      if (fuse && bbox != nullptr) large_tile_box_store(T, bbox, r, a, tid & 31, lo, hi);
    }
    grid.sync();
  }
}

// ---- rigid X-H bonds for solutes above 128 atoms --------------------------------------------------------------------------
// The group solvers of b200md_forces.cuh (Newton iterations on the multipliers of a heavy atom's 1 .. 3 X-H bonds, see
// shake_group / rattle_group there) on registers: a group's atoms belong to no other group, so ONE thread loads the
// group from HBM, solves, and writes its atoms back -- no barrier, no shared memory.  K = 1 or 3 as on chip.
struct LGroup { int p, n, h[3]; float d2[3], wp, wh[3]; };
__device__ __forceinline__ LGroup lgroup_load(const LargeTopo& T, const int gi) {
  const int4 g = __ldg(T.cgroup + gi);
  const float4 d2v = __ldg(T.cdist2 + gi), wv = __ldg(T.cweight + gi);
  LGroup G;
  G.p = g.x & 0xffffff; G.n = g.x >> 24;
  G.h[0] = g.y; G.h[1] = g.z; G.h[2] = g.w;
  G.d2[0] = d2v.x; G.d2[1] = d2v.y; G.d2[2] = d2v.z;
  G.wp = wv.x; G.wh[0] = wv.y; G.wh[1] = wv.z; G.wh[2] = wv.w;
  return G;
}
// s_k = x_p(t) - x_hk(t), u_k = x'_p - x'_hk (unconstrained).  Out: displacements dp of the parent, dh_k of hydrogen k that
// restore |u_k + dp - dh_k|^2 = d_k^2 along the old bond vectors.  Returns false beyond kShakeTol.
template <int K>
__device__ __forceinline__ bool lshake_solve(const LGroup& G, const float (&sx)[K], const float (&sy)[K], const float (&sz)[K],
                                             const float (&ux)[K], const float (&uy)[K], const float (&uz)[K], float (&dp)[3],
                                             float (&dhx)[K], float (&dhy)[K], float (&dhz)[K]) {
  const int n = G.n;
  const float wp = G.wp;
  float lam[K];
#pragma unroll
  for (int k = 0; k < K; ++k) lam[k] = 0.0f;
  float dpx = 0.0f, dpy = 0.0f, dpz = 0.0f;
#pragma unroll
  for (int it = 0; it < kShakeNewtonIters; ++it) {
    float A[K][K], b[K], rx[K], ry[K], rz[K], dl[K];
#pragma unroll
    for (int k = 0; k < K; ++k) {
      rx[k] = fmaf(G.wh[k] * lam[k], sx[k], ux[k] + dpx);
      ry[k] = fmaf(G.wh[k] * lam[k], sy[k], uy[k] + dpy);
      rz[k] = fmaf(G.wh[k] * lam[k], sz[k], uz[k] + dpz);
      b[k] = k < n ? G.d2[k] - (rx[k] * rx[k] + ry[k] * ry[k] + rz[k] * rz[k]) : 0.0f;
    }
#pragma unroll
    for (int k = 0; k < K; ++k)
#pragma unroll
      for (int l = 0; l < K; ++l) {
        const float sr = rx[k] * sx[l] + ry[k] * sy[l] + rz[k] * sz[l];
        A[k][l] = 2.0f * sr * (k == l ? wp + G.wh[k] : wp);
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
    dhx[k] = -G.wh[k] * lam[k] * sx[k]; dhy[k] = -G.wh[k] * lam[k] * sy[k]; dhz[k] = -G.wh[k] * lam[k] * sz[k];
    if (k < n) {
      const float rx = ux[k] + (dpx - dhx[k]), ry = uy[k] + (dpy - dhy[k]), rz = uz[k] + (dpz - dhz[k]);
      ok = ok && fabsf(G.d2[k] - (rx * rx + ry * ry + rz * rz)) <= kShakeTol * G.d2[k];
    }
  }
  dp[0] = dpx; dp[1] = dpy; dp[2] = dpz;
  return ok;
}
// SHAKE after the first half-kick + drift: pold = x(t), pos = unconstrained x(t + dt) -> constrained, vel += d / dt.
// `sweeps` > 1 (b200md_constrain): pold == pos, the linearisation point follows the projection.
template <int K>
__device__ __forceinline__ bool lshake_group(const LargeTopo& T, const int gi, const float* pold, float* pos, float* vel,
                                             const float inv_dt, const int sweeps) {
  const LGroup G = lgroup_load(T, gi);
  float xo[K + 1][3], xn[K + 1][3];  // row 0: the parent, 1 ..: the hydrogens (an unused row repeats the parent)
#pragma unroll
  for (int k = 0; k <= K; ++k) {
    const int a = k == 0 ? G.p : (k - 1 < G.n ? G.h[k - 1] : G.p);
#pragma unroll
    for (int c = 0; c < 3; ++c) { xo[k][c] = pold[3 * a + c]; xn[k][c] = pos[3 * a + c]; }
  }
  bool ok = true;
  float tot[K + 1][3];
#pragma unroll
  for (int k = 0; k <= K; ++k) tot[k][0] = tot[k][1] = tot[k][2] = 0.0f;
  for (int sweep = 0; sweep < sweeps; ++sweep) {
    float sx[K], sy[K], sz[K], ux[K], uy[K], uz[K], dp[3], dhx[K], dhy[K], dhz[K];
#pragma unroll
    for (int k = 0; k < K; ++k) {
      sx[k] = xo[0][0] - xo[k + 1][0]; sy[k] = xo[0][1] - xo[k + 1][1]; sz[k] = xo[0][2] - xo[k + 1][2];
      ux[k] = xn[0][0] - xn[k + 1][0]; uy[k] = xn[0][1] - xn[k + 1][1]; uz[k] = xn[0][2] - xn[k + 1][2];
    }
    ok = lshake_solve<K>(G, sx, sy, sz, ux, uy, uz, dp, dhx, dhy, dhz);
#pragma unroll
    for (int c = 0; c < 3; ++c) { xn[0][c] += dp[c]; tot[0][c] += dp[c]; }
#pragma unroll
    for (int k = 0; k < K; ++k) {
      xn[k + 1][0] += dhx[k]; xn[k + 1][1] += dhy[k]; xn[k + 1][2] += dhz[k];
      tot[k + 1][0] += dhx[k]; tot[k + 1][1] += dhy[k]; tot[k + 1][2] += dhz[k];
    }
    if (sweeps > 1) {
#pragma unroll
      for (int k = 0; k <= K; ++k)
#pragma unroll
        for (int c = 0; c < 3; ++c) xo[k][c] = xn[k][c];
    }
  }
#pragma unroll
  for (int k = 0; k <= K; ++k) {
    if (k == 0 || k - 1 < G.n) {
      const int a = k == 0 ? G.p : G.h[k - 1];
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        pos[3 * a + c] = xn[k][c];
        if (vel != nullptr && inv_dt != 0.0f) vel[3 * a + c] = fmaf(tot[k][c], inv_dt, vel[3 * a + c]);
      }
    }
  }
  return ok;
}
// RATTLE: takes the components along the rigid bonds out of the group's velocities (positions hold the constraints)
template <int K>
__device__ __forceinline__ void lrattle_group(const LargeTopo& T, const int gi, const float* pos, float* vel) {
  const LGroup G = lgroup_load(T, gi);
  const int n = G.n;
  const float wp = G.wp;
  float vp[3] = {vel[3 * G.p], vel[3 * G.p + 1], vel[3 * G.p + 2]};
  float rx[K], ry[K], rz[K], vx[K], vy[K], vz[K], A[K][K], b[K], mu[K];
#pragma unroll
  for (int k = 0; k < K; ++k) {
    const int a = k < n ? G.h[k] : G.p;
    rx[k] = pos[3 * G.p] - pos[3 * a]; ry[k] = pos[3 * G.p + 1] - pos[3 * a + 1]; rz[k] = pos[3 * G.p + 2] - pos[3 * a + 2];
    vx[k] = vel[3 * a]; vy[k] = vel[3 * a + 1]; vz[k] = vel[3 * a + 2];
    b[k] = rx[k] * (vp[0] - vx[k]) + ry[k] * (vp[1] - vy[k]) + rz[k] * (vp[2] - vz[k]);
  }
#pragma unroll
  for (int k = 0; k < K; ++k)
#pragma unroll
    for (int l = 0; l < K; ++l) {
      const float rr = rx[k] * rx[l] + ry[k] * ry[l] + rz[k] * rz[l];
      A[k][l] = rr * (k == l ? wp + G.wh[k] : wp);
      if (k >= n || l >= n) A[k][l] = k == l ? 1.0f : 0.0f;
    }
  solve_k<K>(A, b, mu);
#pragma unroll
  for (int k = 0; k < K; ++k) {
    if (k < n) {
      vp[0] = fmaf(-wp * mu[k], rx[k], vp[0]); vp[1] = fmaf(-wp * mu[k], ry[k], vp[1]); vp[2] = fmaf(-wp * mu[k], rz[k], vp[2]);
      vel[3 * G.h[k]] = fmaf(G.wh[k] * mu[k], rx[k], vx[k]);
      vel[3 * G.h[k] + 1] = fmaf(G.wh[k] * mu[k], ry[k], vy[k]);
      vel[3 * G.h[k] + 2] = fmaf(G.wh[k] * mu[k], rz[k], vz[k]);
    }
  }
  vel[3 * G.p] = vp[0]; vel[3 * G.p + 1] = vp[1]; vel[3 * G.p + 2] = vp[2];
}
// one thread per (replica, group)
__global__ void large_shake_kernel(const LargeTopo T, const long long R, const float* pold, float* pos, float* vel,
                                   const float inv_dt, const int sweeps, uint32_t* __restrict__ status) {
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < R * T.ngroups; g += (long long)gridDim.x * blockDim.x) {
    const long long r = g / T.ngroups;
    const int gi = (int)(g - r * T.ngroups);
    const size_t off = (size_t)r * T.natoms * 3;
    const bool ok = T.cmax > 1 ? lshake_group<3>(T, gi, pold + off, pos + off, vel ? vel + off : nullptr, inv_dt, sweeps)
                               : lshake_group<1>(T, gi, pold + off, pos + off, vel ? vel + off : nullptr, inv_dt, sweeps);
    if (!ok && status != nullptr) atomicOr(status + r, B200MD_FLAG_CONSTRAINT);
  }
}
__global__ void large_rattle_kernel(const LargeTopo T, const long long R, const float* __restrict__ pos, float* vel) {
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < R * T.ngroups; g += (long long)gridDim.x * blockDim.x) {
    const long long r = g / T.ngroups;
    const int gi = (int)(g - r * T.ngroups);
    const size_t off = (size_t)r * T.natoms * 3;
    if (T.cmax > 1) lrattle_group<3>(T, gi, pos + off, vel + off); else lrattle_group<1>(T, gi, pos + off, vel + off);
  }
}
// kinetic energy after RATTLE (report steps of a rigid-bond run): the per-block partials large_vv2_kernel writes otherwise
__global__ void __launch_bounds__(256)
large_ekin_kernel(const LargeTopo T, const long long R, const float* __restrict__ vel, double* __restrict__ ekin) {
  __shared__ double sred[8];
  const int tid = threadIdx.x, a = blockIdx.x * 256 + tid;
  for (long long r = blockIdx.y; r < R; r += gridDim.y) {
    double e[1] = {0.0};
    if (a < T.natoms) {
      const float* v = vel + ((size_t)r * T.natoms + a) * 3;
      e[0] = 0.5 * (double)__ldg(T.mass + a) * ((double)v[0] * v[0] + (double)v[1] * v[1] + (double)v[2] * v[2]);
    }
    lreduce<1, 256>(e, sred, tid);
    if (tid == 0) ekin[r * gridDim.x + blockIdx.x] = e[0];
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
