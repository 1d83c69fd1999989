/*
 * md_oracle.c -- CPU restatement (FP64, scalar C) of the reference's force-field + integrator path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in molecular_dynamics_b200/ may import, link or call this
 * file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * do, and only as the checker.  The product path is the CUDA library and fails loudly without it.
 *
 * What it follows (reference = /root/reference at survey time):
 *   forces      TorchMD_Fix/forces.py  Forces.compute :82-344, wrap_dist :358-363,
 *               calculate_distances :366-370, ELEC_FACTOR :373-376, evaluate_LJ(_internal) :379-413,
 *               evaluate_repulsion(_CG) :416-448, evaluate_electrostatics :451-489,
 *               evaluate_bonds :492-501, evaluate_angles :504-537, evaluate_torsion :540-602
 *   SMD bias    V_Calculations/Test-6_full_system/potcalc.py:75-89 (energy; force = -dE/dr)
 *   integrator  upstream torchmd.integrator (NOT vendored by the reference, version unpinned; used
 *               at V_Calculations/Test-1/tutorial.ipynb:212,313,374): velocity Verlet with an
 *               optional Langevin kick, restated from SURVEY.md section 3.2.
 *
 * Pinning: oracle_forces is checked in tests/test_oracle.py against (1) the notebook known-answer
 * vector V_Calculations/Test-1/tutorial.ipynb:231-271 and (2) fixtures under tests/golden/ produced
 * by the reference's own forces.py (imported unmodified by tools/make_golden.py) in fp32 and fp64.
 * The integrator has no stored output anywhere in the reference: parity for it is pinned only
 * against the restated upstream loop driving the reference's forces.py ("upstream, unpinned").
 *
 * All arithmetic is double on the fp32-rounded inputs the GPU path receives.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/b200md.h"

#define ELEC_FACTOR 332.06371307417066 /* forces.py:373-376 with scipy.constants (CODATA 2018) */
#define BOLTZMAN 0.001987191           /* upstream torchmd.integrator */
#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

typedef struct {
  double x, y, z;
} v3;

static inline v3 sub(v3 a, v3 b) { return (v3){a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline v3 cross(v3 a, v3 b) {
  return (v3){a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
static inline double dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline v3 scale(v3 a, double s) { return (v3){a.x * s, a.y * s, a.z * s}; }
static inline double norm(v3 a) { return sqrt(dot(a, a)); }

/* wrap_dist, forces.py:358-363: no wrap iff all three box diagonal entries are zero */
static inline v3 wrap(v3 d, const double* b, int vacuum) {
  if (vacuum) return d;
  d.x -= b[0] * rint(d.x / b[0]);
  d.y -= b[1] * rint(d.y / b[1]);
  d.z -= b[2] * rint(d.z / b[2]);
  return d;
}

static inline v3 getpos(const double* p, int i) { return (v3){p[3 * i], p[3 * i + 1], p[3 * i + 2]}; }
static inline void addf(double* f, int i, v3 a, double s) {
  f[3 * i] += s * a.x;
  f[3 * i + 1] += s * a.y;
  f[3 * i + 2] += s * a.z;
}

/* evaluate_LJ_internal, forces.py:388-413 */
static void lj_eval(double dist, double aa, double bb, double scl, int use_switch, double rs, double rc,
                    double* E, double* dE) {
  double rinv1 = 1.0 / dist, rinv6 = pow(rinv1, 6), rinv12 = rinv6 * rinv6;
  double pot = (aa * rinv12 - bb * rinv6) / scl;
  double force = (-12.0 * aa * rinv12 + 6.0 * bb * rinv6) * rinv1 / scl;
  if (use_switch && dist > rs) {
    double t = (dist - rs) / (rc - rs);
    double sv = 1.0 + t * t * t * (-10.0 + t * (15.0 - t * 6.0));
    double sd = t * t * (-30.0 + t * (60.0 - t * 30.0)) / (rc - rs);
    force = sv * force + pot * sd / dist;
    pot = pot * sv;
  }
  *E = pot;
  *dE = force;
}

/* evaluate_electrostatics, forces.py:451-489 */
static void coulomb_eval(double dist, double qi, double qj, double scl, int rfa, double rc, double eps,
                         double* E, double* dE) {
  if (rfa) {
    double denom = 2.0 * eps + 1.0;
    double krf = (1.0 / (rc * rc * rc)) * (eps - 1.0) / denom;
    double crf = (1.0 / rc) * (3.0 * eps) / denom;
    double common = ELEC_FACTOR * qi * qj / scl;
    double d2 = dist * dist;
    *E = common * (1.0 / dist + krf * d2 - crf);
    *dE = common * (2.0 * krf * dist - 1.0 / d2);
  } else {
    double pot = ELEC_FACTOR * qi * qj / dist / scl;
    *E = pot;
    *dE = -pot / dist;
  }
}

/* evaluate_torsion, forces.py:540-602, for one term list (dihedrals or impropers) */
static double torsions(const double* p, const double* bx, int vac, int n, const int32_t* idx, int nterms,
                       const int32_t* owner, const float* prm, const int32_t* harmonic, double* f) {
  double etot = 0.0;
  double* pot = (double*)calloc((size_t)(n > 0 ? n : 1), sizeof(double));
  double* coeff = (double*)calloc((size_t)(n > 0 ? n : 1), sizeof(double));
  double* phi = (double*)malloc((size_t)(n > 0 ? n : 1) * sizeof(double));
  for (int d = 0; d < n; ++d) {
    const int32_t* q = idx + 4 * d;
    v3 r12 = wrap(sub(getpos(p, q[0]), getpos(p, q[1])), bx, vac);
    v3 r23 = wrap(sub(getpos(p, q[1]), getpos(p, q[2])), bx, vac);
    v3 r34 = wrap(sub(getpos(p, q[2]), getpos(p, q[3])), bx, vac);
    v3 cA = cross(r12, r23), cB = cross(r23, r34), cC = cross(r23, cA);
    double nA = norm(cA), nB = norm(cB), nC = norm(cC);
    v3 ncB = scale(cB, 1.0 / nB);
    double cosPhi = dot(cA, ncB) / nA, sinPhi = dot(cC, ncB) / nC;
    phi[d] = -atan2(sinPhi, cosPhi);
  }
  for (int t = 0; t < nterms; ++t) {
    int d = owner[t];
    double k0 = prm[3 * t], phi0 = prm[3 * t + 1], per = prm[3 * t + 2];
    if (!harmonic[t]) { /* AMBER branch, forces.py:565-569 */
      double ad = per * phi[d] - phi0;
      pot[d] += k0 * (1.0 + cos(ad));
      coeff[d] += -per * k0 * sin(ad);
    } else { /* CHARMM harmonic branch, forces.py:570-576 */
      double ad = phi[d] - phi0;
      if (ad < -M_PI) ad += 2.0 * M_PI;
      else if (ad > M_PI) ad -= 2.0 * M_PI;
      pot[d] += k0 * ad * ad;
      coeff[d] += 2.0 * k0 * ad;
    }
  }
  for (int d = 0; d < n; ++d) {
    const int32_t* q = idx + 4 * d;
    v3 r12 = wrap(sub(getpos(p, q[0]), getpos(p, q[1])), bx, vac);
    v3 r23 = wrap(sub(getpos(p, q[1]), getpos(p, q[2])), bx, vac);
    v3 r34 = wrap(sub(getpos(p, q[2]), getpos(p, q[3])), bx, vac);
    v3 cA = cross(r12, r23), cB = cross(r23, r34);
    double nA = norm(cA), nB = norm(cB);
    double nd2 = norm(r23), n2d2 = nd2 * nd2; /* OpenMM projection, forces.py:582-600 */
    double ff0 = (-coeff[d] * nd2) / (nA * nA);
    double ff1 = dot(r12, r23) / n2d2;
    double ff2 = dot(r34, r23) / n2d2;
    double ff3 = (coeff[d] * nd2) / (nB * nB);
    v3 f0 = scale(cA, ff0), f3 = scale(cB, ff3);
    v3 s = sub(scale(f0, ff1), scale(f3, ff2));
    addf(f, q[0], f0, -1.0);
    addf(f, q[1], f0, 1.0);
    addf(f, q[1], s, 1.0);
    addf(f, q[2], f3, 1.0);
    addf(f, q[2], s, -1.0);
    addf(f, q[3], f3, -1.0);
    etot += pot[d];
  }
  free(pot);
  free(coeff);
  free(phi);
  return etot;
}

/* One replica of Forces.compute; p: natoms*3 doubles; f zeroed then filled; e[8]. */
static void forces_one(const b200md_topology_desc* T, const b200md_force_opts* O, const b200md_smd_opts* S,
                       double smd_center, const double* p, const double* box9, double* f, double* e) {
  const int N = T->natoms;
  memset(f, 0, sizeof(double) * 3 * (size_t)N);
  memset(e, 0, sizeof(double) * B200MD_NUM_ENERGIES);
  double bx[3] = {0, 0, 0};
  if (box9) { bx[0] = box9[0]; bx[1] = box9[4]; bx[2] = box9[8]; } /* diagonal only, forces.py:107 */
  const int vac = (bx[0] == 0.0 && bx[1] == 0.0 && bx[2] == 0.0);
  const int has_cut = O->cutoff > 0.0f;
  const double rc = O->cutoff, rs = O->switch_dist;
  const int use_switch = has_cut && O->switch_dist > 0.0f; /* forces.py:402 */
  const uint32_t terms = O->terms;

  if ((terms & B200MD_T_BONDS) && T->nbonds) { /* forces.py:111-132 */
    for (int b = 0; b < T->nbonds; ++b) {
      int i = T->bond_idx[2 * b], j = T->bond_idx[2 * b + 1];
      v3 d = wrap(sub(getpos(p, i), getpos(p, j)), bx, vac);
      double dist = norm(d);
      if (has_cut && !(dist <= rc)) continue;
      double k0 = T->bond_prm[2 * b], d0 = T->bond_prm[2 * b + 1], x = dist - d0;
      e[B200MD_E_BONDS] += k0 * x * x;
      double c = 2.0 * k0 * x / dist;
      addf(f, i, d, -c);
      addf(f, j, d, c);
    }
  }
  if ((terms & B200MD_T_ANGLES) && T->nangles) { /* forces.py:134-145, 504-537 */
    for (int a = 0; a < T->nangles; ++a) {
      const int32_t* q = T->angle_idx + 3 * a;
      v3 r21 = wrap(sub(getpos(p, q[0]), getpos(p, q[1])), bx, vac);
      v3 r23 = wrap(sub(getpos(p, q[2]), getpos(p, q[1])), bx, vac);
      double k0 = T->angle_prm[2 * a], th0 = T->angle_prm[2 * a + 1];
      double n23i = 1.0 / norm(r23), n21i = 1.0 / norm(r21);
      double ct = dot(r23, r21) * n21i * n23i;
      ct = ct < -1.0 ? -1.0 : (ct > 1.0 ? 1.0 : ct);
      double th = acos(ct), dth = th - th0;
      e[B200MD_E_ANGLES] += k0 * dth * dth;
      double st = sqrt(1.0 - ct * ct);
      double coef = (st != 0.0) ? -2.0 * k0 * dth / st : 0.0;
      v3 f0 = scale(sub(scale(r21, ct * n21i), scale(r23, n23i)), coef * n21i);
      v3 f2 = scale(sub(scale(r23, ct * n23i), scale(r21, n21i)), coef * n23i);
      addf(f, q[0], f0, 1.0);
      addf(f, q[2], f2, 1.0);
      addf(f, q[1], f0, -1.0);
      addf(f, q[1], f2, -1.0);
    }
  }
  if ((terms & B200MD_T_DIHEDRALS) && T->ndihedrals) /* forces.py:147-174 */
    e[B200MD_E_DIHEDRALS] += torsions(p, bx, vac, T->ndihedrals, T->dihedral_idx, T->ndihedral_terms,
                                      T->dihedral_term_owner, T->dihedral_term_prm, T->dihedral_term_harmonic, f);
  if ((terms & B200MD_T_14) && T->n14) { /* forces.py:176-226: never cut, switched or RF-corrected */
    for (int k = 0; k < T->n14; ++k) {
      int i = T->idx14[2 * k], j = T->idx14[2 * k + 1];
      v3 d = wrap(sub(getpos(p, i), getpos(p, j)), bx, vac);
      double dist = norm(d), E, dE;
      const float* q = T->prm14 + 4 * k; /* A, B, scnb, scee */
      if (terms & B200MD_T_LJ) {
        lj_eval(dist, q[0], q[1], q[2], 0, 0, 0, &E, &dE);
        e[B200MD_E_LJ] += E;
        addf(f, i, d, -dE / dist);
        addf(f, j, d, dE / dist);
      }
      if (terms & B200MD_T_ELECTROSTATICS) {
        coulomb_eval(dist, T->charge ? T->charge[i] : 0.0, T->charge ? T->charge[j] : 0.0, q[3], 0, 0,
                     O->solvent_dielectric, &E, &dE);
        e[B200MD_E_ELECTROSTATICS] += E;
        addf(f, i, d, -dE / dist);
        addf(f, j, d, dE / dist);
      }
    }
  }
  if ((terms & B200MD_T_IMPROPERS) && T->nimpropers) /* forces.py:228-255 */
    e[B200MD_E_IMPROPERS] += torsions(p, bx, vac, T->nimpropers, T->improper_idx, T->nimproper_terms,
                                      T->improper_term_owner, T->improper_term_prm, T->improper_term_harmonic, f);

  const uint32_t nbmask = B200MD_T_ELECTROSTATICS | B200MD_T_LJ | B200MD_T_REPULSION | B200MD_T_REPULSIONCG;
  if ((terms & nbmask) && T->npairs) { /* forces.py:258-316 */
    for (int k = 0; k < T->npairs; ++k) {
      int i = T->pair_idx[2 * k], j = T->pair_idx[2 * k + 1];
      v3 d = wrap(sub(getpos(p, i), getpos(p, j)), bx, vac);
      double dist = norm(d);
      if (has_cut && !(dist <= rc)) continue;
      double E, dE, c = 0.0;
      int ti = T->atom_type[i], tj = T->atom_type[j];
      if (terms & B200MD_T_ELECTROSTATICS) {
        coulomb_eval(dist, T->charge[i], T->charge[j], 1.0, O->rfa, rc, O->solvent_dielectric, &E, &dE);
        e[B200MD_E_ELECTROSTATICS] += E;
        c += dE;
      }
      if (terms & B200MD_T_LJ) {
        lj_eval(dist, T->lj_a[ti * T->ntypes + tj], T->lj_b[ti * T->ntypes + tj], 1.0, use_switch, rs, rc, &E, &dE);
        e[B200MD_E_LJ] += E;
        c += dE;
      }
      /* evaluate_repulsion (forces.py:416-431): A r^-12; evaluate_repulsion_CG (:434-448): B r^-6.  Never switched.
       * The ABI reports the one enabled term of the lj family in the B200MD_E_LJ slot. */
      if (terms & B200MD_T_REPULSION) {
        double r6 = pow(1.0 / dist, 6), aa = T->lj_a[ti * T->ntypes + tj];
        e[B200MD_E_LJ] += aa * r6 * r6;
        c += -12.0 * aa * r6 * r6 / dist;
      }
      if (terms & B200MD_T_REPULSIONCG) {
        double r6 = pow(1.0 / dist, 6), bb = T->lj_b[ti * T->ntypes + tj];
        e[B200MD_E_LJ] += bb * r6;
        c += -6.0 * bb * r6 / dist;
      }
      addf(f, i, d, -c / dist);
      addf(f, j, d, c / dist);
    }
  }
  if (S && S->enabled) { /* potcalc.py:80-85: E = 0.5 fk (c - |r2 - r1|)^2 */
    v3 d = sub(getpos(p, S->atom2), getpos(p, S->atom1));
    double dist = norm(d), dev = smd_center - dist;
    e[B200MD_E_EXTERNAL] += 0.5 * S->fk * dev * dev;
    double c = S->fk * dev / dist; /* force on atom2 = +fk (c - d) * unit(r2 - r1) */
    addf(f, S->atom2, d, c);
    addf(f, S->atom1, d, -c);
  }
}

/* Energy.calc_phi of the ML callers (cGAN/cGAN_D_pos-2.3.2/cgan.py:166-207; same expressions as forces.py:541-552):
 * r12 = x0 - x1, r23 = x1 - x2, r34 = x2 - x3 (wrapped), A = r12 x r23, B = r23 x r34, Cv = r23 x A,
 * cos = A.(B/|B|)/|A|, sin = Cv.(B/|B|)/|Cv|, phi = -atan2(sin, cos).  pos double [R][N][3] so that finite
 * differences of this function check the analytic backward; phi double [R][ndihedrals]. */
int oracle_dihedral_angles(const b200md_topology_desc* T, int64_t R, const double* pos, const float* box, double* phi) {
  const int N = T->natoms;
  for (int64_t r = 0; r < R; ++r) {
    const double* p = pos + r * 3 * N;
    double b[3] = {0, 0, 0};
    if (box) { b[0] = box[r * 9]; b[1] = box[r * 9 + 4]; b[2] = box[r * 9 + 8]; }
    const int vacuum = b[0] == 0.0 && b[1] == 0.0 && b[2] == 0.0;
    for (int d = 0; d < T->ndihedrals; ++d) {
      const int32_t* q = T->dihedral_idx + 4 * d;
      v3 r12 = wrap(sub(getpos(p, q[0]), getpos(p, q[1])), b, vacuum);
      v3 r23 = wrap(sub(getpos(p, q[1]), getpos(p, q[2])), b, vacuum);
      v3 r34 = wrap(sub(getpos(p, q[2]), getpos(p, q[3])), b, vacuum);
      v3 A = cross(r12, r23), B = cross(r23, r34), Cv = cross(r23, A);
      v3 nB = scale(B, 1.0 / norm(B));
      phi[r * T->ndihedrals + d] = -atan2(dot(Cv, nB) / norm(Cv), dot(A, nB) / norm(A));
    }
  }
  return B200MD_OK;
}

/* Forces.compute for R replicas.  pos float [R][N][3] (the fp32 inputs the GPU path sees), box float
 * [R][9] or NULL, smd_centers float [R] or NULL (then S->center0).  forces double [R][N][3],
 * energies double [R][8].  Returns B200MD_ERR_NAN_COORDINATES like forces.py:91-92. */
int oracle_forces(const b200md_topology_desc* T, const b200md_force_opts* O, const b200md_smd_opts* S,
                  int64_t R, const float* pos, const float* box, const float* smd_centers, double* forces,
                  double* energies) {
  const int N = T->natoms;
  for (int64_t i = 0; i < R * N * 3; ++i)
    if (isnan(pos[i])) return B200MD_ERR_NAN_COORDINATES;
#pragma omp parallel for schedule(static)
  for (int64_t r = 0; r < R; ++r) {
    double* p = (double*)malloc(sizeof(double) * 3 * (size_t)N);
    double b9[9];
    for (int k = 0; k < 3 * N; ++k) p[k] = pos[r * 3 * N + k];
    if (box) for (int k = 0; k < 9; ++k) b9[k] = box[r * 9 + k];
    double c = (S && S->enabled) ? (smd_centers ? smd_centers[r] : S->center0) : 0.0;
    forces_one(T, O, S, c, p, box ? b9 : NULL, forces + r * 3 * N, energies + r * B200MD_NUM_ENERGIES);
    free(p);
  }
  return B200MD_OK;
}

/* ---- Philox4x32-10 + Box-Muller: the counter-based stream the CUDA Langevin kick uses ---------- */
static inline void philox_round(uint32_t c[4], const uint32_t k[2]) {
  uint64_t p0 = (uint64_t)0xD2511F53u * c[0], p1 = (uint64_t)0xCD9E8D57u * c[2];
  uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k[0], n1 = (uint32_t)p1;
  uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k[1], n3 = (uint32_t)p0;
  c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
}
void oracle_philox4x32_10(uint64_t seed, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t out[4]) {
  uint32_t k[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
  uint32_t c[4] = {c0, c1, c2, c3};
  for (int r = 0; r < 10; ++r) {
    philox_round(c, k);
    k[0] += 0x9E3779B9u;
    k[1] += 0xBB67AE85u;
  }
  memcpy(out, c, sizeof(uint32_t) * 4);
}
/* three N(0,1) for (seed, replica, atom, step): counter = {atom, replica, step_lo, step_hi};
 * uniforms are the top 24 bits (+0.5) * 2^-24 so the fp32 GPU path sees the same inputs. */
void oracle_normal3(uint64_t seed, uint64_t replica, uint32_t atom, uint64_t step, double n[3]) {
  uint32_t u[4];
  oracle_philox4x32_10(seed, atom, (uint32_t)replica, (uint32_t)step, (uint32_t)(step >> 32), u);
  double u0 = ((u[0] >> 8) + 0.5) * (1.0 / 16777216.0), u1 = ((u[1] >> 8) + 0.5) * (1.0 / 16777216.0);
  double u2 = ((u[2] >> 8) + 0.5) * (1.0 / 16777216.0), u3 = ((u[3] >> 8) + 0.5) * (1.0 / 16777216.0);
  double r0 = sqrt(-2.0 * log(u0)), r1 = sqrt(-2.0 * log(u2));
  n[0] = r0 * cos(2.0 * M_PI * u1);
  n[1] = r0 * sin(2.0 * M_PI * u1);
  n[2] = r1 * cos(2.0 * M_PI * u3);
}

/* Rigid X-H bonds (b200md_integrate_opts.rigid_bonds; NAMD "rigidBonds all" of the reference's data generator,
 * All_ML_Training_Data/210905_SMD_decaalanine/SMD/output/N/smd.namd).  No code of the reference implements this
 * (NAMD is an external program): the constraint SET follows the rule stated in include/b200md.h, the algorithm is
 * textbook SHAKE (Ryckaert, Ciccotti, Berendsen 1977) for the positions and RATTLE (Andersen 1983) for the
 * velocities, swept over a flat list of constraints in FP64 to a tolerance of 1e-12.  Parity "unpinned" for this
 * option: anchored on properties (bond lengths held, energy conserved at 2 fs, the reference's own restart files
 * satisfy the same constraint set). */
typedef struct { int i, j; double d2; } cons_t;

static int build_constraints(const b200md_topology_desc* T, cons_t* out /* [natoms] */) {
  int n = 0;
  for (int a = 0; a < T->natoms; ++a) {
    if (!(T->mass[a] > 0.5f && T->mass[a] < 3.5f)) continue;
    int best = -1;
    for (int b = 0; b < T->nbonds; ++b)
      if ((T->bond_idx[2 * b] == a || T->bond_idx[2 * b + 1] == a) &&
          (best < 0 || T->bond_prm[2 * b + 1] < T->bond_prm[2 * best + 1]))
        best = b;
    if (best < 0) continue;
    int other = T->bond_idx[2 * best] == a ? T->bond_idx[2 * best + 1] : T->bond_idx[2 * best];
    double d = T->bond_prm[2 * best + 1];
    out[n++] = (cons_t){other, a, d * d};
  }
  return n;
}

/* p: unconstrained x(t+dt), p_old: x(t); moves p onto the constraints, corrects v by dx/dt */
static int shake(const cons_t* c, int nc, const float* mass, const double* p_old, double* p, double* v, double dt) {
  for (int it = 0; it < 1000; ++it) {
    int done = 1;
    for (int k = 0; k < nc; ++k) {
      const int i = c[k].i, j = c[k].j;
      v3 r = sub(getpos(p, i), getpos(p, j)), s = sub(getpos(p_old, i), getpos(p_old, j));
      double diff = c[k].d2 - dot(r, r);
      if (fabs(diff) > 1e-12 * c[k].d2) {
        done = 0;
        double wi = 1.0 / mass[i], wj = 1.0 / mass[j];
        double g = diff / (2.0 * dot(s, r) * (wi + wj));
        double di[3] = {g * wi * s.x, g * wi * s.y, g * wi * s.z}, dj[3] = {-g * wj * s.x, -g * wj * s.y, -g * wj * s.z};
        for (int x = 0; x < 3; ++x) {
          p[3 * i + x] += di[x]; v[3 * i + x] += di[x] / dt;
          p[3 * j + x] += dj[x]; v[3 * j + x] += dj[x] / dt;
        }
      }
    }
    if (done) return 0;
  }
  return 1;
}

static int rattle(const cons_t* c, int nc, const float* mass, const double* p, double* v) {
  for (int it = 0; it < 1000; ++it) {
    int done = 1;
    for (int k = 0; k < nc; ++k) {
      const int i = c[k].i, j = c[k].j;
      v3 r = sub(getpos(p, i), getpos(p, j)), u = sub(getpos(v, i), getpos(v, j));
      double rv = dot(r, u);
      if (fabs(rv) > 1e-13 * c[k].d2) {
        done = 0;
        double wi = 1.0 / mass[i], wj = 1.0 / mass[j];
        double g = rv / (c[k].d2 * (wi + wj));
        v[3 * i] -= g * wi * r.x; v[3 * i + 1] -= g * wi * r.y; v[3 * i + 2] -= g * wi * r.z;
        v[3 * j] += g * wj * r.x; v[3 * j + 1] += g * wj * r.y; v[3 * j + 2] += g * wj * r.z;
      }
    }
    if (done) return 0;
  }
  return 1;
}

/* b200md_constrain: start structure onto the constraint surface (positions along their present bond vectors,
 * relinearised 4 times like the kernel; then the velocities).  pos / vel double [R][N][3] in place, vel may be NULL. */
int oracle_constrain(const b200md_topology_desc* T, int64_t R, double* pos, double* vel) {
  const int N = T->natoms;
  cons_t* cons = (cons_t*)malloc(sizeof(cons_t) * (size_t)(N > 0 ? N : 1));
  const int nc = build_constraints(T, cons);
  double* p_old = (double*)malloc(sizeof(double) * 3 * (size_t)N);
  double* vdummy = (double*)calloc(3 * (size_t)N, sizeof(double));
  int failed = 0;
  for (int64_t r = 0; r < R && nc > 0; ++r) {
    double* p = pos + r * 3 * N;
    for (int sweep = 0; sweep < 4; ++sweep) {
      memcpy(p_old, p, sizeof(double) * 3 * (size_t)N);
      if (shake(cons, nc, T->mass, p_old, p, vdummy, 1.0)) failed = 1;
    }
    if (vel && rattle(cons, nc, T->mass, p, vel + r * 3 * N)) failed = 1;
  }
  free(cons); free(p_old); free(vdummy);
  return failed ? B200MD_ERR_UNSUPPORTED : B200MD_OK;
}

/* Integrator.step(niter) of upstream TorchMD, FP64 state.
 * pos/vel/forces double [R][N][3] in/out (forces must be valid at entry), box float [R][9] or NULL.
 * observables double [nreports][R][11] as in b200md_integrate; snapshots double [nsnap][R][N][3]. */
int oracle_integrate(const b200md_topology_desc* T, const b200md_force_opts* O, const b200md_smd_opts* S,
                     const b200md_integrate_opts* I, int64_t R, double* pos, double* vel, double* forces,
                     const float* box, double* observables, double* snapshots) {
  const int N = T->natoms;
  const double dt = I->dt, gamma = I->gamma;
  const int langevin = I->temperature > 0.0f;
  const int nrep = I->energy_every > 0 ? (I->nsteps + I->energy_every - 1) / I->energy_every : 1;
  const int nsnap = I->snapshot_every > 0 ? I->nsteps / I->snapshot_every : 0;
  (void)nrep;
  (void)nsnap;
  cons_t* cons = (cons_t*)malloc(sizeof(cons_t) * (size_t)(N > 0 ? N : 1));
  const int ncons = I->rigid_bonds ? build_constraints(T, cons) : 0;
  int failed = 0;
#pragma omp parallel for schedule(static)
  for (int64_t r = 0; r < R; ++r) {
    double *p = pos + r * 3 * N, *v = vel + r * 3 * N, *f = forces + r * 3 * N;
    double b9[9], e[B200MD_NUM_ENERGIES], work = 0.0;
    double* p_old = ncons ? (double*)malloc(sizeof(double) * 3 * (size_t)N) : NULL;
    if (box) for (int k = 0; k < 9; ++k) b9[k] = box[r * 9 + k];
    for (int s = 0; s < I->nsteps; ++s) {
      const uint64_t gstep = I->step0 + (uint64_t)s;
      if (ncons) memcpy(p_old, p, sizeof(double) * 3 * (size_t)N);
      for (int a = 0; a < N; ++a) { /* _first_VV */
        double m = T->mass[a];
        for (int c = 0; c < 3; ++c) {
          double acc = f[3 * a + c] / m;
          p[3 * a + c] += v[3 * a + c] * dt + 0.5 * acc * dt * dt;
          v[3 * a + c] += 0.5 * dt * acc;
        }
      }
      if (ncons && shake(cons, ncons, T->mass, p_old, p, v, dt)) failed = 1;
      /* SMD centre of the step whose forces are being computed: c0 + rate * (global step + 1) */
      double center = (S && S->enabled) ? (double)S->center0 + (double)S->center_rate * (double)(gstep + 1) : 0.0;
      forces_one(T, O, S, center, p, box ? b9 : NULL, f, e);
      if (S && S->enabled) /* work of the moving restraint: dW = dE/dc dc = fk (c - d) dc */
        work += (double)S->fk * (center - norm(sub(getpos(p, S->atom2), getpos(p, S->atom1)))) * (double)S->center_rate;
      if (langevin) {
        for (int a = 0; a < N; ++a) {
          double m = T->mass[a], n3[3];
          double vc = sqrt(2.0 * gamma / m * BOLTZMAN * (double)I->temperature * dt);
          oracle_normal3(I->seed, (uint64_t)(I->replica0 + r), (uint32_t)a, gstep, n3);
          for (int c = 0; c < 3; ++c) v[3 * a + c] += -gamma * v[3 * a + c] * dt + vc * n3[c];
        }
      }
      for (int a = 0; a < N; ++a) { /* _second_VV */
        double m = T->mass[a];
        for (int c = 0; c < 3; ++c) v[3 * a + c] += 0.5 * dt * f[3 * a + c] / m;
      }
      if (ncons && rattle(cons, ncons, T->mass, p, v)) failed = 1;
      double ekin = 0.0;
      for (int a = 0; a < N; ++a)
        for (int c = 0; c < 3; ++c) ekin += 0.5 * (double)T->mass[a] * v[3 * a + c] * v[3 * a + c];
      const int last = (s == I->nsteps - 1);
      int report = (I->energy_every > 0 && (s + 1) % I->energy_every == 0) || last;
      if (report && observables) {
        int slot = I->energy_every > 0 ? s / I->energy_every : 0;
        double* o = observables + ((int64_t)slot * R + r) * B200MD_NUM_OBSERVABLES;
        for (int k = 0; k < B200MD_NUM_ENERGIES; ++k) o[k] = e[k];
        o[B200MD_OBS_EKIN] = ekin;
        o[B200MD_OBS_E2E_DIST] = (S && S->enabled) ? norm(sub(getpos(p, S->atom2), getpos(p, S->atom1))) : 0.0;
        o[B200MD_OBS_SMD_WORK] = work;
      }
      if (I->snapshot_every > 0 && snapshots && (s + 1) % I->snapshot_every == 0) {
        int slot = (s + 1) / I->snapshot_every - 1;
        memcpy(snapshots + ((int64_t)slot * R + r) * 3 * N, p, sizeof(double) * 3 * (size_t)N);
      }
    }
    free(p_old);
  }
  free(cons);
  return failed ? B200MD_ERR_UNSUPPORTED : B200MD_OK;
}
