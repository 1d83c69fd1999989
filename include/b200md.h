/*
 * b200md.h -- C-ABI of the B200-native TorchMD force-field + integrator path.
 *
 * The reference (johncava/Molecular_Dynamics) has no FFI layer: its boundary for this path is the
 * Python class API of TorchMD_Fix/forces.py (Forces.__init__ :28-73, Forces.compute :82-344),
 * TorchMD_Fix/systems.py (System :5-85) and the upstream torchmd.integrator.Integrator.step that the
 * reference calls at V_Calculations/Test-1/tutorial.ipynb:313,374.  This header is what a
 * replacement binds underneath those classes (ctypes stub in INTEGRATION.md): plain pointers and
 * sizes, no torch types.  All device pointers are CUDA device memory on the device the topology
 * was created on; `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *
 * Units follow the reference: Angstrom, kcal/mol, amu, charges in e; integrator time in TorchMD's
 * internal unit (fs / 48.88821).
 */
#ifndef B200MD_H
#define B200MD_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200MD_MAX_ATOMS 128   /* one CTA per replica; atoms live in 4 x 32-lane blocks          */
/* Above B200MD_MAX_ATOMS a topology takes the large-solute kernels (the adk-1.0.prmtop system named at
 * V_Calculations/Test-10_diALA/potcalc.py:114 is the reference's example): same entry points and semantics; the replica
 * state lives in HBM, pairs run half-shell over rectangles of 32-atom tiles with a dense exclusion bit matrix, bonded
 * terms by windows of 64 atoms, three launches per MD step (one cooperative launch per run of steps for few replicas).
 * Every entry point works there except b200md_integrate_noalloc (B200MD_ERR_UNSUPPORTED). */
#define B200MD_MAX_ATOMS_LARGE 32768
#define B200MD_NUM_ENERGIES 8  /* order of Forces.energies + "external" (forces.py:40-50,100)     */

/* energy slots, in the reference's default dict order (forces.py:41-49, :100) */
enum b200md_energy_slot {
  B200MD_E_ELECTROSTATICS = 0,
  B200MD_E_LJ = 1,
  B200MD_E_BONDS = 2,
  B200MD_E_ANGLES = 3,
  B200MD_E_DIHEDRALS = 4,
  B200MD_E_14 = 5, /* stays 0: 1-4 energies are booked under lj / electrostatics (forces.py:202-226) */
  B200MD_E_IMPROPERS = 6,
  B200MD_E_EXTERNAL = 7 /* SMD end-to-end bias (V_Calculations/Test-6_full_system/potcalc.py:75-89) */
};

/* bits of b200md_force_opts.terms: which Forces.terms are enabled (forces.py:24-26) */
enum b200md_term_bit {
  B200MD_T_ELECTROSTATICS = 1u << 0,
  B200MD_T_LJ = 1u << 1,
  B200MD_T_BONDS = 1u << 2,
  B200MD_T_ANGLES = 1u << 3,
  B200MD_T_DIHEDRALS = 1u << 4,
  B200MD_T_14 = 1u << 5,
  B200MD_T_IMPROPERS = 1u << 6,
  B200MD_T_REPULSION = 1u << 7,   /* forces.py:416-431 (A r^-12 only): energy reported in slot B200MD_E_LJ; */
                                  /* at most one of LJ / REPULSION / REPULSIONCG per call                   */
  B200MD_T_REPULSIONCG = 1u << 8  /* forces.py:434-448 (B r^-6 only)   */
};

/* status codes returned by every entry point (0 = ok) */
enum b200md_status {
  B200MD_OK = 0,
  B200MD_ERR_INVALID_ARGUMENT = 1,
  B200MD_ERR_TOO_MANY_ATOMS = 2,
  B200MD_ERR_CUDA = 3,
  B200MD_ERR_NAN_COORDINATES = 4, /* forces.py:91-92 -> RuntimeError("Found NaN coordinates.") */
  B200MD_ERR_UNSUPPORTED = 5
};

/* per-replica status flags written by the kernels (uint32 per replica) */
#define B200MD_FLAG_NAN_POS 1u
#define B200MD_FLAG_CONSTRAINT 2u      /* integrator, rigid_bonds: SHAKE / RATTLE did not converge for a hydrogen group */
#define B200MD_FLAG_FORCE_RANGE 4u     /* integrator: the bonded force on an atom exceeded 4096 kcal/mol/A (fixed-point range) */

/*
 * Host-side description of one molecule's parameters: the flattened content of the `parameters`
 * object read by Forces (attribute list: SURVEY.md section 8 row a14).  All pointers are HOST
 * memory and only read during b200md_topology_create.
 */
typedef struct b200md_topology_desc {
  int32_t natoms;
  int32_t ntypes;
  const int32_t* atom_type; /* [natoms]  par.mapped_atom_types (forces.py:284)              */
  const float* charge;      /* [natoms]  par.charges (forces.py:215,273); NULL -> zeros     */
  const float* mass;        /* [natoms]  par.masses  (forces.py:60)                         */
  const float* lj_a;        /* [ntypes*ntypes] par.A (forces.py:285)                        */
  const float* lj_b;        /* [ntypes*ntypes] par.B (forces.py:286)                        */

  int32_t nbonds;
  const int32_t* bond_idx; /* [nbonds*2]   par.bonds       (forces.py:111-116) */
  const float* bond_prm;   /* [nbonds*2]   k0, req         (forces.py:495-496) */

  int32_t nangles;
  const int32_t* angle_idx; /* [nangles*3]  par.angles      (forces.py:134-139) */
  const float* angle_prm;   /* [nangles*2]  k0, theta0      (forces.py:505-506) */

  int32_t ndihedrals;
  const int32_t* dihedral_idx; /* [ndihedrals*4] par.dihedrals (forces.py:147-156) */
  /* flattened par.dihedral_params slices (forces.py:559-576): term t acts on dihedral
   * term_owner[t] with (k0, phi0, per); term_harmonic[t] != 0 iff the slice it came from failed
   * torch.all(per > 0), i.e. the whole slice takes the CHARMM harmonic branch (forces.py:565-576) */
  int32_t ndihedral_terms;
  const int32_t* dihedral_term_owner;    /* [ndihedral_terms]   */
  const float* dihedral_term_prm;        /* [ndihedral_terms*3] */
  const int32_t* dihedral_term_harmonic; /* [ndihedral_terms]   */

  int32_t n14;
  const int32_t* idx14; /* [n14*2]  par.idx14               (forces.py:176-180) */
  const float* prm14;   /* [n14*4]  A, B, scnb, scee        (forces.py:197-200) */

  int32_t nimpropers;
  const int32_t* improper_idx; /* [nimpropers*4] par.impropers (forces.py:228-237) */
  int32_t nimproper_terms;
  const int32_t* improper_term_owner;
  const float* improper_term_prm;
  const int32_t* improper_term_harmonic;

  /* non-bonded pair list = Forces.ava_idx (forces.py:62-68, 346-355): all i<j minus exclusions,
   * row-major.  Passed explicitly so the integer artefact the reference builds is the one used. */
  int32_t npairs;
  const int32_t* pair_idx; /* [npairs*2] */
} b200md_topology_desc;

/* Forces(...) constructor kwargs (forces.py:28-38) */
typedef struct b200md_force_opts {
  uint32_t terms;           /* OR of b200md_term_bit                                   */
  float cutoff;             /* <= 0: no cutoff (cutoff=None)                           */
  int32_t rfa;              /* reaction-field electrostatics (needs cutoff)            */
  float solvent_dielectric; /* default 78.5                                            */
  float switch_dist;        /* <= 0: no switching (switch_dist=None)                   */
} b200md_force_opts;

/* Steered-MD end-to-end harmonic bias: E = 0.5 fk (c - |r[a2]-r[a1]|)^2 with
 * c = center0 + center_rate * step (integrator) or per-replica centers[] (single-shot compute). */
typedef struct b200md_smd_opts {
  int32_t enabled;
  int32_t atom1, atom2;
  float fk;
  float center0;
  float center_rate;     /* added per integrator step; ignored by b200md_forces_compute       */
  const float* centers;  /* DEVICE [R] per-replica centre for b200md_forces_compute, or NULL  */
} b200md_smd_opts;

/* Integrator(systems, forces, timestep, device, gamma, T).step(niter) -- upstream TorchMD */
typedef struct b200md_integrate_opts {
  float dt;            /* timestep_fs / 48.88821                                                */
  float gamma;         /* gamma_ps^-1 / (1000/48.88821); 0 with temperature<=0 -> NVE           */
  float temperature;   /* K; <= 0 disables the Langevin kick                                    */
  uint64_t seed;       /* Philox key                                                            */
  uint64_t step0;      /* global index of the first step (Philox counter + SMD centre)          */
  int64_t replica0;    /* global index of replica 0 of this call (multi-GPU shards)             */
  int32_t nsteps;      /* niter                                                                 */
  int32_t energy_every;   /* record observables every this many steps (0: only after the last)  */
  int32_t snapshot_every; /* write positions every this many steps (0: never)                   */
  /* != 0: hold every bond with a hydrogen end at its equilibrium length (NAMD "rigidBonds all", the setting of the
   * reference's data generator, All_ML_Training_Data/210905_SMD_decaalanine/SMD/output/N/smd.namd: it is what lets
   * that run take 2 fs steps).  SHAKE on the positions after the first half step, RATTLE on the velocities after
   * the second.  Constraint set, derived from the topology at b200md_topology_create: an atom with 0.5 < mass < 3.5
   * is a hydrogen; it is tied to the partner of its bond row with the smallest req (so Urey-Bradley 1-3 rows in the
   * bond list stay harmonic), at that row's req.  Hydrogens must hang on heavy atoms, at most 3 per heavy atom
   * (else B200MD_ERR_UNSUPPORTED).  Group vectors are plain differences: keep bonded groups whole under PBC. */
  int32_t rigid_bonds;
} b200md_integrate_opts;


/* observables recorded per replica at each report: 8 energy slots + Ekin + end-to-end distance + the work done by
 * the moving SMD restraint since the first step of the call, W = sum_steps fk (c - d) dc (the W_harmonic column of
 * the data generator's colvars.traj, All_ML_Training_Data/210905_SMD_decaalanine/SMD/output/N/smd_out.colvars.traj) */
#define B200MD_NUM_OBSERVABLES 11
#define B200MD_OBS_EKIN 8
#define B200MD_OBS_E2E_DIST 9
#define B200MD_OBS_SMD_WORK 10

typedef struct b200md_topology b200md_topology; /* opaque */

/* Forces.__init__ (forces.py:28-73): pack topology for `device`; returns NULL on error. */
b200md_topology* b200md_topology_create(const b200md_topology_desc* desc, int device);
void b200md_topology_destroy(b200md_topology* topo);

/* Integer artefacts for the bit-exact checks: the exclusion bitmask rows built from pair_idx
 * (natoms x W uint32, bit j % 32 of word j / 32 of row i set iff pair (i,j) is NOT evaluated), copied to host `out`.
 * W = b200md_topology_mask_words: 4 up to B200MD_MAX_ATOMS atoms, ceil(natoms / 32) above. */
int b200md_topology_exclusion_mask(const b200md_topology* topo, uint32_t* out /* [natoms*W] */);
int b200md_topology_mask_words(const b200md_topology* topo);
int b200md_topology_natoms(const b200md_topology* topo);
/* number of rigid bonds b200md_integrate holds with rigid_bonds != 0 (degrees of freedom = 3 natoms - this),
 * or -1 when the topology has a hydrogen group the constraint kernel does not support */
int b200md_topology_num_constraints(const b200md_topology* topo);

/*
 * Forces.compute (forces.py:82-344) for R replicas: zeroes and fills forces, writes per-term
 * energies.  pos/forces: DEVICE float [R][natoms][3] contiguous; box: DEVICE float [R][3][3] or
 * NULL (= zeros: vacuum); a replica with a non-zero box diagonal is evaluated with the minimum-image convention in
 * every term (wrap_dist, forces.py:107,358-363; off-diagonal entries are ignored like in the reference); energies: DEVICE
 * double [R][8], every term evaluated in FP64 from the fp32 positions (the reference run in torch.double on the same
 * inputs), or NULL: forces only (what a backward pass needs; skips the energy pass); status: DEVICE uint32 [R] or
 * NULL.  Asynchronous on `stream`.
 */
int b200md_forces_compute(const b200md_topology* topo, const b200md_force_opts* opts,
                          const b200md_smd_opts* smd, int64_t nreplicas, const float* pos,
                          const float* box, float* forces, double* energies, uint32_t* status,
                          void* stream);

/*
 * Integrator.step(niter) fused with Forces.compute for R replicas (one CTA per replica keeps
 * pos/vel/forces on chip for all nsteps).  pos/vel/forces are updated in place.  `forces` must hold
 * the forces at `pos` on entry (as left by a previous compute / integrate call).
 * observables: DEVICE double [nreports][R][B200MD_NUM_OBSERVABLES], nreports =
 *   (energy_every > 0 ? ceil(nsteps / energy_every) : 1): one record every energy_every steps and one after the
 *   last step of the call when nsteps is not a multiple; snapshots: DEVICE float
 *   [nsteps / snapshot_every][R][natoms][3] or NULL.
 */
int b200md_integrate(const b200md_topology* topo, const b200md_force_opts* fopts,
                     const b200md_smd_opts* smd, const b200md_integrate_opts* iopts,
                     int64_t nreplicas, float* pos, float* vel, float* forces, const float* box,
                     double* observables, float* snapshots, uint32_t* status, void* stream);

/*
 * b200md_integrate without a single allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same arguments, semantics and results as b200md_integrate (positions, velocities and forces bit-identical;
 * report energies evaluated inside the fused kernel instead of by the energy kernel behind it: same FP64 arithmetic).
 * The scheduler's scratch (ticket counter, per-replica slice counts, carried SMD work) lives in a file-scope device array
 * of 786 688 bytes that every GPU loading the library holds until the process exits; it calls no cudaMalloc / cudaFree or
 * allocating library routine.  Limits: nreplicas <= B200MD_NOALLOC_MAX_REPLICAS and natoms <= B200MD_MAX_ATOMS, else
 * B200MD_ERR_UNSUPPORTED with a message naming the limit.  Every call shares the array: calls that overlap on the GPU
 * (two streams, two threads) race on it -- serialise them.  The first call per device resolves the array's device address
 * (cudaGetSymbolAddress): make that call before starting a graph capture.
 */
#define B200MD_NOALLOC_MAX_REPLICAS 65536
int b200md_integrate_noalloc(const b200md_topology* topo, const b200md_force_opts* fopts,
                             const b200md_smd_opts* smd, const b200md_integrate_opts* iopts,
                             int64_t nreplicas, float* pos, float* vel, float* forces, const float* box,
                             double* observables, float* snapshots, uint32_t* status, void* stream);

/* Energy.calc_phi of the ML callers (cGAN/cGAN_D_pos-2.3.2/cgan.py:166-207; the angle of forces.py:541-552):
 * phi[r][d] = -atan2(sin, cos) of proper dihedral par.dihedrals[d], for every replica; minimum image with the box
 * diagonal when it is not zero.  pos DEVICE float [R][natoms][3], box DEVICE float [R][9] or NULL,
 * phi DEVICE float [R][ndihedrals]. */
int b200md_topology_ndihedrals(const b200md_topology* topo);
int b200md_dihedral_angles(const b200md_topology* topo, int64_t nreplicas, const float* pos, const float* box,
                           float* phi, void* stream);
/* its backward for autograd (cgan.py:351-360 differentiates through the Energy adapter):
 * grad_pos[r][a][:] = sum_d grad_phi[r][d] * d phi[r][d] / d pos[r][a]; at most 512 dihedrals. */
int b200md_dihedral_angles_backward(const b200md_topology* topo, int64_t nreplicas, const float* pos, const float* box,
                                    const float* grad_phi, float* grad_pos, void* stream);

/* Puts a start structure on the constraint surface of b200md_integrate's rigid_bonds option: positions are projected
 * along their present X-H bond vectors (SHAKE, relinearised 4 times), the bond-parallel components are taken out of
 * the velocities (RATTLE).  In place; vel and status may be NULL.  (The first rigid step of b200md_integrate does
 * the same projection on a start that is off the surface; this call does it without stepping.) */
int b200md_constrain(const b200md_topology* topo, int64_t nreplicas, float* pos, float* vel, uint32_t* status,
                     void* stream);

/* Host-buffer convenience (the end-to-end call a non-torch host would make): copies pos in,
 * runs b200md_forces_compute, copies forces/energies out; synchronous. */
int b200md_forces_compute_host(const b200md_topology* topo, const b200md_force_opts* opts,
                               const b200md_smd_opts* smd, int64_t nreplicas, const float* pos_host,
                               const float* box_host, float* forces_host, double* energies_host,
                               uint32_t* status_host);

/* Batched FIRE minimiser, the update between two force evaluations as one kernel (the batched form of the tutorial's
 * minimize_bfgs call, V_Calculations/Test-1/tutorial.ipynb cell 14; upstream torchmd.minimizers is not part of the
 * reference).  state: DEVICE float [R][4] = (dt, alpha, steps since the last uphill move, max|F| of this call's forces):
 * initialise to (dt_start, alpha_start, 0, 0).  pos / vel are updated in place from `forces` (= forces at pos); the
 * caller evaluates the forces at the new positions (b200md_forces_compute with energies = NULL) and calls again, reading
 * state[:, 3] every few iterations for the convergence test max|F| <= fmax. */
typedef struct b200md_fire_opts {
  float dt_max, max_move, f_inc, f_dec, alpha_start, f_alpha;
  int32_t n_min;
} b200md_fire_opts;
int b200md_fire_update(int device, int64_t nreplicas, int32_t natoms, float* pos, float* vel, const float* forces,
                       float* state, const b200md_fire_opts* opts, void* stream);

/* FP32 FMA throughput micro-benchmark used as the roofline denominator: runs `iters` dependent
 * FMA chains on every SM and returns achieved TFLOP/s (<0 on error). */
double b200md_fp32_peak_tflops(int device, int iters);

/* number of kernels this library has launched since load (bench.py's gpu_launches claim) */
int64_t b200md_launch_count(void);

const char* b200md_last_error(void);
const char* b200md_version(void);

#ifdef __cplusplus
}
#endif
#endif /* B200MD_H */
