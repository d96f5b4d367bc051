/*
 * mpmc_b200.h -- C ABI of the B200-native energy engine for MPMC's per-step
 * energy hot path (LJ + Ewald + iterative Thole polarization).
 *
 * Plain pointers and sizes only; no torch / CUDA types appear in any signature,
 * so the library can be bound from C (the reference's host language), ctypes,
 * cgo, JNI...  All arithmetic is IEEE fp64.  Units follow the reference:
 * energies in K, lengths in Angstrom, charges in sqrt(K*A) (e * 408.7816,
 * reference src/include/defines.h:45), polarizabilities in A^3.
 *
 * Which reference interface each entry point replaces (paths relative to the
 * reference tree):
 *
 *   mpmc_b200_set_params     system_t scalar options read by the hot path
 *                            (src/include/structs.h:347-513) and pbc_t
 *                            (structs.h:100-105)
 *   mpmc_b200_set_atoms      the molecule_t -> atom_t linked lists as walked by
 *                            rebuild_arrays() (src/energy/pairs.c:12-38); the
 *                            per-pair pair_t list (structs.h:21-37) is NOT
 *                            materialised
 *   mpmc_b200_set_positions  atom_t.pos[] after an MC move
 *                            (src/mc/mc_moves.c:567-741)
 *   mpmc_b200_energy         energy()            (src/energy/energy.c:67-226)
 *                            = pairs()           (src/energy/pairs.c:293-361)
 *                            + polar()           (src/energy/polar.c:31-135)
 *                            + lj()              (src/energy/lj.c:165-276)
 *                            + coulombic()       (src/energy/coulombic.c:22-39)
 *   mpmc_b200_polar          polar() alone       (src/energy/polar.c:31-135)
 *   mpmc_b200_set_incremental, mpmc_b200_set_warm_start
 *                            no reference counterpart: opt-in accelerations of the same energy() (the
 *                            reference's TODO:2; its pair_t cache src/energy/pairs.c:243-250 is the idea)
 *   mpmc_b200_mc_steps       the body of mc()'s loop for displacement moves (src/mc/mc.c:294-353): move, energy,
 *                            Metropolis test, restore -- a batch of steps resident on the device
 *   mpmc_b200_get_atom_arrays  the per-atom fields polar() writes back:
 *                            mu, ef_static, ef_induced, ef_induced_change,
 *                            dipole_rrms (structs.h:47-49), rank_metric
 *
 * The drop-in `double energy(system_t*)` / `double polar(system_t*)` symbols of
 * the reference (src/include/function_prototypes.h:5,193) are provided by the
 * shim in mpmc_b200/shim/ on top of this ABI; see INTEGRATION.md.
 *
 * Error convention: every function returns 0 on success, non-zero on failure;
 * mpmc_b200_last_error() returns a static, human-readable message.  CUDA errors
 * are never swallowed.  There is NO CPU fallback: without a CUDA device
 * mpmc_b200_create() fails.
 */
#ifndef MPMC_B200_H
#define MPMC_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#define MPMC_B200_ABI_VERSION 3

/* damp_type: src/include/defines.h DAMPING_* enum */
enum { MPMC_DAMP_OFF = 0, MPMC_DAMP_LINEAR = 1, MPMC_DAMP_EXPONENTIAL = 2 };
/* LJ mixing rule (src/energy/pairs.c:84-211) */
enum { MPMC_MIX_LORENTZ_BERTHELOT = 0, MPMC_MIX_WALDMAN_HAGLER = 1, MPMC_MIX_HALGREN = 2, MPMC_MIX_C6 = 3 };

/* Scalar options of system_t / pbc_t that the hot path reads. */
typedef struct mpmc_b200_params {
    /* pbc_t (structs.h:100-105): basis[q][p] = component p of lattice vector q */
    double basis[3][3];
    double reciprocal_basis[3][3];
    double cutoff;
    double volume;
    /* Ewald (structs.h:424-427) */
    double ewald_alpha;
    int ewald_kmax;
    /* Feynman-Hibbs (structs.h:442) */
    int feynman_hibbs;
    int feynman_hibbs_order;
    double temperature;
    /* repulsion / dispersion */
    int rd_lrc;  /* structs.h:444 */
    int mixing;  /* MPMC_MIX_* */
    int rd_only; /* structs.h:418: skip electrostatics and polarization */
    /* cavity_autoreject_absolute (structs.h:395,406; energy.c:49-64) */
    int cavity_autoreject_absolute;
    double cavity_autoreject_scale;
    /* Thole polarization (structs.h:429-435) */
    int polarization;
    int polar_iterative;
    int damp_type;
    double polar_damp;
    int polar_max_iter;
    double polar_precision;
    double polar_gamma;
    int polar_gs, polar_gs_ranked;
    int polar_sor, polar_esor;
    int polar_palmo, polar_rrms, polar_zodid;
    int polar_ewald, polar_ewald_full, polar_wolf, polar_wolf_full;
    int polar_wolf_alpha_lookup;
    double polar_wolf_alpha;
    double polar_wolf_alpha_lookup_cutoff;
    double polar_ewald_alpha;
    /* Wolf-summed electrostatics (system_t.wolf, structs.h; coulombic.c:27-28,269-308): coulombic()
     * = coulombic_wolf() alone -- shifted pair sum, no reciprocal / self / intramolecular term.
     * Appended last so the layout of the earlier members is unchanged. */
    int es_wolf;
    /* rd_crystal (structs.h; lj.c:109-162,175-216,264-267): LJ summed over the periodic images
     * (-(order-1)..order-1)^3 of every pair (same-molecule pairs included, without their home image) inside the
     * wider cutoff 2*cutoff*(order-0.5), plus each atom's interaction with its own images.  The image terms of
     * the pairs and of the atoms themselves are reported inside energies.rd_pair. */
    int rd_crystal;
    int rd_crystal_order;
    int reserved0;
} mpmc_b200_params;

/* What energy() puts into observables_t (structs.h:152-162) plus its parts. */
typedef struct mpmc_b200_energies {
    double total;       /* observables->energy */
    double rd;          /* observables->rd_energy  (pair LJ [+FH] + LRC pair + LRC self) */
    double coulombic;   /* observables->coulombic_energy = real + reciprocal + self */
    double polar;       /* observables->polarization_energy */
    double rd_pair;     /* sum over in-cutoff pairs only */
    double rd_lrc;      /* pair + self long-range correction */
    double es_real;     /* coulombic_real(): erfc pairs [+FH] - intramolecular erf term */
    double es_recip;    /* coulombic_reciprocal() */
    double es_self;     /* coulombic_self() */
    double dipole_rrms; /* observables->dipole_rrms (get_dipole_rrms, polar.c:13-28) */
    double min_intermolecular_r; /* smallest r_img between atoms of different molecules */
    int polar_iterations;        /* nodestats->polarization_iterations */
    int iter_success;            /* system->iter_success: 1 = solver failed to converge */
    int autorejected;            /* 1 = cavity_autoreject_absolute fired; total = 1e40 */
    int reserved;
} mpmc_b200_energies;

/* Work counters of the last evaluation (for Gpair/s and roofline arithmetic). */
typedef struct mpmc_b200_counters {
    double pairs_tested;      /* unordered pairs whose geometry was evaluated by the pair kernel */
    double pairs_lj;          /* pairs inside the cutoff contributing an LJ term */
    double pairs_es;          /* pairs inside the cutoff contributing an erfc term */
    double pairs_field;       /* unordered pairs contributing to the static field */
    double pairs_dipole;      /* ordered polarizable pairs per contraction */
    double contractions;      /* dipole-tensor contractions done (iterations [+1 palmo]) */
    double kvectors;          /* reciprocal vectors */
    double recip_atoms;       /* atoms in the structure factor */
    int n_atoms, n_polarizable;
    int kernel_launches;      /* kernels of THIS library launched by the last evaluation */
    int incremental;          /* 1 = the last evaluation took the incremental path (mpmc_b200_set_incremental) */
    double pairs_induced_list; /* polar_ewald_full: ordered in-cutoff polarizable pairs held in the pair list (0: on the fly) */
} mpmc_b200_counters;

typedef struct mpmc_b200_ctx mpmc_b200_ctx;

int mpmc_b200_abi_version(void);
const char *mpmc_b200_last_error(void);

/* device = CUDA ordinal.  Fails (non-zero) when no CUDA device is usable. */
int mpmc_b200_create(mpmc_b200_ctx **ctx, int device);
void mpmc_b200_destroy(mpmc_b200_ctx *ctx);
/* One process driving `ndev` (1..8) GPUs of one NVLink / NVSwitch box: the single-process form of the multi-GPU path
 * SURVEY 7.2 / 8b plan ("one stream per device, created once at first energy()") and the reference's own open TODO
 * (src/polarization_gpu/polar_cuda.cu:32).  Rank i of the sharded evaluation lives on devices[i] (NULL: 0..ndev-1).
 * The returned context leads the group: mpmc_b200_set_params / _set_atoms / _set_positions / _energy / _polar /
 * _get_atom_arrays / _get_counters / _destroy on it act on the whole group, so the drop-in shim only changes its
 * create call (MPMC_B200_DEVICES).  Every exchange (static-field rows, induced-field partial sums of
 * every solver pass, the final scalars) is done by the library's own kernels over peer memory; there is no NCCL call.
 * Jacobi-family solvers only; Gauss-Seidel, polar_ewald_full and cavity_autoreject_absolute are refused (error 3). */
int mpmc_b200_create_group(mpmc_b200_ctx **ctx, int ndev, const int *devices);
int mpmc_b200_group_size(mpmc_b200_ctx *ctx);

/* Optional: run on a caller-owned cudaStream_t (passed as void*). NULL = own stream. */
int mpmc_b200_set_stream(mpmc_b200_ctx *ctx, void *cuda_stream);

int mpmc_b200_set_params(mpmc_b200_ctx *ctx, const mpmc_b200_params *p);

/*
 * Full (re)definition of the atom set, in the reference's list order
 * (molecule list order, then atom list order: pairs.c:28-34).  HOST pointers.
 *   pos[3*n]       atom_t.pos
 *   charge[n]      atom_t.charge          polarizability[n] atom_t.polarizability
 *   epsilon[n]     atom_t.epsilon         sigma[n]          atom_t.sigma
 *   mol_mass[n]    molecule_t.mass of the atom's molecule (Feynman-Hibbs reduced mass)
 *   mol_id[n]      index of the atom's molecule (same molecule <=> same id)
 *   frozen[n]      atom_t.frozen
 */
int mpmc_b200_set_atoms(mpmc_b200_ctx *ctx, int n, const double *pos, const double *charge,
                        const double *polarizability, const double *epsilon, const double *sigma,
                        const double *mol_mass, const int *mol_id, const int *frozen);

/* Overwrite positions of atoms [first, first+count) -- one MC displacement.  HOST pointer. */
int mpmc_b200_set_positions(mpmc_b200_ctx *ctx, int first, int count, const double *pos);

/*
 * Row sharding for multi-GPU runs: this context computes only its share of the
 * pair tiles, field rows, dipole rows and k-vector/atom chunks.  rank in
 * [0, world).  Partial results are combined by the caller (allreduce of
 * mpmc_b200_energies fields; allgather of dipole rows, see *_polar_stage).
 */
int mpmc_b200_set_shard(mpmc_b200_ctx *ctx, int rank, int world);

/* energy(): all kernels + device->host copy of the scalars; synchronous.  On a sharded context (set_shard with
 * world > 1 after mpmc_b200_peer_export / _connect, or the leader of a group) the whole staged evaluation described
 * below runs inside the library with every exchange over peer memory, and `out` / the counters are the totals over
 * all ranks (identical bits on every rank). */
int mpmc_b200_energy(mpmc_b200_ctx *ctx, mpmc_b200_energies *out);
/* Opt-in incremental evaluation (the reference recomputes everything every step, TODO:2; its own pair_t cache,
 * src/energy/pairs.c:243-250, is the same idea): with on != 0 -- or MPMC_B200_INCREMENTAL=1 in the environment when the
 * context is created -- mpmc_b200_energy() after mpmc_b200_set_positions() of at most 128 atoms re-evaluates only the
 * pairs that contain a moved atom and the moved atoms' terms of S(k), and adds the change to the last result.  Applies
 * to non-polarizable runs (or rd_only) without cavity_autoreject_absolute and rd_crystal on one GPU; everything else,
 * every set_params / set_atoms, and every 256th call (MPMC_B200_INCREMENTAL_REFRESH) is a full evaluation.  Results
 * agree with the full evaluation to rounding, not bit for bit, hence opt-in.  mpmc_b200_counters.incremental is 1 when
 * the last evaluation was incremental; energies.min_intermolecular_r is then not maintained (0). */
int mpmc_b200_set_incremental(mpmc_b200_ctx *ctx, int on);
/* Opt-in warm start of the dipole solve (the other half of the same idea): with on != 0 -- or MPMC_B200_WARM_START=1 in
 * the environment -- the tolerance-mode iteration (polar_precision != 0; thole_iterative.c:168-259) starts from the
 * converged dipoles of the previous evaluation instead of init_dipoles()'s gamma alpha E_s (thole_iterative.c:13-25).
 * It converges to the same tolerance in fewer passes; dipoles and U_pol then agree with the cold start to the solver
 * tolerance, not bit for bit, and nodestats->polarization_iterations is smaller -- hence opt-in.  Fixed-iteration
 * runs (polar_max_iter), polar_zodid, polar_ewald_full and sharded contexts are never warm-started. */
int mpmc_b200_set_warm_start(mpmc_b200_ctx *ctx, int on);
/* polar() alone: static field + solver + U_pol; synchronous. */
int mpmc_b200_polar(mpmc_b200_ctx *ctx, mpmc_b200_energies *out);

/* Enqueue energy() on the stream without any host synchronisation or copy
 * (fixed-iteration solver modes only); pair with mpmc_b200_energy_finish(). */
int mpmc_b200_energy_enqueue(mpmc_b200_ctx *ctx);
int mpmc_b200_energy_finish(mpmc_b200_ctx *ctx, mpmc_b200_energies *out);

/* Staged polarization for sharded (multi-GPU) runs, Jacobi family only.  Each call enqueues on the
 * context's stream and returns; between stages the CALLER runs the collectives (on the same
 * stream) over the buffers returned by mpmc_b200_device_buffer():
 *   mpmc_b200_energy_enqueue()                    pair tiles + k-sum partials of this shard
 *   stage FIELD                                   static field at the atoms of this shard's spatial clusters
 *        -> all-reduce(sum) of ef_static (buffer 3; rows other ranks own are zero)
 *   stage INIT                                    mu0 = gamma alpha E_s for all rows (replicated)
 *   per iteration it = 1, 2, ...:
 *     stage PASS(it)                              symmetric dipole-tensor contraction over this shard's
 *                                                 tile pairs: its contribution to E_ind of ALL rows
 *        -> all-reduce(sum) of buffer 4 (the one real exchange of the path, 3 n_pol doubles)
 *     stage UPDATE(it)                            new_mu + residual (slot 12 of buffer 2) for all rows,
 *                                                 identical on every rank
 *     on the last iteration, when polar_palmo is set:
 *       stage PALMO -> all-reduce(sum) of buffer 4 -> stage PALMO_FINISH
 *     stage COMMIT(it)                            mu = new_mu | SOR | ESOR, all rows (no dipole all-gather)
 *     (stage FAIL instead, when it reaches 128 in precision mode: mu = alpha E_s, iter_success = 1)
 *   stage ENERGY(iterations)                      U_pol partial over own rows
 *   mpmc_b200_energy_finish()                     partial energies; the caller sums them over ranks
 * This is thole_iterative() (src/polarization/thole_iterative.c:168-259) cut at the points where
 * partial induced fields have to be combined.  With MPMC_B200_K4=rows in the environment the
 * row-by-row contraction is used instead: PASS updates own rows and the caller all-gathers the dipole
 * rows of buffer 0 after COMMIT (UPDATE / PALMO_FINISH are then no-ops). */
enum { MPMC_STAGE_FIELD = 0, MPMC_STAGE_INIT = 1, MPMC_STAGE_PASS = 2, MPMC_STAGE_PALMO = 3, MPMC_STAGE_COMMIT = 4,
       MPMC_STAGE_ENERGY = 5, MPMC_STAGE_FAIL = 6, MPMC_STAGE_UPDATE = 7, MPMC_STAGE_PALMO_FINISH = 8 };
int mpmc_b200_polar_stage(mpmc_b200_ctx *ctx, int stage, int iteration);
/* 1 when the symmetric contraction (all-reduce of buffer 4) is active, 0 for the row form (all-gather of buffer 0) */
int mpmc_b200_uses_symmetric_contraction(mpmc_b200_ctx *ctx);
/* 1 when the static field is computed per spatial cluster (every rank writes scattered rows of buffer 3: combine
 * with an all-reduce(sum)), 0 for the row-tile form (contiguous row blocks: all-gather) */
int mpmc_b200_uses_cluster_field(mpmc_b200_ctx *ctx);
/* Peer-memory exchange of the symmetric contraction's partial sums (optional; one process per GPU on ONE node,
 * NVLink / NVSwitch peer access).  After set_shard + set_atoms every rank calls peer_export (allocates the
 * exchange buffer, returns its 64-byte CUDA IPC handle), the caller all-gathers the handles (rank order) and
 * every rank calls peer_connect.  While peer_active() is 1 the staged protocol needs NO collective between
 * MPMC_STAGE_PASS and MPMC_STAGE_UPDATE (nor PALMO / PALMO_FINISH): the reduce kernel publishes this rank's
 * partial field with a flag in every peer's memory and the update kernel waits for all flags, adds the partial
 * fields in rank order and updates (and, without palmo, commits) the dipoles -- one kernel instead of
 * all-reduce + update + commit.  peer_active() drops to 0 when the padded polarizable count changes (uVT):
 * export / connect again or keep using the all-reduce of buffer 4. */
int mpmc_b200_peer_export(mpmc_b200_ctx *ctx, unsigned char *handle64);
int mpmc_b200_peer_connect(mpmc_b200_ctx *ctx, int world, const unsigned char *handles /* [world][64] */);
int mpmc_b200_peer_active(mpmc_b200_ctx *ctx);
/* atom rows [begin,end) (padded, tile aligned) whose static field this shard computes */
int mpmc_b200_shard_field_rows(mpmc_b200_ctx *ctx, int *begin, int *end, int *n_atoms_padded);
/* Raw device pointer + element count of an internal buffer, for collectives
 * issued by the caller on the same stream.  which: 0 = packed dipoles mu
 * [3][n_pol_padded] (SoA), 1 = packed new dipoles, 2 = the 32 device scalars (slot 12 = max squared
 * dipole change of the last pass, as a double), 3 = static field [3][n_atoms_padded], 4 = induced field
 * partial sums [3][n_pol_padded] of the symmetric contraction. */
int mpmc_b200_device_buffer(mpmc_b200_ctx *ctx, int which, void **dev_ptr, long long *n_doubles);
/* rows [begin,end) of the packed polarizable index space owned by this shard */
int mpmc_b200_shard_rows(mpmc_b200_ctx *ctx, int *begin, int *end, int *n_pol, int *n_pol_padded);

/* Copy per-atom results back (HOST pointers, any may be NULL), reference list order:
 * mu, ef_static, ef_induced, ef_induced_change : [3*n] ; dipole_rrms, rank_metric : [n] */
int mpmc_b200_get_atom_arrays(mpmc_b200_ctx *ctx, double *mu, double *ef_static, double *ef_induced,
                              double *ef_induced_change, double *dipole_rrms, double *rank_metric);
/* The same results without the copy: pointers into the library's pinned host image ([3*n] / [n] doubles, reference
 * list order), valid until the next call on this context.  The drop-in shim scatters from here straight into atom_t
 * (src/include/structs.h:47-50) with its host threads. */
int mpmc_b200_atom_arrays_view(mpmc_b200_ctx *ctx, const double **mu, const double **ef_static, const double **ef_induced,
                               const double **ef_induced_change, const double **dipole_rrms, const double **rank_metric);
/* Sharded runs: 1 when every rank's get_atom_arrays() is complete (symmetric contraction: the all-reduced
 * field lets every rank update every row; only the write-back-only palmo rows of non-polarizable atoms stay
 * per rank), 0 when the row-owned arrays hold this rank's rows only and must be summed over ranks. */
int mpmc_b200_atom_arrays_complete(mpmc_b200_ctx *ctx);

int mpmc_b200_get_counters(mpmc_b200_ctx *ctx, mpmc_b200_counters *out);

/* Measured device time (ms, CUDA events on the library's stream) of the named
 * phase in the last evaluation: "pair", "recip", "field", "dipole", "total"; "contract" = mean
 * time of one dipole-tensor contraction (K4) launch. */
int mpmc_b200_phase_ms(mpmc_b200_ctx *ctx, const char *phase, double *ms);

/* fp64 FMA throughput probe (dependent-chain DFMA kernel), TFLOP/s; used as the
 * roofline denominator because MEASURED_PEAKS.json carries no fp64 figure. */
int mpmc_b200_measure_fp64_peak(mpmc_b200_ctx *ctx, double *tflops);

/*
 * Resident MC steps: a batch of displacement moves evaluated AND decided on the device, with no host
 * round trip between steps.  Replaces, for the steps it covers, the loop body of mc() (src/mc/mc.c:294-353):
 *   make_move() DISPLACE = translate() + rotate()   (src/mc/mc_moves.c:376-464,713-731; the random numbers -- hence the
 *                                                    translation vector, the rotation quaternion and the Metropolis
 *                                                    draw -- come from the caller, who keeps the reference's RNG)
 *   energy()                                         (src/energy/energy.c:67-226)
 *   boltzmann_factor() NVT + the Metropolis test     (src/mc/mc.c:106-114,320)
 *   restore() of the coordinates on a reject         (src/mc/mc_moves.c:744-804)
 * Step j applies move j to the CURRENT device positions exactly as translate() / rotate() do on the host (centre of
 * mass as update_com() computes it, src/energy/pairs.c:364-385; quaternion products in the reference's operation order,
 * src/main/quaternion.c:76-86; no fused multiply-add), evaluates the energy, accepts iff
 * u_accept < exp(-(E_new - E_cur) / temperature) and the solver did not fail, and on a reject puts the old coordinates
 * back.  The caller replays the same steps on its lists from the returned decisions; the drop-in shim uses this
 * speculatively -- the unmodified mc() stays in charge and every served step is checked against what make_move() really
 * did (MPMC_B200_RESIDENT_STEPS).
 * Single GPU; fixed-iteration solvers (or polarization off); no cavity_autoreject_absolute.  At most
 * MPMC_B200_MC_MAX_STEPS steps per call and MPMC_B200_MC_MAX_ATOMS atoms per moved molecule.
 */
#define MPMC_B200_MC_MAX_ATOMS 16
#define MPMC_B200_MC_MAX_STEPS 64
typedef struct mpmc_b200_mc_move {
    int first, count;        /* the molecule's atoms [first, first+count) in list order */
    int explicit_positions;  /* 1: new_pos / old_pos are given (a move the caller already made and uploaded) */
    int reserved;
    double u_accept;         /* the Metropolis draw of this step (mc.c:320) */
    double trans[3];         /* translate(): signed displacement (mc_moves.c:380-385) */
    double quat[4];          /* rotate(): x, y, z, w of quaternion_construct_axis_angle_degree() (mc_moves.c:415-420) */
    double mass[MPMC_B200_MC_MAX_ATOMS];         /* atom_t.mass, for the centre of mass */
    double new_pos[3 * MPMC_B200_MC_MAX_ATOMS];  /* explicit_positions only */
    double old_pos[3 * MPMC_B200_MC_MAX_ATOMS];  /* explicit_positions only: what a reject restores */
} mpmc_b200_mc_move;
typedef struct mpmc_b200_mc_result {
    mpmc_b200_energies e;    /* what mpmc_b200_energy() would have returned for the trial configuration */
    int accepted;            /* the device's decision */
    int reserved;
    double new_pos[3 * MPMC_B200_MC_MAX_ATOMS];  /* the trial coordinates of the moved atoms */
    double old_pos[3 * MPMC_B200_MC_MAX_ATOMS];  /* their coordinates before the move */
} mpmc_b200_mc_result;
int mpmc_b200_mc_steps(mpmc_b200_ctx *ctx, int nsteps, const mpmc_b200_mc_move *moves, double temperature,
                       double current_energy, mpmc_b200_mc_result *results);
/* Per-atom results of step `step` of the last mpmc_b200_mc_steps() call (what mpmc_b200_atom_arrays_view() would have
 * shown after that step's evaluation); pointers into pinned host memory, valid until the next call on this context. */
int mpmc_b200_mc_step_arrays_view(mpmc_b200_ctx *ctx, int step, const double **mu, const double **ef_static,
                                  const double **ef_induced, const double **ef_induced_change, const double **dipole_rrms,
                                  const double **rank_metric);

/* No counterpart in the reference.  For callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it): the reciprocal-space
 * Ewald energy -- the quantity of coulombic_reciprocal() (src/energy/coulombic.c:42-95) that mpmc_b200_energy()
 * reports as energies.es_recip -- computed by a variant of the structure-factor kernel K2 whose scratch (per-chunk
 * partial structure factors and S(k)) is a file-scope `static __device__` array of 4.9 MB instead of context
 * buffers.  The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine) and enqueues three
 * kernels on `cuda_stream` (a cudaStream_t passed as void*, NULL = the default stream) of the CURRENT device;
 * it does not synchronise.  The first call on a device resolves the array's address (cudaGetSymbolAddress): make
 * that call outside any capture.
 *   DEVICE pointers:  d_pos[3*n] atom_t.pos, d_charge[n] atom_t.charge, d_frozen[n] atom_t.frozen (NULL: none
 *                     frozen; frozen and uncharged atoms are skipped as coulombic.c:73-74 does), d_energy[1] (K).
 *   HOST values:      reciprocal_basis[9] = pbc_t.reciprocal_basis[p][q] row-major, volume, ewald_alpha, ewald_kmax.
 * Limits (the scratch is sized for them; beyond them the call fails with error 3 and a message):
 *   0 <= n <= 1,048,576 atoms, 1 <= ewald_kmax <= 7 (the reference's default).
 * Every GPU that loads the library holds the array until exit.  All calls share it and it keeps the last call's
 * contents (each call writes every slot it reads; nothing has to be cleared by the caller): calls that overlap on
 * the GPU -- different streams, concurrent graph launches -- race on it and must be serialised by the caller. */
int mpmc_b200_recip_energy_static(int n, const double *d_pos, const double *d_charge, const int *d_frozen,
                                  const double *reciprocal_basis, double volume, double ewald_alpha, int ewald_kmax,
                                  double *d_energy, void *cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* MPMC_B200_H */
