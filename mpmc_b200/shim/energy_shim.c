/*
 * energy_shim.c -- the drop-in boundary: MPMC's own entry points
 *
 *     double energy(system_t *)                 reference src/energy/energy.c:67-226
 *     double energy_no_observables(system_t *)  reference src/energy/energy.c:232-318
 *     double polar(system_t *)                  reference src/energy/polar.c:31-135
 *     void   pairs(system_t *) and the pair-list maintenance calls   src/energy/pairs.c
 *     thole_resize_matrices / thole_amatrix     src/polarization/thole_matrix.c
 *
 * implemented over the C ABI of libmpmc_b200 (include/mpmc_b200.h).  The file is compiled against
 * the reference's OWN headers (-I<reference>/src/include, same cmake_config.h) and linked with
 * the reference's unmodified host sources (main, io, mc, mersenne, histogram, ...) in place of
 * src/energy/{energy,pairs,polar}.c and src/polarization/*.c; see mpmc_b200/shim/Makefile and
 * INTEGRATION.md.  The MC driver, the parser, the RNG and the writers are the reference's.
 *
 * What the shim does per call: walk the molecule/atom lists in list order (the order
 * rebuild_arrays() defines, src/energy/pairs.c:12-38), pack the fields the hot path reads into
 * flat arrays, send only what changed since the previous call to the device, evaluate, and write
 * the results back into system_t / atom_t exactly where the reference leaves them.
 * The O(N^2) pair_t lists are never allocated: every atom->pairs stays NULL.
 *
 * No CPU fallback: options the CUDA engine does not implement end in error() + die(-1).
 */
#include <mc.h>
#include <quaternion.h>

#include "mpmc_b200.h"

static mpmc_b200_ctx *g_ctx = NULL;

/* host image of what was last sent to the device */
static int g_n = -1;
static int g_cap = 0;
static double *g_pos, *g_q, *g_alpha, *g_eps, *g_sig, *g_mmass;
static int *g_mol, *g_frozen;
static double *n_pos, *n_q, *n_alpha, *n_eps, *n_sig, *n_mmass; /* freshly packed */
static int *n_mol, *n_frozen;
static double *w_mu, *w_es, *w_ei, *w_eic, *w_rrms, *w_rank; /* write-back staging */
static mpmc_b200_params g_params;
static int g_have_params = 0;

static void shim_die(const char *what) {
    char buf[MAXLINE];
    snprintf(buf, MAXLINE, "B200: %s: %s\n", what, mpmc_b200_last_error());
    error(buf);
    die(-1);
}

static void shim_unsupported(const char *what) {
    char buf[MAXLINE];
    snprintf(buf, MAXLINE,
             "B200: %s is not implemented by the B200 energy engine (and there is no CPU fallback in this build)\n", what);
    error(buf);
    die(-1);
}

static void shim_shutdown(void) {
    if (g_ctx) mpmc_b200_destroy(g_ctx);
    g_ctx = NULL;
}

/* "0,1,2,3", "0-3" or a mix -> device ordinals; returns how many */
static int shim_parse_devices(const char *spec, int *dev, int max) {
    int n = 0;
    const char *p = spec;
    while (*p && n < max) {
        char *end;
        long a = strtol(p, &end, 10), b;
        if (end == p) break;
        b = a;
        p = end;
        if (*p == '-') {
            b = strtol(p + 1, &end, 10);
            if (end == p + 1) break;
            p = end;
        }
        for (; a <= b && n < max; a++) dev[n++] = (int)a;
        if (*p == ',') p++;
    }
    return n;
}

/* Which GPU(s) this process evaluates on:
 *   MPMC_B200_DEVICE=k        one GPU (one MC chain per GPU: the ensemble launcher sets this per process)
 *   MPMC_B200_DEVICES=0-7     ONE chain sharded over several GPUs of the box (static-field rows, dipole-tensor tile
 *                             pairs of every solver pass, pair clusters and k-vectors split by rank; exchanges over
 *                             NVLink peer memory inside the library).  MPMC_B200_GPUS=n is short for 0..n-1. */
static void shim_init(void) {
    int dev[8], ndev = 0;
    const char *env;
    if (g_ctx) return;
    if ((env = getenv("MPMC_B200_DEVICES")) != NULL)
        ndev = shim_parse_devices(env, dev, 8);
    else if ((env = getenv("MPMC_B200_GPUS")) != NULL) {
        int k, n = atoi(env);
        for (k = 0; k < n && k < 8; k++) dev[ndev++] = k;
    }
    if (ndev > 1) {
        if (mpmc_b200_create_group(&g_ctx, ndev, dev)) shim_die("cannot create the CUDA contexts of the device group");
    } else {
        int d = ndev == 1 ? dev[0] : 0;
        if (ndev == 0 && (env = getenv("MPMC_B200_DEVICE")) != NULL) d = atoi(env);
        if (mpmc_b200_create(&g_ctx, d)) shim_die("cannot create the CUDA context");
    }
    atexit(shim_shutdown);
}

/* Host threads for the O(N) duties around a call (one pass over the lists, the scatter of the results into atom_t): the
 * reference's data model -- 912-byte atoms allocated one by one -- makes both memory-latency bound, which threads hide.
 * MPMC_B200_HOST_THREADS (default: up to 8 of the cores this process may run on; 1 = the single-threaded passes). */
#ifdef _OPENMP
#include <omp.h>
#endif
static int g_threads = 0, g_threads_min = 4096; /* below this many atoms the sequential passes are faster */
static int shim_threads(void) {
    if (g_threads <= 0) {
        const char *env = getenv("MPMC_B200_HOST_THREADS");
        if (getenv("MPMC_B200_HOST_THREADS_MIN")) g_threads_min = atoi(getenv("MPMC_B200_HOST_THREADS_MIN"));
        g_threads = 1;
#ifdef _OPENMP
        g_threads = env ? atoi(env) : omp_get_num_procs();
        if (!env && g_threads > 8) g_threads = 8;
        if (g_threads < 1) g_threads = 1;
#else
        (void)env;
#endif
    }
    return g_threads;
}
/* writers.c formats with the same threads, above the same threshold */
int mpmc_b200_shim_host_threads(int *min_atoms) {
    const int t = shim_threads();
    if (min_atoms) *min_atoms = g_threads_min;
    return t;
}

static void *xrealloc(void *p, size_t bytes) {
    void *q = realloc(p, bytes ? bytes : 1);
    memnullcheck(q, bytes, __LINE__ - 1, __FILE__);
    return q;
}

/* MPMC_B200_SHIM_TIMING=1: where a drop-in energy() call spends its host time (printed at exit) */
#include <time.h>
static double t_acc[5];
static long t_calls = 0, t_fast = 0;
static int t_on = -1;
static double shim_now(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}
static double t_first_call = 0.0, t_last_call = 0.0; /* end of the 1st / of the last energy() call (ms) */
static void shim_timing_report(void) {
    /* start-up (parse, CUDA context, the first evaluation's allocations) lies before the end of the 1st call -- in an MC run
     * the initial energy.  The interval starts there, so a resident batch started by the 2nd call is inside it as a whole. */
    if (t_on == 1 && t_calls > 1)
        fprintf(stderr, "MPMC_B200_SHIM_TIMING: %.1f energy() calls per second of wall clock between the 1st and the last call "
                        "(%ld calls: the driver's own work per step included)\n",
                1e3 * (double)(t_calls - 1) / (t_last_call - t_first_call), t_calls);
    if (t_on == 1 && t_calls > 0)
        fprintf(stderr,
                "MPMC_B200_SHIM_TIMING: %ld energy() calls (%ld on the cached-sequence pass); ms per call: host lists %.3f, "
                "upload %.3f, device energy %.3f, write-back + epilogue %.3f, whole call %.3f\n",
                t_calls, t_fast, t_acc[0] / t_calls, t_acc[1] / t_calls, t_acc[2] / t_calls, t_acc[3] / t_calls, t_acc[4] / t_calls);
}

static void shim_timing_init(void) {
    if (t_on >= 0) return;
    t_on = getenv("MPMC_B200_SHIM_TIMING") ? 1 : 0;
    if (t_on) atexit(shim_timing_report);
}

static void shim_reserve(int n) {
    int cap;
    if (n <= g_cap) return;
    cap = n + n / 4 + 64;
#define GROW(ptr, type, mult) ptr = (type *)xrealloc(ptr, sizeof(type) * (size_t)(mult) * cap)
    GROW(g_pos, double, 3);
    GROW(g_q, double, 1);
    GROW(g_alpha, double, 1);
    GROW(g_eps, double, 1);
    GROW(g_sig, double, 1);
    GROW(g_mmass, double, 1);
    GROW(g_mol, int, 1);
    GROW(g_frozen, int, 1);
    GROW(n_pos, double, 3);
    GROW(n_q, double, 1);
    GROW(n_alpha, double, 1);
    GROW(n_eps, double, 1);
    GROW(n_sig, double, 1);
    GROW(n_mmass, double, 1);
    GROW(n_mol, int, 1);
    GROW(n_frozen, int, 1);
    GROW(w_mu, double, 3);
    GROW(w_es, double, 3);
    GROW(w_ei, double, 3);
    GROW(w_eic, double, 3);
    GROW(w_rrms, double, 1);
    GROW(w_rank, double, 1);
#undef GROW
    g_cap = cap;
}

/* system_t / pbc_t scalars the hot path reads -> mpmc_b200_params */
static void shim_fill_params(const system_t *s, mpmc_b200_params *P) {
    int p, q;
    memset(P, 0, sizeof(*P));
    for (p = 0; p < 3; p++)
        for (q = 0; q < 3; q++) {
            P->basis[p][q] = s->pbc->basis[p][q];
            P->reciprocal_basis[p][q] = s->pbc->reciprocal_basis[p][q];
        }
    P->cutoff = s->pbc->cutoff;
    P->volume = s->pbc->volume;
    P->ewald_alpha = s->ewald_alpha;
    P->ewald_kmax = s->ewald_kmax;
    P->feynman_hibbs = s->feynman_hibbs;
    P->feynman_hibbs_order = s->feynman_hibbs_order;
    P->temperature = s->temperature;
    P->rd_lrc = s->rd_lrc;
    P->mixing = s->waldmanhagler ? MPMC_MIX_WALDMAN_HAGLER
                                 : s->halgren_mixing ? MPMC_MIX_HALGREN : s->c6_mixing ? MPMC_MIX_C6 : MPMC_MIX_LORENTZ_BERTHELOT;
    P->rd_only = s->rd_only;
    P->es_wolf = s->wolf; /* coulombic(): coulombic.c:27-28 */
    P->rd_crystal = s->rd_crystal; /* lj(): lj.c:175-216,264-267 */
    P->rd_crystal_order = s->rd_crystal_order;
    P->cavity_autoreject_absolute = s->cavity_autoreject_absolute;
    P->cavity_autoreject_scale = s->cavity_autoreject_scale;
    P->polarization = s->polarization;
    P->polar_iterative = s->polar_iterative;
    P->damp_type = s->damp_type == DAMPING_OFF ? MPMC_DAMP_OFF : s->damp_type == DAMPING_LINEAR ? MPMC_DAMP_LINEAR : MPMC_DAMP_EXPONENTIAL;
    P->polar_damp = s->polar_damp;
    P->polar_max_iter = s->polar_max_iter;
    P->polar_precision = s->polar_precision;
    P->polar_gamma = s->polar_gamma;
    P->polar_gs = s->polar_gs;
    P->polar_gs_ranked = s->polar_gs_ranked;
    P->polar_sor = s->polar_sor;
    P->polar_esor = s->polar_esor;
    P->polar_palmo = s->polar_palmo;
    P->polar_rrms = s->polar_rrms;
    P->polar_zodid = s->polar_zodid;
    P->polar_ewald = s->polar_ewald;
    P->polar_ewald_full = s->polar_ewald_full;
    P->polar_wolf = s->polar_wolf;
    P->polar_wolf_full = s->polar_wolf_full;
    P->polar_wolf_alpha_lookup = s->polar_wolf_alpha_lookup;
    P->polar_wolf_alpha = s->polar_wolf_alpha;
    P->polar_wolf_alpha_lookup_cutoff = s->polar_wolf_alpha_lookup_cutoff;
    P->polar_ewald_alpha = s->polar_ewald_alpha;
}

/* the rd_energy / electrostatics switches of energy() (energy.c:132-176) the engine does not cover */
static void shim_check_supported(const system_t *s) {
    if (s->rd_anharmonic) shim_unsupported("rd_anharmonic");
    if (s->sg) shim_unsupported("sg (Silvera-Goldman)");
    if (s->dreiding) shim_unsupported("dreiding");
    if (s->lj_buffered_14_7) shim_unsupported("lj_buffered_14_7");
    if (s->disp_expansion) shim_unsupported("disp_expansion");
    /* all three change the LJ mixing inside pair_exclusions() (pairs.c:100-126) */
    if (s->cdvdw_exp_repulsion || s->cdvdw_sig_repulsion || s->cdvdw_9th_repulsion) shim_unsupported("cdvdw repulsion");
    /* surface() / surf_fit() evaluate lj_nopbc() / coulombic_nopbc() over atom->pairs (mc/surface.c:26-60), which this
     * build never allocates: they would return 0.0 without a word */
    if (s->ensemble == ENSEMBLE_SURF || s->ensemble == ENSEMBLE_SURF_FIT)
        shim_unsupported("ensemble surf / surf_fit (potential-energy-surface scans over the pair lists)");
    if (s->gwp) shim_unsupported("gwp");
    if (s->axilrod_teller) shim_unsupported("axilrod_teller");
    if (s->spectre) shim_unsupported("spectre");
    if (s->polarvdw) shim_unsupported("polarvdw");
}

/* ------------------------------------------------------------------------------------------------
 * O(N) duties of the reference's pair bookkeeping (src/energy/pairs.c, src/energy/energy.c:16-46)
 * ---------------------------------------------------------------------------------------------- */
int countNatoms(system_t *system) {
    molecule_t *m;
    atom_t *a;
    int N = 0;
    for (m = system->molecules; m; m = m->next)
        for (a = m->atoms; a; a = a->next) N++;
    return N;
}

void countN(system_t *system) {
    molecule_t *m;
    system->observables->N = 0;
    system->observables->spin_ratio = 0;
    for (m = system->molecules; m; m = m->next) {
        if (!(m->frozen || m->adiabatic || m->target)) {
            system->observables->N += 1.0;
            if (m->nuclear_spin == NUCLEAR_SPIN_PARA) system->observables->spin_ratio += 1.0;
        }
        if (system->ensemble == ENSEMBLE_NVE) system->N = system->observables->N;
    }
}

/* atom_array[] / molecule_array[] in list order: other reference code indexes them */
void rebuild_arrays(system_t *system) {
    molecule_t *m;
    atom_t *a;
    int n = system->natoms;
    free(system->atom_array);
    free(system->molecule_array);
    system->molecule_array = malloc((n ? n : 1) * sizeof(molecule_t *));
    memnullcheck(system->molecule_array, n * sizeof(molecule_t *), __LINE__ - 1, __FILE__);
    system->atom_array = malloc((n ? n : 1) * sizeof(atom_t *));
    memnullcheck(system->atom_array, n * sizeof(atom_t *), __LINE__ - 1, __FILE__);
    n = 0;
    for (m = system->molecules; m; m = m->next)
        for (a = m->atoms; a; a = a->next) {
            system->molecule_array[n] = m;
            system->atom_array[n] = a;
            n++;
        }
    system->natoms = n;
}

/* molecular centre of mass; also refreshes molecule->mass, which the Feynman-Hibbs reduced mass
 * reads (src/energy/pairs.c:364-385, src/energy/lj.c:22-23) */
void update_com(molecule_t *molecules) {
    int i;
    molecule_t *m;
    atom_t *a;
    for (m = molecules; m; m = m->next) {
        for (i = 0; i < 3; i++) m->com[i] = 0;
        if (!(m->spectre || m->target)) {
            for (a = m->atoms, m->mass = 0; a; a = a->next) {
                m->mass += a->mass;
                for (i = 0; i < 3; i++) m->com[i] += a->mass * a->pos[i];
            }
            for (i = 0; i < 3; i++) m->com[i] /= m->mass;
        }
    }
}

/* pairs(): the reference fills N(N-1)/2 pair_t here (src/energy/pairs.c:293-361).  The device
 * kernels recompute exclusions, mixing and minimum images in registers, so only the O(N) duties
 * remain: pointer arrays, centres of mass, wrapped coordinates.  (The GS rank metric is computed
 * on the device and written back by the evaluation.) */
void pairs(system_t *system) {
    shim_check_supported(system); /* also reached from surface() (mc/surface.c:26), which never calls energy() */
    system->natoms = countNatoms(system);
    rebuild_arrays(system);
    update_com(system->molecules);
    wrapall(system->molecules, system->pbc);
}

/* pair-list maintenance entry points called by the MC driver (src/io/input.c:1783-1792,
 * src/mc/mc_moves.c:664-775): nothing to maintain, every atom->pairs stays NULL */
void setup_pairs(system_t *system) { shim_check_supported(system); }
void flag_all_pairs(system_t *system) { (void)system; }
void update_pairs_insert(system_t *system) { (void)system; }
void update_pairs_remove(system_t *system) { (void)system; }
void unupdate_pairs_insert(system_t *system) { (void)system; }
void unupdate_pairs_remove(system_t *system) { (void)system; }

/* minimum_image() keeps its reference semantics for the few callers outside the hot path
 * (src/energy/axilrod_teller.cpp:132-146 passes a scratch pair_t): src/energy/pairs.c:230-290 */
void minimum_image(system_t *system, atom_t *atom_i, atom_t *atom_j, pair_t *pair_ptr) {
    int p, q;
    double img[3], d[3], di[3], r2 = 0, ri2 = 0;
    pair_ptr->recalculate_energy = 1;
    for (p = 0; p < 3; p++) {
        d[p] = atom_i->pos[p] - atom_j->pos[p];
        pair_ptr->d_prev[p] = d[p];
    }
    for (p = 0; p < 3; p++) {
        for (q = 0, img[p] = 0; q < 3; q++) img[p] += system->pbc->reciprocal_basis[q][p] * d[q];
        img[p] = rint(img[p]);
    }
    for (p = 0; p < 3; p++) {
        for (q = 0, di[p] = 0; q < 3; q++) di[p] += system->pbc->basis[q][p] * img[q];
        di[p] = d[p] - di[p];
        r2 += d[p] * d[p];
        ri2 += di[p] * di[p];
    }
    pair_ptr->r = sqrt(r2);
    if (isnan(sqrt(ri2)) != 0) {
        pair_ptr->rimg = pair_ptr->r;
        for (p = 0; p < 3; p++) pair_ptr->dimg[p] = d[p];
    } else {
        pair_ptr->rimg = sqrt(ri2);
        for (p = 0; p < 3; p++) pair_ptr->dimg[p] = di[p];
    }
}

/* The dense 3N x 3N matrices are never built (src/polarization/thole_matrix.c:149-191): leaving
 * both NULL makes free_matrices() a no-op (src/main/cleanup.c:138-139). */
void thole_resize_matrices(system_t *system) {
    system->A_matrix = NULL;
    system->B_matrix = NULL;
    system->checkpoint->thole_N_atom = system->natoms;
}
void thole_amatrix(system_t *system) {
    (void)system;
    shim_unsupported("thole_amatrix (dense A matrix; used only by disp_expansion / vdw)");
}

/* ------------------------------------------------------------------------------------------------
 * host lists -> device image
 * ---------------------------------------------------------------------------------------------- */
/* flat image of the lists -> n_* arrays (list walk; the fast path of shim_fast_pass() fills the same arrays) */
static int shim_pack(system_t *system) {
    molecule_t *m;
    atom_t *a;
    int n = 0, molidx = 0;
    shim_reserve(system->natoms);
    for (m = system->molecules; m; m = m->next, molidx++)
        for (a = m->atoms; a; a = a->next, n++) {
            n_pos[3 * n] = a->pos[0];
            n_pos[3 * n + 1] = a->pos[1];
            n_pos[3 * n + 2] = a->pos[2];
            n_q[n] = a->charge;
            n_alpha[n] = a->polarizability;
            n_eps[n] = a->epsilon;
            n_sig[n] = a->sigma;
            n_mmass[n] = m->mass;
            n_mol[n] = molidx;
            n_frozen[n] = a->frozen;
        }
    return n;
}

/* parameters + the packed image (n atoms in n_*) -> device, sending only what changed */
static void shim_upload(system_t *system, int n) {
    mpmc_b200_params P;
    int i, same, first, last;

    shim_init();
    shim_check_supported(system);

    shim_fill_params(system, &P);
    if (!g_have_params || memcmp(&P, &g_params, sizeof(P)) != 0) {
        if (mpmc_b200_set_params(g_ctx, &P)) shim_die("set_params");
        g_params = P;
        g_have_params = 1;
    }

    /* a translation / rotation changes positions only; an insertion, a removal or a restored
     * molecule changes the atom count or the per-atom parameters and needs the full image */
    same = (n == g_n) && !memcmp(n_q, g_q, sizeof(double) * n) && !memcmp(n_alpha, g_alpha, sizeof(double) * n) &&
           !memcmp(n_eps, g_eps, sizeof(double) * n) && !memcmp(n_sig, g_sig, sizeof(double) * n) &&
           !memcmp(n_mmass, g_mmass, sizeof(double) * n) && !memcmp(n_mol, g_mol, sizeof(int) * n) &&
           !memcmp(n_frozen, g_frozen, sizeof(int) * n);
    if (!same) {
        if (mpmc_b200_set_atoms(g_ctx, n, n_pos, n_q, n_alpha, n_eps, n_sig, n_mmass, n_mol, n_frozen)) shim_die("set_atoms");
        memcpy(g_q, n_q, sizeof(double) * n);
        memcpy(g_alpha, n_alpha, sizeof(double) * n);
        memcpy(g_eps, n_eps, sizeof(double) * n);
        memcpy(g_sig, n_sig, sizeof(double) * n);
        memcpy(g_mmass, n_mmass, sizeof(double) * n);
        memcpy(g_mol, n_mol, sizeof(int) * n);
        memcpy(g_frozen, n_frozen, sizeof(int) * n);
        memcpy(g_pos, n_pos, sizeof(double) * 3 * n);
        g_n = n;
        return;
    }
    first = -1;
    last = -1;
    for (i = 0; i < n; i++)
        if (n_pos[3 * i] != g_pos[3 * i] || n_pos[3 * i + 1] != g_pos[3 * i + 1] || n_pos[3 * i + 2] != g_pos[3 * i + 2]) {
            if (first < 0) first = i;
            last = i;
        }
    if (first >= 0) {
        if (mpmc_b200_set_positions(g_ctx, first, last - first + 1, n_pos + 3 * first)) shim_die("set_positions");
        memcpy(g_pos + 3 * first, n_pos + 3 * first, sizeof(double) * 3 * (last - first + 1));
    }
}

/* ------------------------------------------------------------------------------------------------
 * Fast path of the O(N) host duties.  The reference's lists are chains of individually allocated 912-byte
 * atoms: every walk is a chain of dependent cache misses (~5 ms per walk at 64k atoms), and energy() used to
 * make eight of them (count x2, rebuild_arrays, update_com, wrapall, pack, self points, write-back).  The
 * pointer sequence is cached after a full walk; later calls check that the lists still are that sequence
 * (molecule k -> k+1, first atom, atom i -> i+1: independent loads, which the core overlaps) and do count,
 * COM (pairs.c:364-385), wrapping (output.c:142-183), self points (coulombic.c:103-108) and packing in that one
 * pass.  Any relinking (uVT insert / remove / restore, replay) fails the check and takes the full walks.
 * ---------------------------------------------------------------------------------------------- */
static molecule_t **c_mol = NULL;   /* molecules in list order */
static atom_t **c_atom = NULL;      /* atoms in list order */
static int *c_first = NULL;         /* index of a molecule's first atom */
static int c_nmol = 0, c_natom = 0, c_molcap = 0, c_atomcap = 0, c_valid = 0;
static int *t_slot = NULL;          /* threaded pass: molecule slot of atom i */
static double *t_mass = NULL;       /* threaded pass: atom masses, contiguous */
static double *t_shift = NULL;      /* threaded pass: [nmol][3] wrapping shift of a molecule */
static int t_atomcap = 0, t_molcap = 0;

static void shim_cache_topology(system_t *system) {
    molecule_t *m;
    atom_t *a;
    int nm = 0, na = 0;
    for (m = system->molecules; m; m = m->next) nm++;
    na = system->natoms;
    if (nm + 1 > c_molcap) {
        c_molcap = nm + nm / 4 + 64;
        c_mol = (molecule_t **)xrealloc(c_mol, sizeof(molecule_t *) * c_molcap);
        c_first = (int *)xrealloc(c_first, sizeof(int) * c_molcap);
    }
    if (na > c_atomcap) {
        c_atomcap = na + na / 4 + 64;
        c_atom = (atom_t **)xrealloc(c_atom, sizeof(atom_t *) * c_atomcap);
    }
    nm = 0;
    na = 0;
    for (m = system->molecules; m; m = m->next) {
        c_mol[nm] = m;
        c_first[nm++] = na;
        for (a = m->atoms; a; a = a->next) c_atom[na++] = a;
    }
    c_first[nm] = na;
    c_nmol = nm;
    c_natom = na;
    c_valid = (na == system->natoms);
}

/* Keeping the cached sequence across the MC driver's relinking (mc_moves.c:567-741 make_move(), :744-804 restore()).  At slot k
 * the real list holds `fresh` where the cache expects c_mol[k] (NULL past the end).  Three single-node changes are recognised
 * and the cache is edited in place; anything else returns 0 and the caller falls back to the full walks:
 *   removal      fresh is the cached successor of slot k               (uVT remove; a rejected insertion)
 *   insertion    fresh->next is the node the cache has in slot k       (uVT insert; a rejected removal relinks the backup)
 *   replacement  fresh->next is the cached successor, same atom count  (a rejected displacement relinks the backup copy)
 * Several of them in one MC step (restore + new move) are met one after the other as the pass walks the list.  Everything
 * compared is a node reached from system->molecules through real `next` links, so a match always means "this node sits here".
 * *resized is set when the atom count changed (atom_array / molecule_array are then rebuilt at the end of the pass). */
static molecule_t *shim_cached(int k) { return k < c_nmol ? c_mol[k] : NULL; }

static int shim_resync(system_t *system, int k, molecule_t *fresh, int *resized) {
    atom_t *a;
    int i, cnt;
    if (k < c_nmol && fresh == shim_cached(k + 1)) { /* removal of slot k */
        const int b = c_first[k], e = c_first[k + 1];
        cnt = e - b;
        memmove(c_atom + b, c_atom + e, sizeof(atom_t *) * (size_t)(c_natom - e));
        memmove(c_mol + k, c_mol + k + 1, sizeof(molecule_t *) * (size_t)(c_nmol - k - 1));
        for (i = k + 1; i <= c_nmol; i++) c_first[i - 1] = c_first[i] - cnt;
        c_nmol--;
        c_natom -= cnt;
        *resized = 1;
        return 1;
    }
    if (!fresh) return 0;
    if (fresh->next == shim_cached(k)) { /* insertion before slot k */
        const int b = c_first[k];
        for (a = fresh->atoms, cnt = 0; a; a = a->next) cnt++;
        if (c_nmol + 2 > c_molcap) {
            c_molcap = c_nmol + c_nmol / 4 + 64;
            c_mol = (molecule_t **)xrealloc(c_mol, sizeof(molecule_t *) * c_molcap);
            c_first = (int *)xrealloc(c_first, sizeof(int) * c_molcap);
        }
        if (c_natom + cnt > c_atomcap) {
            c_atomcap = c_natom + cnt + c_natom / 4 + 64;
            c_atom = (atom_t **)xrealloc(c_atom, sizeof(atom_t *) * c_atomcap);
        }
        memmove(c_atom + b + cnt, c_atom + b, sizeof(atom_t *) * (size_t)(c_natom - b));
        memmove(c_mol + k + 1, c_mol + k, sizeof(molecule_t *) * (size_t)(c_nmol - k));
        for (i = c_nmol; i >= k; i--) c_first[i + 1] = c_first[i] + cnt;
        c_mol[k] = fresh;
        for (a = fresh->atoms, i = b; a; a = a->next, i++) c_atom[i] = a;
        c_nmol++;
        c_natom += cnt;
        shim_reserve(c_natom);
        *resized = 1;
        return 1;
    }
    if (k < c_nmol && fresh->next == shim_cached(k + 1)) { /* replacement of slot k by a node with as many atoms */
        const int b = c_first[k], e = c_first[k + 1];
        for (a = fresh->atoms, i = b; a && i < e; a = a->next) i++;
        if (a || i != e) return 0;
        c_mol[k] = fresh;
        for (a = fresh->atoms, i = b; a; a = a->next, i++) {
            c_atom[i] = a;
            if (!*resized) {                       /* both arrays are per ATOM: molecule_array[i] is the molecule of atom i */
                system->atom_array[i] = a;         /* (pairs.c:28-34); after a resize they are rebuilt as a whole */
                system->molecule_array[i] = fresh;
            }
        }
        return 1;
    }
    return 0;
}

/* The one pass with host threads.  Taken when the cached sequence IS the list, molecule by molecule (checked first, in
 * order, stopping at the first relinked node, so nothing that make_move() / restore() freed is ever read); any
 * relinking falls through to the sequential pass below, which edits the cache.  Three sweeps:
 *   atoms      verify the atom links, copy what the hot path reads into the flat arrays (912-byte stride -> contiguous)
 *   molecules  mass, centre of mass (pairs.c:364-385) and wrapping shift (output.c:142-183) from the contiguous copies,
 *              summed in list order exactly as the reference does
 *   atoms      wrapped_pos, molecule mass / index per atom, point self energies (coulombic.c:103-108)
 * Same results as the sequential pass bit for bit (every sum keeps its order); returns the atom count or -1. */
static int shim_threaded_pass(system_t *system, int es_on) {
    const pbc_t *pbc = system->pbc;
    const double self_c = system->ewald_alpha / sqrt(M_PI);
    const int nthreads = shim_threads();
    int k, i, bad = 0;
    molecule_t *real = system->molecules;
    if (nthreads <= 1 || !c_valid || c_nmol <= 0 || c_natom < g_threads_min || !system->atom_array || !system->molecule_array) return -1;
    for (k = 0; k < c_nmol; k++) { /* the addresses come from the cache, so the loads overlap; the order keeps it safe */
        if (real != c_mol[k]) return -1;
        real = c_mol[k]->next;
    }
    if (real) return -1;
    shim_reserve(c_natom);
    if (c_natom > t_atomcap) {
        t_atomcap = c_natom + c_natom / 4 + 64;
        t_slot = (int *)xrealloc(t_slot, sizeof(int) * t_atomcap);
        t_mass = (double *)xrealloc(t_mass, sizeof(double) * t_atomcap);
    }
    if (c_nmol > t_molcap) {
        t_molcap = c_nmol + c_nmol / 4 + 64;
        t_shift = (double *)xrealloc(t_shift, sizeof(double) * 3 * t_molcap);
    }
#pragma omp parallel for schedule(static) num_threads(nthreads)
    for (k = 0; k < c_nmol; k++) {
        int j;
        if (c_mol[k]->atoms != (c_first[k + 1] > c_first[k] ? c_atom[c_first[k]] : NULL)) {
#pragma omp atomic write
            bad = 1;
        }
        for (j = c_first[k]; j < c_first[k + 1]; j++) t_slot[j] = k;
    }
    if (bad) return -1;
#pragma omp parallel for schedule(static) num_threads(nthreads)
    for (i = 0; i < c_natom; i++) {
        const atom_t *a = c_atom[i];
        const int last = (i + 1 == c_first[t_slot[i] + 1]);
        if (a->next != (last ? NULL : c_atom[i + 1])) {
#pragma omp atomic write
            bad = 1;
        }
        n_pos[3 * i] = a->pos[0];
        n_pos[3 * i + 1] = a->pos[1];
        n_pos[3 * i + 2] = a->pos[2];
        n_q[i] = a->charge;
        n_alpha[i] = a->polarizability;
        n_eps[i] = a->epsilon;
        n_sig[i] = a->sigma;
        n_frozen[i] = a->frozen;
        n_mol[i] = t_slot[i];
        t_mass[i] = a->mass;
    }
    if (bad) return -1;
#pragma omp parallel for schedule(dynamic, 64) num_threads(nthreads)
    for (k = 0; k < c_nmol; k++) {
        molecule_t *m = c_mol[k];
        const int b = c_first[k], e = c_first[k + 1];
        double com[3] = {0, 0, 0}, mass = 0, dimg[3] = {0, 0, 0}, d[3];
        int j, p, q;
        for (j = b; j < e; j++) {
            mass += t_mass[j];
            for (p = 0; p < 3; p++) com[p] += t_mass[j] * n_pos[3 * j + p];
        }
        for (p = 0; p < 3; p++) m->com[p] = 0; /* update_com(): pairs.c:364-385 */
        if (!(m->spectre || m->target)) {
            m->mass = mass;
            for (p = 0; p < 3; p++) m->com[p] = com[p] / mass;
        }
        if (!m->frozen) { /* wrapall(): output.c:142-183 */
            for (p = 0; p < 3; p++) {
                for (q = 0, d[p] = 0; q < 3; q++) d[p] += pbc->reciprocal_basis[q][p] * m->com[q];
                d[p] = rint(d[p]);
            }
            for (p = 0; p < 3; p++) {
                for (q = 0, dimg[p] = 0; q < 3; q++) dimg[p] += pbc->basis[q][p] * d[q];
                m->wrapped_com[p] = dimg[p];
            }
        }
        for (p = 0; p < 3; p++) t_shift[3 * k + p] = dimg[p];
    }
#pragma omp parallel for schedule(static) num_threads(nthreads)
    for (i = 0; i < c_natom; i++) {
        atom_t *a = c_atom[i];
        const int ks = t_slot[i];
        const molecule_t *m = c_mol[ks];
        int p;
        for (p = 0; p < 3; p++) a->wrapped_pos[p] = m->frozen ? n_pos[3 * i + p] : n_pos[3 * i + p] - t_shift[3 * ks + p];
        n_mmass[i] = m->mass;
        if (es_on && !n_frozen[i]) a->es_self_point_energy = self_c * n_q[i] * n_q[i]; /* coulombic.c:103-108 */
    }
    system->natoms = c_natom;
    return c_natom;
}

/* returns the atom count when the cached sequence still is the list (everything done), -1 otherwise */
static int shim_fast_pass(system_t *system, int es_on) {
    const pbc_t *pbc = system->pbc;
    const double self_c = system->ewald_alpha / sqrt(M_PI);
    int k, i, p, q;
    molecule_t *real = system->molecules;
    int resized = 0;
    if (!c_valid || c_nmol <= 0 || !system->atom_array || !system->molecule_array) return -1;
    shim_reserve(c_natom);
    for (k = 0;; k++) {
        molecule_t *m;
        int b, e, tries = 0;
        double com[3] = {0, 0, 0}, mass = 0, dimg[3] = {0, 0, 0}, d[3];
        while (real != shim_cached(k)) /* the MC driver relinked something here: edit the cache or give up */
            if (++tries > 4 || !shim_resync(system, k, real, &resized)) return -1;
        if (k >= c_nmol) break; /* real == NULL: the end of both */
        m = c_mol[k];
        b = c_first[k];
        e = c_first[k + 1];
        real = m->next;
        if (m->atoms != (e > b ? c_atom[b] : NULL)) return -1;
        for (i = b; i < e; i++) {
            const atom_t *a = c_atom[i];
            if (a->next != (i + 1 < e ? c_atom[i + 1] : NULL)) return -1;
            mass += a->mass;
            for (p = 0; p < 3; p++) com[p] += a->mass * a->pos[p];
        }
        for (p = 0; p < 3; p++) m->com[p] = 0; /* update_com(): pairs.c:364-385 */
        if (!(m->spectre || m->target)) {
            m->mass = mass;
            for (p = 0; p < 3; p++) m->com[p] = com[p] / mass;
        }
        if (!m->frozen) { /* wrapall(): output.c:142-183 */
            for (p = 0; p < 3; p++) {
                for (q = 0, d[p] = 0; q < 3; q++) d[p] += pbc->reciprocal_basis[q][p] * m->com[q];
                d[p] = rint(d[p]);
            }
            for (p = 0; p < 3; p++) {
                for (q = 0, dimg[p] = 0; q < 3; q++) dimg[p] += pbc->basis[q][p] * d[q];
                m->wrapped_com[p] = dimg[p];
            }
        }
        for (i = b; i < e; i++) {
            atom_t *a = c_atom[i];
            for (p = 0; p < 3; p++) {
                a->wrapped_pos[p] = m->frozen ? a->pos[p] : a->pos[p] - dimg[p];
                n_pos[3 * i + p] = a->pos[p];
            }
            n_q[i] = a->charge;
            n_alpha[i] = a->polarizability;
            n_eps[i] = a->epsilon;
            n_sig[i] = a->sigma;
            n_mmass[i] = m->mass;
            n_mol[i] = k;
            n_frozen[i] = a->frozen;
            if (es_on && !a->frozen) a->es_self_point_energy = self_c * a->charge * a->charge; /* coulombic.c:103-108 */
        }
    }
    if (resized) { /* rebuild_arrays(): pairs.c:12-38, from the sequence just verified */
        free(system->atom_array);
        free(system->molecule_array);
        system->atom_array = (atom_t **)malloc((c_natom ? c_natom : 1) * sizeof(atom_t *));
        system->molecule_array = (molecule_t **)malloc((c_natom ? c_natom : 1) * sizeof(molecule_t *));
        if (!system->atom_array || !system->molecule_array) shim_die("out of memory");
        for (k = 0; k < c_nmol; k++)
            for (i = c_first[k]; i < c_first[k + 1]; i++) {
                system->atom_array[i] = c_atom[i];
                system->molecule_array[i] = c_mol[k];
            }
    }
    system->natoms = c_natom;
    return c_natom;
}

/* what polar() leaves in atom_t (read by write_dipole / write_field, src/io/output.c:1029-1088) */
static int g_view_step = -1; /* >= 0: the per-atom results of that step of the last resident batch (mpmc_b200_mc_steps) */
static void shim_write_back_atoms(system_t *system) {
    molecule_t *m;
    atom_t *a;
    int n = 0, p;
    const double *v_mu, *v_es, *v_ei, *v_eic, *v_rrms, *v_rank;
    const int ranked = system->polar_gs_ranked;
    if (g_view_step >= 0) {
        if (mpmc_b200_mc_step_arrays_view(g_ctx, g_view_step, &v_mu, &v_es, &v_ei, &v_eic, &v_rrms, &v_rank)) shim_die("mc_step_arrays_view");
    } else if (c_valid && c_natom == system->natoms && shim_threads() > 1 && c_natom >= g_threads_min) {
        /* straight from the library's pinned host image into atom_t (no staging copy) */
        if (mpmc_b200_atom_arrays_view(g_ctx, &v_mu, &v_es, &v_ei, &v_eic, &v_rrms, &v_rank)) shim_die("atom_arrays_view");
    } else {
        if (mpmc_b200_get_atom_arrays(g_ctx, w_mu, w_es, w_ei, w_eic, w_rrms, ranked ? w_rank : NULL)) shim_die("get_atom_arrays");
        v_mu = w_mu, v_es = w_es, v_ei = w_ei, v_eic = w_eic, v_rrms = w_rrms, v_rank = w_rank;
    }
    if (c_valid && c_natom == system->natoms) { /* the sequence was checked (or rebuilt, or repointed) by this call */
        const int nthreads = (c_natom >= g_threads_min) ? shim_threads() : 1;
        int i;
#pragma omp parallel for schedule(static) num_threads(nthreads) if (nthreads > 1)
        for (i = 0; i < c_natom; i++) {
            atom_t *t = c_atom[i];
            int pp;
            for (pp = 0; pp < 3; pp++) {
                t->mu[pp] = v_mu[3 * i + pp];
                t->ef_static[pp] = v_es[3 * i + pp];
                t->ef_induced[pp] = v_ei[3 * i + pp];
                t->ef_induced_change[pp] = v_eic[3 * i + pp];
            }
            t->dipole_rrms = v_rrms[i];
            if (ranked) t->rank_metric = v_rank[i];
        }
        return;
    }
    for (m = system->molecules; m; m = m->next)
        for (a = m->atoms; a; a = a->next, n++) {
            for (p = 0; p < 3; p++) {
                a->mu[p] = v_mu[3 * n + p];
                a->ef_static[p] = v_es[3 * n + p];
                a->ef_induced[p] = v_ei[3 * n + p];
                a->ef_induced_change[p] = v_eic[3 * n + p];
            }
            a->dipole_rrms = v_rrms[n];
            if (ranked) a->rank_metric = v_rank[n];
        }
}

static void shim_polar_epilogue(system_t *system, const mpmc_b200_energies *e) {
    char linebuf[MAXLINE];
    {
        const double t0 = (t_on == 1) ? shim_now() : 0.0;
        shim_write_back_atoms(system);
        if (t_on == 1) t_acc[3] += shim_now() - t0;
    }
    if (system->polar_ewald_full) { /* polar.c:54-57: ewald_full() reports a failed solve only (polar_ewald.c:391-394) */
        if (e->iter_success) system->iter_success = 1;
        return;
    }
    if (!system->polar_iterative) return;
    system->nodestats->polarization_iterations = (double)e->polar_iterations;
    system->observables->dipole_rrms = e->dipole_rrms;
    if (e->iter_success) system->iter_success = 1; /* thole_iterative.c:207; the MC driver resets it (mc.c:347) */
    if (system->iter_success) {                    /* polar.c:66-89 */
        switch (system->ensemble) {
            case ENSEMBLE_UVT:
            case ENSEMBLE_NVT:
            case ENSEMBLE_NVE:
            case ENSEMBLE_NPT:
                sprintf(linebuf, "POLAR: polarization iterative solver convergence failure on mc step %d.\n", system->step);
                break;
            case ENSEMBLE_REPLAY:
                sprintf(linebuf, "POLAR: polarization iterative solver convergence failure on configuration %d.\n", system->step);
                break;
            default:
                sprintf(linebuf, "POLAR: polarization iterative solver failed to reach convergence.\n");
        }
        error(linebuf);
    }
}

/* coulombic_self() also stores the per-atom point self energy (src/energy/coulombic.c:103-108) */
static void shim_es_self_points(system_t *system) {
    molecule_t *m;
    atom_t *a;
    for (m = system->molecules; m; m = m->next)
        for (a = m->atoms; a; a = a->next)
            if (!a->frozen) a->es_self_point_energy = system->ewald_alpha * a->charge * a->charge / sqrt(M_PI);
}

/* MPMC_B200_SHIM_CHECK=1 (tests): the cached sequence, atom_array and molecule_array against a real walk of the lists */
static void shim_check_cache(system_t *system, int n) {
    static int check = -1;
    molecule_t *m;
    atom_t *a;
    int i = 0, k = 0;
    if (check < 0) check = getenv("MPMC_B200_SHIM_CHECK") ? 1 : 0;
    if (!check) return;
    for (m = system->molecules; m; m = m->next, k++) {
        if (k >= c_nmol || c_mol[k] != m || c_first[k] != i) shim_die("SHIM_CHECK: molecule sequence");
        for (a = m->atoms; a; a = a->next, i++)
            if (i >= n || c_atom[i] != a || system->atom_array[i] != a || system->molecule_array[i] != m)
                shim_die("SHIM_CHECK: atom sequence / atom_array / molecule_array");
    }
    if (i != n || k != c_nmol || system->natoms != n) shim_die("SHIM_CHECK: counts");
}

/* host lists -> O(N) duties + packed image + device: the one-pass fast path when the lists are the cached
 * sequence, else the reference-shaped walks (which also refresh the cache) */
static void shim_prepare(system_t *system, int es_on) {
    static int fast = -1;
    double t0 = 0, t1;
    int n;
    if (fast < 0) fast = getenv("MPMC_B200_SHIM_SLOW") ? 0 : 1; /* MPMC_B200_SHIM_SLOW=1: always the reference-shaped walks */
    shim_timing_init();
    if (t_on) t0 = shim_now();
    n = fast ? shim_threaded_pass(system, es_on) : -1;
    if (n < 0 && fast) n = shim_fast_pass(system, es_on);
    if (n >= 0) t_fast++;
    if (n >= 0) shim_check_cache(system, n);
    if (n < 0) {
        system->natoms = countNatoms(system);
        pairs(system);
        n = shim_pack(system);
        shim_cache_topology(system);
        if (es_on) shim_es_self_points(system);
    }
    if (t_on) {
        t1 = shim_now();
        t_acc[0] += t1 - t0;
        t0 = t1;
    }
    shim_upload(system, n);
    if (t_on) t_acc[1] += shim_now() - t0;
}

/* ------------------------------------------------------------------------------------------------
 * entry points
 * ---------------------------------------------------------------------------------------------- */
double polar(system_t *system) {
    mpmc_b200_energies e;
    if (system->polarization && !system->polar_iterative && !system->polar_zodid && !system->polar_ewald_full)
        shim_unsupported("matrix-inversion polarization (polar_iterative off)");
    shim_prepare(system, 0);
    if (mpmc_b200_polar(g_ctx, &e)) shim_die("polar");
    shim_polar_epilogue(system, &e);
    return e.polar;
}

/* ------------------------------------------------------------------------------------------------
 * Resident MC steps (opt-in: MPMC_B200_RESIDENT_STEPS=B).  The reference's mc() (src/mc/mc.c:294-353)
 * stays in charge, unmodified: make_move() -> energy() -> Metropolis test -> checkpoint() / restore().  In an NVT run of
 * displacement moves (or between the volume moves of an NPT run) everything the NEXT steps will do follows from random numbers that do not depend on the
 * configuration: per step one Metropolis draw (mc.c:320), one draw that picks the molecule (checkpoint.c:126), six for
 * translate() and four for rotate() (mc_moves.c:380-385,415-418).  So when energy() is called for step s the shim reads
 * the next draws AHEAD of the driver (get_rand() below serves them to the driver later, in the same order, from the
 * reference's own generator), turns them into the moves of steps s+1 .. s+B-1, and has the library run all B steps on
 * the device -- move, energy, decision, restore -- without coming back to the host (mpmc_b200_mc_steps).  The calls of
 * the following steps are then answered from that batch, after checking that make_move() really produced the predicted
 * trial coordinates (bit for bit) and that the driver's previous decision was the device's; any difference (it takes a
 * last-digit disagreement between glibc's exp and the device's inside the Metropolis test) drops the rest of the batch
 * and the call is evaluated the normal way -- the device image is then corrected by the usual diffed upload.
 * What a served call still does on the host: the duties of the ONE moved molecule (centre of mass, wrapping, pointer
 * arrays), the scatter of that step's per-atom results, the observables epilogue.
 * ---------------------------------------------------------------------------------------------- */
double mpmc_ref_get_rand(system_t *system); /* get_rand() of src/mersenne/mersenne.cpp, renamed in its object (Makefile) */
static double *r_fifo = NULL;
static int r_head = 0, r_tail = 0, r_cap = 0;
double get_rand(system_t *system) {
    if (r_head < r_tail) return r_fifo[r_head++];
    return mpmc_ref_get_rand(system);
}
/* the k-th draw the driver will make from now on */
static double shim_peek_rand(system_t *system, int k) {
    if (r_head == r_tail) r_head = r_tail = 0;
    while (r_tail - r_head <= k) {
        if (r_tail >= r_cap) {
            if (r_head > 0) {
                memmove(r_fifo, r_fifo + r_head, sizeof(double) * (size_t)(r_tail - r_head));
                r_tail -= r_head;
                r_head = 0;
            }
            if (r_tail >= r_cap) {
                r_cap = 2 * r_cap + 1024;
                r_fifo = (double *)xrealloc(r_fifo, sizeof(double) * (size_t)r_cap);
            }
        }
        r_fifo[r_tail++] = mpmc_ref_get_rand(system);
    }
    return r_fifo[r_head + k];
}

static int s_batch = -1;      /* steps per batch; 0 = off */
static int s_n = 0, s_pos = 0; /* steps in the current batch, next one to serve */
static mpmc_b200_mc_move *s_move = NULL;
static mpmc_b200_mc_result *s_res = NULL;
static int s_k[MPMC_B200_MC_MAX_STEPS]; /* list position of each step's molecule */
static int *s_exch = NULL, s_nexch = 0, s_exchcap = 0; /* checkpoint()'s ptr_array_exchange as list positions */
static long s_batches = 0, s_served = 0, s_dropped = 0;

static void shim_resident_report(void) {
    fprintf(stderr, "MPMC_B200_RESIDENT_STEPS: %ld batches, %ld steps answered from a batch, %ld batches cut short\n", s_batches,
            s_served, s_dropped);
}

static int shim_resident_eligible(const system_t *system, int with_observables) {
    if (s_batch < 0) {
        const char *v = getenv("MPMC_B200_RESIDENT_STEPS");
        s_batch = v ? atoi(v) : 0;
        if (s_batch > MPMC_B200_MC_MAX_STEPS) s_batch = MPMC_B200_MC_MAX_STEPS;
        if (s_batch < 2) s_batch = 0;
        if (s_batch) {
            s_move = (mpmc_b200_mc_move *)xrealloc(NULL, sizeof(mpmc_b200_mc_move) * MPMC_B200_MC_MAX_STEPS);
            s_res = (mpmc_b200_mc_result *)xrealloc(NULL, sizeof(mpmc_b200_mc_result) * MPMC_B200_MC_MAX_STEPS);
            if (getenv("MPMC_B200_SHIM_TIMING")) atexit(shim_resident_report);
        }
    }
    if (!s_batch || !with_observables) return 0;
    /* NVT, and the displacement steps of NPT runs (a batch ends where the driver will draw a volume move) */
    if ((system->ensemble != ENSEMBLE_NVT && system->ensemble != ENSEMBLE_NPT) || system->step < 1 || !system->checkpoint) return 0;
    if (system->checkpoint->movetype != MOVETYPE_DISPLACE || !system->checkpoint->molecule_altered || !system->checkpoint->molecule_backup) return 0;
    if (system->quantum_rotation || system->spectre || system->gwp || system->rd_anharmonic || system->cavity_bias) return 0;
    if (system->parallel_tempering || system->simulated_annealing || system->cavity_autoreject_absolute || system->sg) return 0;
    if (system->polarization && !system->rd_only) {
        if (system->polar_ewald_full || !system->polar_iterative) return 0;
        if (!system->polar_zodid && system->polar_precision != 0.0) return 0; /* tolerance mode reads a scalar per pass */
    }
    if (mpmc_b200_group_size(g_ctx) > 1) return 0;
    return 1;
}

/* molecule at list position k according to the lists themselves (the cache entry of k may be a freed object) */
static molecule_t *shim_molecule_at(system_t *system, int k) { return k > 0 ? c_mol[k - 1]->next : system->molecules; }

/* make position k of the cached sequence point at `m` (restore() relinks the backup copy in place of the moved molecule,
 * mc_moves.c:790-799); 0 when the atom count does not fit */
static int shim_repoint(system_t *system, int k, molecule_t *m) {
    const int b = c_first[k], e = c_first[k + 1];
    atom_t *a;
    int i;
    if (!m) return 0;
    for (a = m->atoms, i = b; a && i < e; a = a->next) i++;
    if (a || i != e) return 0;
    c_mol[k] = m;
    for (a = m->atoms, i = b; a; a = a->next, i++) {
        c_atom[i] = a;
        system->atom_array[i] = a;
        system->molecule_array[i] = m;
    }
    return 1;
}

static int shim_same_positions(int k, const double *want) {
    int i, p;
    for (i = c_first[k]; i < c_first[k + 1]; i++)
        for (p = 0; p < 3; p++)
            if (c_atom[i]->pos[p] != want[3 * (i - c_first[k]) + p]) return 0;
    return 1;
}

/* the host duties of pairs() for ONE molecule: update_com() (pairs.c:364-385) and wrapall() (output.c:142-183), as the
 * one-pass forms above do them for all */
static void shim_molecule_duties(system_t *system, int k) {
    const pbc_t *pbc = system->pbc;
    molecule_t *m = c_mol[k];
    const int b = c_first[k], e = c_first[k + 1];
    double com[3] = {0, 0, 0}, mass = 0, dimg[3] = {0, 0, 0}, d[3];
    int i, p, q;
    for (i = b; i < e; i++) {
        const atom_t *a = c_atom[i];
        mass += a->mass;
        for (p = 0; p < 3; p++) com[p] += a->mass * a->pos[p];
    }
    for (p = 0; p < 3; p++) m->com[p] = 0;
    if (!(m->spectre || m->target)) {
        m->mass = mass;
        for (p = 0; p < 3; p++) m->com[p] = com[p] / mass;
    }
    if (!m->frozen) {
        for (p = 0; p < 3; p++) {
            for (q = 0, d[p] = 0; q < 3; q++) d[p] += pbc->reciprocal_basis[q][p] * m->com[q];
            d[p] = rint(d[p]);
        }
        for (p = 0; p < 3; p++) {
            for (q = 0, dimg[p] = 0; q < 3; q++) dimg[p] += pbc->basis[q][p] * d[q];
            m->wrapped_com[p] = dimg[p];
        }
    }
    for (i = b; i < e; i++) {
        atom_t *a = c_atom[i];
        for (p = 0; p < 3; p++) {
            a->wrapped_pos[p] = m->frozen ? a->pos[p] : a->pos[p] - dimg[p];
            n_pos[3 * i + p] = a->pos[p];
        }
    }
}

/* Answer this call from the batch; 0 = not possible (the rest of the batch is dropped). */
static int shim_resident_serve(system_t *system, mpmc_b200_energies *e) {
    const int j = s_pos, k = s_k[j], kp = s_k[j - 1];
    const mpmc_b200_mc_result *prev = &s_res[j - 1];
    molecule_t *alt = system->checkpoint->molecule_altered;
#define DROP(why)                                                                                    \
    do {                                                                                             \
        if (getenv("MPMC_B200_RESIDENT_DEBUG")) fprintf(stderr, "resident: step %d of a batch dropped: %s\n", j, why); \
        return 0;                                                                                    \
    } while (0)
    if (!c_valid || c_natom != system->natoms || !system->atom_array || !system->molecule_array) DROP("no cached sequence");
    /* the molecule of the previous step: the driver kept it (accept) or relinked its backup copy (reject) */
    if (!shim_repoint(system, kp, shim_molecule_at(system, kp))) DROP("previous molecule not where it was");
    if (k != kp && !shim_same_positions(kp, prev->accepted ? prev->new_pos : prev->old_pos)) DROP("the driver decided the previous step otherwise");
    /* this step's molecule and its trial coordinates */
    if (shim_molecule_at(system, k) != alt || !shim_repoint(system, k, alt)) DROP("another molecule was picked");
    if (!shim_same_positions(k, s_res[j].new_pos)) DROP("trial coordinates differ");
#undef DROP
    shim_molecule_duties(system, kp);
    shim_molecule_duties(system, k);
    shim_check_cache(system, c_natom);
    *e = s_res[j].e;
    g_view_step = j;
    s_pos++;
    s_served++;
    return 1;
}

/* Start a batch at this call: the move the driver just made is step 0, the following ones come from the draws ahead.
 * The host image and the device are up to date (shim_prepare).  0 = no batch (evaluate the normal way). */
static int shim_resident_start(system_t *system, mpmc_b200_energies *e) {
    const molecule_t *alt = system->checkpoint->molecule_altered;
    const atom_t *a;
    int B = s_batch, j, i, k0 = -1, kk, draw;
    const double N = system->observables->N;
    const long left = (long)system->numsteps - (long)system->step + 1;
    if (left < B) B = (int)left;
    if (B < 2 || !c_valid || c_natom != system->natoms) return 0;
    /* checkpoint()'s eligibility list (checkpoint.c:34-46) as list positions */
    if (c_nmol > s_exchcap) {
        s_exchcap = c_nmol + c_nmol / 4 + 64;
        s_exch = (int *)xrealloc(s_exch, sizeof(int) * (size_t)s_exchcap);
    }
    for (kk = 0, s_nexch = 0; kk < c_nmol; kk++) {
        if (c_mol[kk] == alt) k0 = kk;
        if (!(c_mol[kk]->frozen || c_mol[kk]->adiabatic || c_mol[kk]->target)) s_exch[s_nexch++] = kk;
        else if (c_mol[kk]->adiabatic) return 0;
    }
    if (k0 < 0 || (double)s_nexch != N) return 0;
    /* a batch keeps one write-back image per step on the device and in pinned memory: bound it */
    {
        static long ring_mb = -1; /* MPMC_B200_RESIDENT_RING_MB (default 64) */
        if (ring_mb < 0) ring_mb = getenv("MPMC_B200_RESIDENT_RING_MB") ? atol(getenv("MPMC_B200_RESIDENT_RING_MB")) : 64;
        if (ring_mb < 1) ring_mb = 1;
        if (system->polarization && (long)B * c_natom * 14 * 8 > (ring_mb << 20)) B = (int)((ring_mb << 20) / ((long)c_natom * 14 * 8));
    }
    if (B < 2) return 0;
    memset(s_move, 0, sizeof(mpmc_b200_mc_move) * (size_t)B);
    /* step 0: the trial coordinates are in the lists (and on the device); a reject restores the backup's */
    s_k[0] = k0;
    s_move[0].first = c_first[k0];
    s_move[0].count = c_first[k0 + 1] - c_first[k0];
    if (s_move[0].count < 1 || s_move[0].count > MPMC_B200_MC_MAX_ATOMS) return 0;
    s_move[0].explicit_positions = 1;
    for (i = 0, a = system->checkpoint->molecule_backup->atoms; i < s_move[0].count; i++, a = a->next) {
        int p;
        if (!a) return 0;
        for (p = 0; p < 3; p++) {
            s_move[0].new_pos[3 * i + p] = c_atom[c_first[k0] + i]->pos[p];
            s_move[0].old_pos[3 * i + p] = a->pos[p];
        }
    }
    if (a) return 0;
    s_move[0].u_accept = shim_peek_rand(system, 0);
    /* steps 1..B-1: per step the driver will draw u (mc.c:320), in NPT the volume-or-displace draw (checkpoint.c:83-97),
     * the pick (checkpoint.c:126), translate() x6, rotate() x4 */
    for (j = 1, draw = 1; j < B; j++, draw += 12) {
        struct quaternion q;
        mpmc_b200_mc_move *mv = &s_move[j];
        int altered;
        if (system->ensemble == ENSEMBLE_NPT) {
            const double pv = system->volume_probability == 0.0 ? 1.0 / N : system->volume_probability;
            if (shim_peek_rand(system, draw) < pv) break; /* a volume move comes next: the batch ends before it */
            draw++;
        }
        altered = (int)floor(shim_peek_rand(system, draw) * N);
        double t[3], sgn[3], ax[3], angle;
        int p;
        if (altered < 0 || altered >= s_nexch) break;
        kk = s_exch[altered];
        mv->first = c_first[kk];
        mv->count = c_first[kk + 1] - c_first[kk];
        if (mv->count < 1 || mv->count > MPMC_B200_MC_MAX_ATOMS) break;
        s_k[j] = kk;
        for (i = 0; i < mv->count; i++) mv->mass[i] = c_atom[mv->first + i]->mass;
        for (p = 0; p < 3; p++) t[p] = system->move_factor * shim_peek_rand(system, draw + 1 + p) * system->pbc->cutoff; /* mc_moves.c:380-382 */
        for (p = 0; p < 3; p++) sgn[p] = shim_peek_rand(system, draw + 4 + p);
        for (p = 0; p < 3; p++) {
            if (sgn[p] < 0.5) t[p] *= -1.0;
            mv->trans[p] = t[p];
        }
        for (p = 0; p < 3; p++) ax[p] = shim_peek_rand(system, draw + 7 + p) - 0.5; /* mc_moves.c:415-418 */
        angle = shim_peek_rand(system, draw + 10) * 360 * system->rot_factor;
        quaternion_construct_axis_angle_degree(&q, ax[0], ax[1], ax[2], angle);
        mv->quat[0] = q.x, mv->quat[1] = q.y, mv->quat[2] = q.z, mv->quat[3] = q.w;
        mv->u_accept = shim_peek_rand(system, draw + 11);
    }
    B = j;
    if (B < 2) return 0;
    if (mpmc_b200_mc_steps(g_ctx, B, s_move, system->temperature, system->observables->energy, s_res)) {
        s_batch = 0; /* refused (options the batch form does not cover): stay on the normal path for good */
        return 0;
    }
    /* the device now holds the configuration after all B decisions: that is what later uploads are diffed against */
    for (j = 0; j < B; j++) {
        const double *fin = s_res[j].accepted ? s_res[j].new_pos : s_res[j].old_pos;
        memcpy(g_pos + 3 * (size_t)s_move[j].first, fin, sizeof(double) * 3 * (size_t)s_move[j].count);
    }
    s_n = B;
    s_pos = 1;
    s_batches++;
    s_served++;
    *e = s_res[0].e;
    g_view_step = 0;
    return 1;
}

static double shim_energy(system_t *system, int with_observables) {
    mpmc_b200_energies e;
    double potential_energy = 0, rd_energy = 0, coulombic_energy = 0, polar_energy = 0;
    const int es_on = !(system->sg || system->rd_only);
    int have = 0;

    if (es_on && system->polarization && !system->polar_iterative && !system->polar_zodid && !system->polar_ewald_full)
        shim_unsupported("matrix-inversion polarization (polar_iterative off)");
    shim_timing_init();
    g_view_step = -1;
    {
        const double tw = (t_on == 1) ? shim_now() : 0.0;
        double te;
        const int resident = g_ctx ? shim_resident_eligible(system, with_observables) : 0;
        if (s_pos < s_n) { /* a batch is open: answer from it, or drop what is left of it */
            if (resident) have = shim_resident_serve(system, &e);
            if (!have) {
                s_n = s_pos = 0;
                s_dropped++;
            }
        }
        if (!have) {
            shim_prepare(system, es_on);
            te = (t_on == 1) ? shim_now() : 0.0;
            if (g_ctx && shim_resident_eligible(system, with_observables)) have = shim_resident_start(system, &e);
            if (!have && mpmc_b200_energy(g_ctx, &e)) shim_die("energy");
            if (t_on == 1) t_acc[2] += shim_now() - te;
        }
        if (t_on == 1) {
            t_calls++;
            t_acc[4] -= tw; /* closed at the end of this function */
        }
    }

    if (e.autorejected) { /* cavity_autoreject_absolute found a bad contact: energy.c:99-102,187-191 */
        potential_energy = MAXVALUE;
        if (with_observables) system->count_autorejects++;
    } else {
        if (es_on && system->polarization) {
            polar_energy = e.polar;
            if (with_observables) system->observables->polarization_energy = polar_energy;
            shim_polar_epilogue(system, &e);
        }
        rd_energy = e.rd;
        if (with_observables) system->observables->rd_energy = rd_energy;
        if (es_on) {
            coulombic_energy = e.coulombic;
            if (with_observables) system->observables->coulombic_energy = coulombic_energy;
        }
    }
    g_view_step = -1;
    potential_energy += rd_energy + coulombic_energy + polar_energy; /* + vdw + three-body: unsupported => 0 */
    if (t_on == 1) {
        t_last_call = shim_now();
        t_acc[4] += t_last_call;
        if (t_calls == 1) t_first_call = t_last_call;
    }
    if (!with_observables) return potential_energy;

    system->observables->energy = potential_energy;
    countN(system);
    system->observables->spin_ratio /= system->observables->N;
    if (system->ensemble == ENSEMBLE_NVE) {
        system->observables->kinetic_energy = system->total_energy - potential_energy;
        system->observables->temperature = (2.0 / 3.0) * system->observables->kinetic_energy / system->observables->N;
    }
    system->observables->NU = system->observables->N * system->observables->energy;
    system->last_volume = system->pbc->volume;
    return potential_energy;
}

double energy(system_t *system) { return shim_energy(system, 1); }
double energy_no_observables(system_t *system) { return shim_energy(system, 0); }
