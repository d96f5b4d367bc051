/*
 * writers.c -- SURVEY 8f f4, writer half: the per-corrtime writers of the MC driver for large N.
 *
 * The reference formats every field of every atom with its own fprintf (src/io/output.c:315-564 write_molecules,
 * :674-852 write_states: ~20 calls per atom, 1.3 million per file at 64k atoms), and the driver calls both plus
 * write_dipole / write_field (:1029-1088) at every corrtime (src/mc/mc.c:385-412).  Here the lists are walked once
 * into an array, the atom lines are formatted by the shim's host threads -- one snprintf per atom into per-thread
 * buffers -- and each file is written with a few fwrite calls.  The conversions are printf's own with the reference's
 * format strings, so the files are byte for byte the reference's (tests/test_shim_host_cpu.py compares all four with
 * the unmodified reference's, the fixed-seed trajectory tests the final configuration).
 *
 * Link: the reference's output.o is compiled unmodified; mpmc_b200/shim/Makefile weakens these four symbols in it so
 * that the definitions below win (also for write_molecules_wrapper's call inside output.o), and keeps a second copy
 * of the object with the four renamed mpmc_ref_write_* : MPMC_B200_WRITERS=reference routes to them (A/B timing, and a
 * run-time way back).  MPMC_B200_SHIM_TIMING=1 reports the time spent in here at exit.
 */
#include <mc.h>
#include <time.h>

int mpmc_ref_write_molecules(system_t *system, FILE *fp);
void mpmc_ref_write_states(system_t *system);
void mpmc_ref_write_dipole(system_t *system);
void mpmc_ref_write_field(system_t *system);
int mpmc_b200_shim_host_threads(int *min_atoms); /* energy_shim.c: MPMC_B200_HOST_THREADS, MPMC_B200_HOST_THREADS_MIN */
/* output.c:566-672 (not in function_prototypes.h) */
FILE *open_traj_file(system_t *system);
FILE *open_dipole_file(system_t *system);
FILE *open_field_file(system_t *system);

static double w_now(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

static int w_mode = -1; /* 0 = the writers of this file, 1 = the reference's */
static int w_timing = 0;
static double w_acc[4];
static long w_calls[4];
static const char *const w_name[4] = {"write_molecules", "write_states", "write_dipole", "write_field"};

static void w_report(void) {
    int k;
    for (k = 0; k < 4; k++)
        if (w_calls[k])
            fprintf(stderr, "MPMC_B200_SHIM_TIMING: %s (%s): %ld calls, %.3f ms per call\n", w_name[k],
                    w_mode ? "reference" : "buffered, host threads", w_calls[k], 1e3 * w_acc[k] / (double)w_calls[k]);
}

static void w_init(void) {
    const char *v;
    if (w_mode >= 0) return;
    v = getenv("MPMC_B200_WRITERS");
    w_mode = (v && !strcmp(v, "reference")) ? 1 : 0;
    w_timing = getenv("MPMC_B200_SHIM_TIMING") ? 1 : 0;
    if (w_timing) atexit(w_report);
}

/* the lists as arrays: atom, its molecule, its 1-based molecule number */
static atom_t **l_atom = NULL;
static molecule_t **l_mol = NULL;
static int *l_molno = NULL, *l_first = NULL;
static int l_cap = 0, l_mcap = 0, l_natom = 0, l_nmol = 0;

static void w_die(const char *what) {
    fprintf(stderr, "mpmc_b200 writers: %s\n", what);
    exit(1);
}

static void w_collect(system_t *system) {
    molecule_t *m;
    atom_t *a;
    int i = 0, j = 0;
    for (m = system->molecules; m; m = m->next, j++) {
        if (j + 1 >= l_mcap) {
            l_mcap = 2 * l_mcap + 1024;
            l_first = (int *)realloc(l_first, (size_t)l_mcap * sizeof(int));
            if (!l_first) w_die("out of memory");
        }
        l_first[j] = i;
        for (a = m->atoms; a; a = a->next, i++) {
            if (i >= l_cap) {
                l_cap = 2 * l_cap + 4096;
                l_atom = (atom_t **)realloc(l_atom, (size_t)l_cap * sizeof(atom_t *));
                l_mol = (molecule_t **)realloc(l_mol, (size_t)l_cap * sizeof(molecule_t *));
                l_molno = (int *)realloc(l_molno, (size_t)l_cap * sizeof(int));
                if (!l_atom || !l_mol || !l_molno) w_die("out of memory");
            }
            l_atom[i] = a;
            l_mol[i] = m;
            l_molno[i] = j + 1;
        }
    }
    if (!l_first) {
        l_mcap = 1024;
        l_first = (int *)malloc((size_t)l_mcap * sizeof(int));
    }
    l_first[j] = i;
    l_natom = i;
    l_nmol = j;
}

/* per-thread text buffers */
#define W_MAX_THREADS 64
#define W_LINE_MAX 640 /* an atom line is ~190 characters; %-4.45s and wide numbers can stretch it */
static char *t_buf[W_MAX_THREADS];
static size_t t_cap[W_MAX_THREADS], t_len[W_MAX_THREADS];

static char *w_room(int t, size_t need) {
    if (t_len[t] + need > t_cap[t]) {
        t_cap[t] = 2 * (t_len[t] + need) + (1 << 16);
        t_buf[t] = (char *)realloc(t_buf[t], t_cap[t]);
        if (!t_buf[t]) w_die("out of memory");
    }
    return t_buf[t] + t_len[t];
}

static const char *w_flag(const atom_t *a) { /* output.c:361-380, 770-789 */
    return a->adiabatic ? "A" : a->frozen ? "F" : a->spectre ? "S" : a->target ? "T" : "M";
}

static int w_threads(int units) {
    int min_atoms = 0;
    int t = mpmc_b200_shim_host_threads(&min_atoms);
    if (t > W_MAX_THREADS) t = W_MAX_THREADS;
    if (units < min_atoms) t = 1;
    return t < 1 ? 1 : t;
}

/* kind 0: write_molecules (output.c:352-430), kind 1: write_states (:761-826) */
static void w_atom_lines(system_t *system, int kind, int ext_output, int nthreads) {
    int t;
    const int wrap = system->wrapall, indep = system->independent_particle;
#pragma omp parallel for schedule(static, 1) num_threads(nthreads)
    for (t = 0; t < nthreads; t++) {
        const int a0 = (int)((long long)l_natom * t / nthreads), a1 = (int)((long long)l_natom * (t + 1) / nthreads);
        int i;
        t_len[t] = 0;
        for (i = a0; i < a1; i++) {
            const atom_t *a = l_atom[i];
            const double *r = (kind == 1 || wrap) ? a->wrapped_pos : a->pos;
            char *o = w_room(t, W_LINE_MAX);
            int n;
#define W_ADD(...)                                                                  \
    do {                                                                            \
        n += snprintf(o + n, (size_t)(W_LINE_MAX - n), __VA_ARGS__);                \
        if (n >= W_LINE_MAX) w_die("atom line longer than the line buffer");        \
    } while (0)
            n = 0;
            W_ADD("ATOM  %5d %-4.45s %-3.3s %-1.1s", i + 1, a->atomtype, l_mol[i]->moleculetype, w_flag(a));
            if (kind == 0)
                W_ADD(" %4d   ", indep ? i + 1 : l_molno[i]);
            else
                W_ADD("%4d    ", l_molno[i]);
            if (ext_output == 0)
                W_ADD("%8.3f%8.3f%8.3f", r[0], r[1], r[2]);
            else
                W_ADD("%11.6f %11.6f %11.6f ", r[0], r[1], r[2]);
            if (kind == 0)
                W_ADD(" %8.5f %8.5f", a->mass, a->charge / E2REDUCED);
            else
                W_ADD(" %8.4f %8.4f", a->mass, a->charge / E2REDUCED);
            W_ADD(" %8.5f %8.5f %8.5f %8.5f %8.5f %8.5f %8.5f %8.5f %8.5f\n", a->polarizability, a->epsilon, a->sigma, a->omega,
                  a->gwp_alpha, a->c6, a->c8, a->c10, a->c9);
#undef W_ADD
            t_len[t] += (size_t)n;
        }
    }
}

static void w_flush_threads(FILE *fp, int nthreads) {
    int t;
    for (t = 0; t < nthreads; t++)
        if (t_len[t] && fwrite(t_buf[t], 1, t_len[t], fp) != t_len[t]) w_die("short write");
}

static void w_cryst1(FILE *fp, const pbc_t *pbc, int traj) {
    const double a = sqrt(dddotprod((double *)pbc->basis[0], (double *)pbc->basis[0]));
    const double b = sqrt(dddotprod((double *)pbc->basis[1], (double *)pbc->basis[1]));
    const double c = sqrt(dddotprod((double *)pbc->basis[2], (double *)pbc->basis[2]));
    const double g20 = 180.0 / M_PI * acos(dddotprod((double *)pbc->basis[2], (double *)pbc->basis[0]) /
                                           sqrt(dddotprod((double *)pbc->basis[0], (double *)pbc->basis[0]) * dddotprod((double *)pbc->basis[2], (double *)pbc->basis[2])));
    const double g12 = 180.0 / M_PI * acos(dddotprod((double *)pbc->basis[1], (double *)pbc->basis[2]) /
                                           sqrt(dddotprod((double *)pbc->basis[1], (double *)pbc->basis[1]) * dddotprod((double *)pbc->basis[2], (double *)pbc->basis[2])));
    const double g01 = 180.0 / M_PI * acos(dddotprod((double *)pbc->basis[0], (double *)pbc->basis[1]) /
                                           sqrt(dddotprod((double *)pbc->basis[1], (double *)pbc->basis[1]) * dddotprod((double *)pbc->basis[0], (double *)pbc->basis[0])));
    /* the restart file swaps the first two angles (VMD's convention, output.c:337-345); the trajectory does not (:740-745) */
    if (traj)
        fprintf(fp, "CRYST1%9.3f%9.3f%9.3f%7.2f%7.2f%7.2f\n", a, b, c, g12, g20, g01);
    else
        fprintf(fp, "CRYST1%9.3f%9.3f%9.3f%7.2f%7.2f%7.2f\n", a, b, c, g20, g12, g01);
}

static void w_basis_remarks(FILE *fp, const pbc_t *pbc) {
    int q;
    for (q = 0; q < 3; q++)
        fprintf(fp, "REMARK BOX BASIS[%d] = %20.14lf %20.14lf %20.14lf\n", q, pbc->basis[q][0], pbc->basis[q][1], pbc->basis[q][2]);
}

static int w_any_basis_at_least(const pbc_t *pbc, double lim) {
    int p, q;
    for (p = 0; p < 3; p++)
        for (q = 0; q < 3; q++)
            if (pbc->basis[p][q] >= lim) return 1;
    return 0;
}

int write_molecules(system_t *system, FILE *fp) {
    double t0;
    int ext_output, nthreads;
    w_init();
    t0 = w_timing ? w_now() : 0.0;
    if (w_mode) {
        const int rv = mpmc_ref_write_molecules(system, fp);
        if (w_timing) w_acc[0] += w_now() - t0, w_calls[0]++;
        return rv;
    }
    ext_output = system->long_output || w_any_basis_at_least(system->pbc, 100.0); /* output.c:329-337 */
    w_cryst1(fp, system->pbc, 0);
    w_collect(system);
    nthreads = w_threads(l_natom);
    w_atom_lines(system, 0, ext_output, nthreads);
    w_flush_threads(fp, nthreads);

    if (system->wrapall) { /* the box corners as virtual particles + their edges: output.c:432-523 */
        int i, j, k, l, m, n, p, q, box_labels[2][2][2];
        int atom_box = l_natom + 1;
        const int molecule_box = l_nmol + 1;
        double box_pos[3], box_occupancy[3];
        for (i = 0; i < 2; i++)
            for (j = 0; j < 2; j++)
                for (k = 0; k < 2; k++) {
                    fprintf(fp, "ATOM  %5d %-4.45s %-3.3s %-1.1s %4d   ", atom_box, "X", "BOX", "F", molecule_box);
                    box_occupancy[0] = ((double)i) - 0.5;
                    box_occupancy[1] = ((double)j) - 0.5;
                    box_occupancy[2] = ((double)k) - 0.5;
                    for (p = 0; p < 3; p++)
                        for (q = 0, box_pos[p] = 0; q < 3; q++) box_pos[p] += system->pbc->basis[q][p] * box_occupancy[q];
                    for (p = 0; p < 3; p++) fprintf(fp, ext_output == 0 ? "%8.3f" : "%11.6f ", box_pos[p]);
                    fprintf(fp, " %8.4f %8.4f %8.5f %8.5f %8.5f\n", 0.0, 0.0, 0.0, 0.0, 0.0);
                    box_labels[i][j][k] = atom_box;
                    ++atom_box;
                }
        for (i = 0; i < 2; i++)
            for (j = 0; j < 2; j++)
                for (k = 0; k < 2; k++)
                    for (l = 0; l < 2; l++)
                        for (m = 0; m < 2; m++)
                            for (n = 0; n < 2; n++)
                                if (abs(i - l) + abs(j - m) + abs(k - n) == 1)
                                    fprintf(fp, "CONECT %4d %4d\n", box_labels[i][j][k], box_labels[l][m][n]);
    }
    w_basis_remarks(fp, system->pbc);
    if (system->ensemble == ENSEMBLE_SURF_FIT) { /* output.c:537-547 */
        if (system->fit_boltzmann_weight)
            fprintf(fp, "REMARK SURFACE FITTING BOLTZMANN TEMPERATURE = %lf\n", system->fit_max_energy);
        else
            fprintf(fp, "REMARK SURFACE FITTING MAX_ENERGY = %lf\n", system->fit_max_energy);
        fprintf(fp, "REMARK SURFACE FITTING SQUARE_ERROR = %lf\n", system->fit_best_square_error);
    }
    fprintf(fp, "END\n");
    fflush(fp);
    if (w_timing) w_acc[0] += w_now() - t0, w_calls[0]++;
    return 0;
}

void write_states(system_t *system) {
    FILE *fp;
    double t0;
    int k, ext_output, nthreads;
    int num_frozen_molecules = 0, num_moveable_molecules = 0, num_frozen_atoms = 0, num_moveable_atoms = 0;
    w_init();
    t0 = w_timing ? w_now() : 0.0;
    if (w_mode) {
        mpmc_ref_write_states(system);
        if (w_timing) w_acc[1] += w_now() - t0, w_calls[1]++;
        return;
    }
    if (!strncmp("/dev/null", system->traj_output, 9)) return; /* output.c:690-692 */
    fp = open_traj_file(system);
    w_collect(system);
    for (k = 0; k < l_nmol; k++) {
        const int cnt = l_first[k + 1] - l_first[k];
        const molecule_t *m = cnt ? l_mol[l_first[k]] : NULL;
        if (!m) { /* a molecule without atoms: find it by walking (never the case in practice) */
            molecule_t *w = system->molecules;
            int q;
            for (q = 0; q < k; q++) w = w->next;
            m = w;
        }
        if (m->frozen) {
            ++num_frozen_molecules;
            num_frozen_atoms += cnt;
        } else {
            ++num_moveable_molecules;
            num_moveable_atoms += cnt;
        }
    }
    fprintf(fp, "REMARK step=%d\n", system->step);
    fprintf(fp, "REMARK total_molecules=%d, total_atoms=%d\n", num_frozen_molecules + num_moveable_molecules,
            num_frozen_atoms + num_moveable_atoms);
    fprintf(fp, "REMARK frozen_molecules=%d, moveable_molecules=%d\n", num_frozen_molecules, num_moveable_molecules);
    fprintf(fp, "REMARK frozen_atoms=%d, moveable_atoms=%d\n", num_frozen_atoms, num_moveable_atoms);
    ext_output = system->long_output || w_any_basis_at_least(system->pbc, 200.0); /* output.c:731-737 */
    w_cryst1(fp, system->pbc, 1);
    nthreads = w_threads(l_natom);
    w_atom_lines(system, 1, ext_output, nthreads);
    w_flush_threads(fp, nthreads);
    w_basis_remarks(fp, system->pbc);
    fprintf(fp, "ENDMDL\n");
    fflush(fp);
    fclose(fp);
    if (w_timing) w_acc[1] += w_now() - t0, w_calls[1]++;
}

/* kind 0: molecular dipoles in Debye (output.c:1029-1056), kind 1: molecular fields in e/A^2 (:1059-1088) */
static void w_molecule_lines(int kind, int nthreads) {
    int t;
#pragma omp parallel for schedule(static, 1) num_threads(nthreads)
    for (t = 0; t < nthreads; t++) {
        const int m0 = (int)((long long)l_nmol * t / nthreads), m1 = (int)((long long)l_nmol * (t + 1) / nthreads);
        int k, i, p;
        t_len[t] = 0;
        for (k = m0; k < m1; k++) {
            double v[3] = {0, 0, 0};
            char *o;
            int n;
            if (l_first[k + 1] == l_first[k] || l_mol[l_first[k]]->frozen) continue;
            for (i = l_first[k]; i < l_first[k + 1]; i++)
                for (p = 0; p < 3; p++) v[p] += kind == 0 ? l_atom[i]->mu[p] : l_atom[i]->ef_static[p] + l_atom[i]->ef_induced[p];
            o = w_room(t, W_LINE_MAX);
            if (kind == 0)
                n = snprintf(o, W_LINE_MAX, "%f %f %f\n", v[0] / DEBYE2SKA, v[1] / DEBYE2SKA, v[2] / DEBYE2SKA);
            else
                n = snprintf(o, W_LINE_MAX, "%f %f %f\n", v[0] / E2REDUCED, v[1] / E2REDUCED, v[2] / E2REDUCED);
            if (n >= W_LINE_MAX) w_die("line longer than the line buffer");
            t_len[t] += (size_t)n;
        }
    }
}

/* a molecule without atoms that is not frozen prints a line of zeros in the reference; keep that exact by routing
 * such lists (never produced by the parser) to the reference's writer */
static int w_has_empty_molecule(void) {
    int k;
    for (k = 0; k < l_nmol; k++)
        if (l_first[k + 1] == l_first[k]) return 1;
    return 0;
}

void write_dipole(system_t *system) {
    FILE *fp;
    double t0;
    int nthreads;
    w_init();
    t0 = w_timing ? w_now() : 0.0;
    if (!w_mode) w_collect(system);
    if (w_mode || w_has_empty_molecule()) {
        mpmc_ref_write_dipole(system);
        if (w_timing) w_acc[2] += w_now() - t0, w_calls[2]++;
        return;
    }
    if (!strncmp("/dev/null", system->dipole_output, 9)) return;
    fp = open_dipole_file(system);
    nthreads = w_threads(l_natom);
    w_molecule_lines(0, nthreads);
    w_flush_threads(fp, nthreads);
    fflush(fp);
    fclose(fp);
    if (w_timing) w_acc[2] += w_now() - t0, w_calls[2]++;
}

void write_field(system_t *system) {
    FILE *fp;
    double t0;
    int nthreads;
    w_init();
    t0 = w_timing ? w_now() : 0.0;
    if (!w_mode) w_collect(system);
    if (w_mode || w_has_empty_molecule()) {
        mpmc_ref_write_field(system);
        if (w_timing) w_acc[3] += w_now() - t0, w_calls[3]++;
        return;
    }
    if (!strncmp("/dev/null", system->field_output, 9)) return;
    fp = open_field_file(system);
    nthreads = w_threads(l_natom);
    w_molecule_lines(1, nthreads);
    w_flush_threads(fp, nthreads);
    fflush(fp);
    fclose(fp);
    if (w_timing) w_acc[3] += w_now() - t0, w_calls[3]++;
}
