/*
 * oracle/ref_driver.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Full-precision dump driver around the UNMODIFIED reference library
 * (oracle/_ref/libmpmc_ref.so, built from /root/reference by oracle/Makefile).
 * It does what the reference's main() does up to the first energy() call
 * (reference src/main/main.c:135-177), calls energy() once
 * (src/energy/energy.c:67) and prints every input the hot path reads and every
 * output it writes with %.17g, so that golden fixtures carry all 53 bits.
 *
 * usage: ref_driver <mpmc input file> [repeat [warmup [flags]]]
 *   repeat > 1 re-evaluates energy() that many times (after `warmup` untimed calls) and
 *   reports seconds per call (CPU baseline timing of the real reference).
 *   flags: a string of letters -- q: no per-atom lines in the snapshot (timing runs of large systems);
 *          n: call energy_no_observables() (src/energy/energy.c:232-318) instead of energy().
 * Output goes to stdout after the marker line "MPMC_SNAPSHOT 1"; the
 * reference's own chatter goes to stdout before it / stderr.
 */
#include <mc.h>

/* not in function_prototypes.h (declared locally in src/energy/coulombic.c:19) */
double coulombic_self(system_t *);

static double now_s(void) {
    struct timeval tv;
    gettimeofday(&tv, NULL);
    return tv.tv_sec + 1e-6 * tv.tv_usec;
}

static void *zalloc(size_t n) {
    void *p = calloc(1, n);
    if (!p) {
        fprintf(stderr, "ref_driver: out of memory\n");
        exit(2);
    }
    return p;
}

static void p3x3(const char *name, double m[3][3]) {
    printf("param %s %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g\n", name, m[0][0], m[0][1], m[0][2], m[1][0],
           m[1][1], m[1][2], m[2][0], m[2][1], m[2][2]);
}
#define PD(field) printf("param " #field " %.17g\n", (double)(s->field))
#define PI(field) printf("param " #field " %d\n", (int)(s->field))

int main(int argc, char **argv) {
    system_t *s;
    int repeat = 1, warmup = 0, i, n, molidx, quiet = 0, no_obs = 0;
    double e = 0, t0, t1;
    molecule_t *m;
    atom_t *a;

    if (argc < 2) {
        fprintf(stderr, "usage: %s <input> [repeat]\n", argv[0]);
        return 1;
    }
    if (argc > 2) repeat = atoi(argv[2]);
    if (argc > 3) warmup = atoi(argv[3]);
    if (argc > 4) {
        quiet = strchr(argv[4], 'q') != NULL;
        no_obs = strchr(argv[4], 'n') != NULL;
    }
    if (repeat < 1) repeat = 1;
    rank = 0;
    size = 1;

    s = setup_system(argv[1]);
    if (!s) {
        fprintf(stderr, "ref_driver: setup_system failed\n");
        return 1;
    }
    s->nodestats = zalloc(sizeof(nodestats_t));
    s->avg_nodestats = zalloc(sizeof(avg_nodestats_t));
    s->observables = zalloc(sizeof(observables_t));
    s->avg_observables = zalloc(sizeof(avg_observables_t));
    s->checkpoint = zalloc(sizeof(checkpoint_t));
    s->checkpoint->observables = zalloc(sizeof(observables_t));
    if (s->polarization && !s->polar_zodid) thole_resize_matrices(s);

    /* timing mode (repeat > 1 or warmup > 0): between calls the last molecule is displaced by
     * +-1e-3 A along x, as one MC translation would, so the reference's own pair cache
     * (src/energy/pairs.c:243-250) behaves as it does inside mc(); the snapshot printed below
     * is then of the LAST call.  Golden fixtures are generated with repeat = 1, warmup = 0. */
    t0 = now_s();
    for (i = 0; i < warmup + repeat; i++) {
        if (i == warmup) t0 = now_s();
        if (i) {
            for (m = s->molecules; m->next; m = m->next)
                ;
            for (a = m->atoms; a; a = a->next) a->pos[0] += (i & 1) ? 1e-3 : -1e-3;
        }
        e = no_obs ? energy_no_observables(s) : energy(s);
    }
    t1 = now_s();

    printf("MPMC_SNAPSHOT 1\n");
    printf("timing seconds_per_energy %.6g repeat %d\n", (t1 - t0) / repeat, repeat);
    p3x3("basis", s->pbc->basis);
    p3x3("reciprocal_basis", s->pbc->reciprocal_basis);
    printf("param cutoff %.17g\n", s->pbc->cutoff);
    printf("param volume %.17g\n", s->pbc->volume);
    PD(ewald_alpha);
    PI(ewald_kmax);
    PI(feynman_hibbs);
    PI(feynman_hibbs_order);
    PD(temperature);
    PI(rd_lrc);
    PI(waldmanhagler);
    PI(halgren_mixing);
    PI(c6_mixing);
    PI(rd_only);
    PI(cavity_autoreject_absolute);
    PD(cavity_autoreject_scale);
    PI(polarization);
    PI(polar_iterative);
    PI(damp_type);
    PD(polar_damp);
    PI(polar_max_iter);
    PD(polar_precision);
    PD(polar_gamma);
    PI(polar_gs);
    PI(polar_gs_ranked);
    PI(polar_sor);
    PI(polar_esor);
    PI(polar_palmo);
    PI(polar_rrms);
    PI(polar_zodid);
    PI(polar_ewald);
    PI(polar_ewald_full);
    PI(polar_wolf);
    PI(polar_wolf_full);
    PI(polar_wolf_alpha_lookup);
    PD(polar_wolf_alpha);
    PD(polar_wolf_alpha_lookup_cutoff);
    PD(polar_ewald_alpha);
    PI(wolf);
    PI(rd_crystal);
    PI(rd_crystal_order);
    PI(ensemble);

    printf("energy returned %.17g\n", e);
    printf("energy total %.17g\n", s->observables->energy);
    printf("energy rd %.17g\n", s->observables->rd_energy);
    printf("energy coulombic %.17g\n", s->observables->coulombic_energy);
    printf("energy polar %.17g\n", s->observables->polarization_energy);
    printf("energy dipole_rrms %.17g\n", s->observables->dipole_rrms);
    printf("energy N %.17g\n", s->observables->N);
    printf("energy polar_iterations %.17g\n", s->nodestats->polarization_iterations);
    printf("energy iter_success %d\n", s->iter_success);
    /* the three Ewald pieces separately (pure functions of the pair lists).  Not available when
     * this driver is linked against the B200 drop-in (mpmc_b200/shim, -DSHIM_DRIVER): there are
     * no pair lists there. */
#ifndef SHIM_DRIVER
    if (!s->rd_only && !s->wolf) {
        printf("energy es_recip %.17g\n", coulombic_reciprocal(s));
        printf("energy es_self %.17g\n", coulombic_self(s));
        printf("energy es_real %.17g\n", coulombic_real(s));
    } else if (!s->rd_only && s->wolf) { /* coulombic(): coulombic.c:27-28 */
        printf("energy es_recip 0\n");
        printf("energy es_self 0\n");
        printf("energy es_real %.17g\n", coulombic_wolf(s));
    }
#endif

    n = 0;
    for (m = s->molecules; m; m = m->next)
        for (a = m->atoms; a; a = a->next) n++;
    printf("natoms %d\n", n);
    molidx = 0;
    for (m = s->molecules; m && !quiet; m = m->next, molidx++) {
        for (a = m->atoms; a; a = a->next) {
            printf("atom %d %d", molidx, a->frozen);
            printf(" %.17g %.17g %.17g", a->pos[0], a->pos[1], a->pos[2]);
            printf(" %.17g %.17g %.17g %.17g %.17g %.17g", a->charge, a->polarizability, a->epsilon, a->sigma, a->mass,
                   m->mass);
            printf(" %.17g %.17g %.17g", a->ef_static[0], a->ef_static[1], a->ef_static[2]);
            printf(" %.17g %.17g %.17g", a->ef_induced[0], a->ef_induced[1], a->ef_induced[2]);
            printf(" %.17g %.17g %.17g", a->ef_induced_change[0], a->ef_induced_change[1], a->ef_induced_change[2]);
            printf(" %.17g %.17g %.17g", a->mu[0], a->mu[1], a->mu[2]);
            printf(" %.17g %.17g\n", a->dipole_rrms, a->rank_metric);
        }
    }
    printf("END\n");
    return 0;
}
