/* tests/hoststub/abi_over_oracle.c -- TEST INFRASTRUCTURE ONLY, never part of the product.
 *
 * The C-ABI entry points the drop-in shim calls (include/mpmc_b200.h), answered by the CPU oracle
 * (oracle/mpmc_oracle.c) instead of the CUDA engine.  Linked with the shim's own objects into
 * tests/hoststub/_build/mpmc_hoststub so that the HOST logic of mpmc_b200/shim/energy_shim.c -- the list walks and the
 * cached-sequence pass, change detection, uVT relinking, the observables epilogue, the write-back into atom_t -- is
 * exercised by the CPU test suite against the reference's fixed-seed trajectories (tests/test_shim_host_cpu.py).
 * libmpmc_b200.so itself has no CPU path: mpmc_b200_create() fails without a CUDA device (tests/test_abi_cpu.py). */
#include <stdlib.h>
#include <string.h>

#include "mpmc_b200.h"

int mpmc_oracle_energy(const mpmc_b200_params *P, int n, const double *pos, const double *charge, const double *polarizability,
                       const double *epsilon, const double *sigma, const double *mol_mass, const int *mol_id,
                       const int *frozen, mpmc_b200_energies *out, mpmc_b200_counters *cnt, double *mu, double *ef_static,
                       double *ef_induced, double *ef_induced_change, double *dipole_rrms, double *rank_metric, int nthreads);
int mpmc_oracle_polar(const mpmc_b200_params *P, int n, const double *pos, const double *charge, const double *polarizability,
                      const double *epsilon, const double *sigma, const double *mol_mass, const int *mol_id, const int *frozen,
                      mpmc_b200_energies *out, mpmc_b200_counters *cnt, double *mu, double *ef_static, double *ef_induced,
                      double *ef_induced_change, double *dipole_rrms, double *rank_metric, int nthreads);

struct mpmc_b200_ctx {
    mpmc_b200_params P;
    int n, cap;
    double *pos, *q, *alpha, *eps, *sig, *mm;
    int *mol, *frozen;
    double *mu, *es, *ei, *eic, *rrms, *rank;
};

static const char *g_err = "";
const char *mpmc_b200_last_error(void) { return g_err; }

int mpmc_b200_create(mpmc_b200_ctx **ctx, int device) {
    (void)device;
    *ctx = (mpmc_b200_ctx *)calloc(1, sizeof(mpmc_b200_ctx));
    return *ctx ? 0 : 11;
}

/* a "group" on the host stand-in is one context: what is under test is the shim's parsing of MPMC_B200_DEVICES */
static int g_group = 1;
int mpmc_b200_create_group(mpmc_b200_ctx **ctx, int ndev, const int *devices) {
    (void)devices;
    g_group = ndev;
    return mpmc_b200_create(ctx, 0);
}
int mpmc_b200_group_size(mpmc_b200_ctx *c) { return c ? g_group : 0; }

void mpmc_b200_destroy(mpmc_b200_ctx *c) {
    if (!c) return;
    free(c->pos), free(c->q), free(c->alpha), free(c->eps), free(c->sig), free(c->mm), free(c->mol), free(c->frozen);
    free(c->mu), free(c->es), free(c->ei), free(c->eic), free(c->rrms), free(c->rank);
    free(c);
}

int mpmc_b200_set_params(mpmc_b200_ctx *c, const mpmc_b200_params *p) {
    c->P = *p;
    return 0;
}

#define GROW(field, type, count) c->field = (type *)realloc(c->field, sizeof(type) * (size_t)(count))
int mpmc_b200_set_atoms(mpmc_b200_ctx *c, int n, const double *pos, const double *charge, const double *polarizability,
                        const double *epsilon, const double *sigma, const double *mol_mass, const int *mol_id,
                        const int *frozen) {
    if (n > c->cap) {
        c->cap = n + n / 4 + 16;
        GROW(pos, double, 3 * c->cap), GROW(q, double, c->cap), GROW(alpha, double, c->cap), GROW(eps, double, c->cap);
        GROW(sig, double, c->cap), GROW(mm, double, c->cap), GROW(mol, int, c->cap), GROW(frozen, int, c->cap);
        GROW(mu, double, 3 * c->cap), GROW(es, double, 3 * c->cap), GROW(ei, double, 3 * c->cap), GROW(eic, double, 3 * c->cap);
        GROW(rrms, double, c->cap), GROW(rank, double, c->cap);
    }
    c->n = n;
    memcpy(c->pos, pos, sizeof(double) * 3 * n), memcpy(c->q, charge, sizeof(double) * n);
    memcpy(c->alpha, polarizability, sizeof(double) * n), memcpy(c->eps, epsilon, sizeof(double) * n);
    memcpy(c->sig, sigma, sizeof(double) * n), memcpy(c->mm, mol_mass, sizeof(double) * n);
    memcpy(c->mol, mol_id, sizeof(int) * n), memcpy(c->frozen, frozen, sizeof(int) * n);
    memset(c->mu, 0, sizeof(double) * 3 * n), memset(c->es, 0, sizeof(double) * 3 * n), memset(c->ei, 0, sizeof(double) * 3 * n);
    memset(c->eic, 0, sizeof(double) * 3 * n), memset(c->rrms, 0, sizeof(double) * n), memset(c->rank, 0, sizeof(double) * n);
    return 0;
}

int mpmc_b200_set_positions(mpmc_b200_ctx *c, int first, int count, const double *pos) {
    if (first < 0 || count < 0 || first + count > c->n) {
        g_err = "position range out of bounds";
        return 1;
    }
    memcpy(c->pos + 3 * first, pos, sizeof(double) * 3 * count);
    return 0;
}

int mpmc_b200_energy(mpmc_b200_ctx *c, mpmc_b200_energies *out) {
    mpmc_b200_counters cnt;
    int rc = mpmc_oracle_energy(&c->P, c->n, c->pos, c->q, c->alpha, c->eps, c->sig, c->mm, c->mol, c->frozen, out, &cnt, c->mu,
                                c->es, c->ei, c->eic, c->rrms, c->rank, c->n < 4000 ? 1 : 0 /* threads: small systems run serially */);
    if (rc) g_err = "oracle refused the parameters";
    return rc;
}

int mpmc_b200_polar(mpmc_b200_ctx *c, mpmc_b200_energies *out) {
    mpmc_b200_counters cnt;
    int rc = mpmc_oracle_polar(&c->P, c->n, c->pos, c->q, c->alpha, c->eps, c->sig, c->mm, c->mol, c->frozen, out, &cnt, c->mu,
                               c->es, c->ei, c->eic, c->rrms, c->rank, c->n < 4000 ? 1 : 0 /* threads: small systems run serially */);
    if (rc) g_err = "oracle refused the parameters";
    return rc;
}

int mpmc_b200_get_atom_arrays(mpmc_b200_ctx *c, double *mu, double *ef_static, double *ef_induced, double *ef_induced_change,
                              double *dipole_rrms, double *rank_metric) {
    const size_t n = (size_t)c->n;
    if (mu) memcpy(mu, c->mu, sizeof(double) * 3 * n);
    if (ef_static) memcpy(ef_static, c->es, sizeof(double) * 3 * n);
    if (ef_induced) memcpy(ef_induced, c->ei, sizeof(double) * 3 * n);
    if (ef_induced_change) memcpy(ef_induced_change, c->eic, sizeof(double) * 3 * n);
    if (dipole_rrms) memcpy(dipole_rrms, c->rrms, sizeof(double) * n);
    if (rank_metric) memcpy(rank_metric, c->rank, sizeof(double) * n);
    return 0;
}

int mpmc_b200_atom_arrays_view(mpmc_b200_ctx *c, const double **mu, const double **ef_static, const double **ef_induced,
                               const double **ef_induced_change, const double **dipole_rrms, const double **rank_metric) {
    if (mu) *mu = c->mu;
    if (ef_static) *ef_static = c->es;
    if (ef_induced) *ef_induced = c->ei;
    if (ef_induced_change) *ef_induced_change = c->eic;
    if (dipole_rrms) *dipole_rrms = c->rrms;
    if (rank_metric) *rank_metric = c->rank;
    return 0;
}
