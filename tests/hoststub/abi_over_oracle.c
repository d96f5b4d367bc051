/* tests/hoststub/abi_over_oracle.c -- TEST INFRASTRUCTURE ONLY, never part of the product.
 *
 * The C-ABI entry points the drop-in shim calls (include/mpmc_b200.h), answered by the CPU oracle
 * (oracle/mpmc_oracle.c) instead of the CUDA engine.  Linked with the shim's own objects into
 * tests/hoststub/_build/mpmc_hoststub so that the HOST logic of mpmc_b200/shim/energy_shim.c -- the list walks and the
 * cached-sequence pass, change detection, uVT relinking, the observables epilogue, the write-back into atom_t -- is
 * exercised by the CPU test suite against the reference's fixed-seed trajectories (tests/test_shim_host_cpu.py).
 * libmpmc_b200.so itself has no CPU path: mpmc_b200_create() fails without a CUDA device (tests/test_abi_cpu.py). */
#include <math.h>
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
    double *ring; /* mpmc_b200_mc_steps: per-step copies of the six per-atom arrays, 14 n doubles each */
    int ring_steps, ring_n;
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
    free(c->mu), free(c->es), free(c->ei), free(c->eic), free(c->rrms), free(c->rank), free(c->ring);
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

/* Resident MC steps on the CPU stand-in: the plain-C restatement of what mc_step_kernel.cu does on the device --
 * translate() / rotate() (src/mc/mc_moves.c:376-464) about the centre of mass of update_com() (src/energy/pairs.c:364-385)
 * with quaternion_multiplication()'s operation order (src/main/quaternion.c:76-86), the oracle's energy, the NVT
 * Metropolis test (src/mc/mc.c:106-114,320), coordinates put back on a reject.  gcc on x86-64 without -mfma rounds every
 * operation separately, as the reference's own objects do. */
typedef struct {
    double x, y, z, w;
} stub_quat;
static stub_quat stub_qmul(stub_quat a, stub_quat b) {
    stub_quat r;
    r.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
    r.x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
    r.y = a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x;
    r.z = a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w;
    return r;
}

int mpmc_b200_mc_steps(mpmc_b200_ctx *c, int nsteps, const mpmc_b200_mc_move *moves, double temperature, double current_energy,
                       mpmc_b200_mc_result *results) {
    const size_t n = (size_t)c->n;
    int j, i, p;
    if (nsteps < 1 || nsteps > MPMC_B200_MC_MAX_STEPS) {
        g_err = "nsteps out of range";
        return 1;
    }
    c->ring = (double *)realloc(c->ring, sizeof(double) * 14 * n * (size_t)nsteps);
    c->ring_steps = nsteps;
    c->ring_n = c->n;
    for (j = 0; j < nsteps; j++) {
        const mpmc_b200_mc_move *m = &moves[j];
        mpmc_b200_mc_result *r = &results[j];
        double *pos = c->pos + 3 * (size_t)m->first;
        double *slot = c->ring + 14 * n * (size_t)j;
        int rc;
        memset(r, 0, sizeof(*r));
        if (m->explicit_positions) {
            memcpy(r->old_pos, m->old_pos, sizeof(double) * 3 * m->count);
            memcpy(r->new_pos, m->new_pos, sizeof(double) * 3 * m->count);
            memcpy(pos, m->new_pos, sizeof(double) * 3 * m->count);
        } else {
            double mass = 0, com[3] = {0, 0, 0};
            stub_quat q = {m->quat[0], m->quat[1], m->quat[2], m->quat[3]}, qc = {-m->quat[0], -m->quat[1], -m->quat[2], m->quat[3]};
            memcpy(r->old_pos, pos, sizeof(double) * 3 * m->count);
            for (i = 0; i < m->count; i++) {
                mass += m->mass[i];
                for (p = 0; p < 3; p++) com[p] += m->mass[i] * pos[3 * i + p];
            }
            for (p = 0; p < 3; p++) com[p] = com[p] / mass;
            for (p = 0; p < 3; p++) com[p] += m->trans[p];
            for (i = 0; i < m->count; i++)
                for (p = 0; p < 3; p++) pos[3 * i + p] += m->trans[p];
            for (i = 0; i < m->count; i++) {
                stub_quat pv = {pos[3 * i] - com[0], pos[3 * i + 1] - com[1], pos[3 * i + 2] - com[2], 0.}, a;
                a = stub_qmul(pv, qc);
                a = stub_qmul(q, a);
                pos[3 * i] = a.x + com[0];
                pos[3 * i + 1] = a.y + com[1];
                pos[3 * i + 2] = a.z + com[2];
            }
            memcpy(r->new_pos, pos, sizeof(double) * 3 * m->count);
        }
        rc = mpmc_b200_energy(c, &r->e);
        if (rc) return rc;
        memcpy(slot, c->mu, sizeof(double) * 3 * n), memcpy(slot + 3 * n, c->es, sizeof(double) * 3 * n);
        memcpy(slot + 6 * n, c->ei, sizeof(double) * 3 * n), memcpy(slot + 9 * n, c->eic, sizeof(double) * 3 * n);
        memcpy(slot + 12 * n, c->rrms, sizeof(double) * n), memcpy(slot + 13 * n, c->rank, sizeof(double) * n);
        {
            double bf = 0.0;
            if (isfinite(r->e.total)) bf = exp(-(r->e.total - current_energy) / temperature);
            r->accepted = (m->u_accept < bf) && r->e.iter_success == 0;
        }
        if (r->accepted)
            current_energy = r->e.total;
        else
            memcpy(pos, r->old_pos, sizeof(double) * 3 * m->count);
    }
    return 0;
}

int mpmc_b200_mc_step_arrays_view(mpmc_b200_ctx *c, int step, const double **mu, const double **ef_static, const double **ef_induced,
                                  const double **ef_induced_change, const double **dipole_rrms, const double **rank_metric) {
    const size_t n = (size_t)c->ring_n;
    const double *h;
    if (!c->ring || step < 0 || step >= c->ring_steps) {
        g_err = "no per-atom results of that step";
        return 2;
    }
    h = c->ring + 14 * n * (size_t)step;
    if (mu) *mu = h;
    if (ef_static) *ef_static = h + 3 * n;
    if (ef_induced) *ef_induced = h + 6 * n;
    if (ef_induced_change) *ef_induced_change = h + 9 * n;
    if (dipole_rrms) *dipole_rrms = h + 12 * n;
    if (rank_metric) *rank_metric = h + 13 * n;
    return 0;
}
