/*
 * oracle/mpmc_oracle.c -- TEST INFRASTRUCTURE ONLY.  Never linked into, imported
 * by, or called from the product (mpmc_b200/); only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may use it.
 *
 * A CPU restatement of MPMC's per-step energy hot path on flat arrays:
 * matrix-free (no pair_t lists, no 3N x 3N A matrix, O(N) memory) but with the
 * reference's own loop order and operation order, so that with nthreads == 1 the
 * sums are accumulated in exactly the order the reference accumulates them.
 * Parity status: PINNED -- tests/test_oracle_golden.py checks this file against
 * fixtures generated from the unmodified reference (oracle/_ref, see
 * oracle/make_golden.py) and against the reference's own golden vectors
 * (mpmc_testing/tests/*.txt, sample_configs_gpu/.../*.energy.dat:2).
 *
 * Each function cites the reference file:line (relative to the reference tree)
 * it follows.  Nothing here is copied: the reference walks linked lists of
 * pair_t and a dense matrix; this walks index ranges and recomputes geometry.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "mpmc_b200.h"

/* reference src/include/defines.h */
#define O_HBAR2 1.11211999e-68
#define O_HBAR4 1.23681087e-136
#define O_KB 1.3806503e-23
#define O_KB2 1.90619525e-46
#define O_M2A2 1.0e20
#define O_M2A4 1.0e40
#define O_AMU2KG 1.66053873e-27
#define O_DEBYE2SKA 85.10597636
#define O_MAXVALUE 1.0e40
#define O_SMALL_dR 1.0e-12
#define O_MAX_ITERATION_COUNT 128

typedef struct {
    const mpmc_b200_params *p;
    int n;
    const double *pos, *q, *alpha, *eps, *sig, *molmass;
    const int *mol, *frozen;
} osys_t;

typedef struct {
    double r, rimg, dimg[3];
} ogeom_t;

typedef struct {
    int rd_excluded, es_excluded, frozen, attractive_only;
    double sigma, epsilon;
} opair_t;

/* minimum_image(): src/energy/pairs.c:230-290.  d = pos_i - pos_j. */
static void o_geom(const osys_t *s, int i, int j, ogeom_t *g) {
    const mpmc_b200_params *P = s->p;
    double d[3], img[3], di[3], r2 = 0, ri2 = 0;
    int p, q;
    for (p = 0; p < 3; p++) d[p] = s->pos[3 * i + p] - s->pos[3 * j + p];
    for (p = 0; p < 3; p++) {
        img[p] = 0;
        for (q = 0; q < 3; q++) img[p] += P->reciprocal_basis[q][p] * d[q];
        img[p] = rint(img[p]);
    }
    for (p = 0; p < 3; p++) {
        di[p] = 0;
        for (q = 0; q < 3; q++) di[p] += P->basis[q][p] * img[q];
    }
    for (p = 0; p < 3; p++) di[p] = d[p] - di[p];
    for (p = 0; p < 3; p++) {
        r2 += d[p] * d[p];
        ri2 += di[p] * di[p];
    }
    g->r = sqrt(r2);
    double ri = sqrt(ri2);
    if (isnan(ri)) { /* pairs.c:279-283 */
        g->rimg = g->r;
        for (p = 0; p < 3; p++) g->dimg[p] = d[p];
    } else {
        g->rimg = ri;
        for (p = 0; p < 3; p++) g->dimg[p] = di[p];
    }
}

/* pair_exclusions(): src/energy/pairs.c:56-227 (LJ-family mixing rules only). */
static void o_excl(const osys_t *s, int i, int j, opair_t *e) {
    double si = s->sig[i], sj = s->sig[j], ei = s->eps[i], ej = s->eps[j];
    e->attractive_only = 0;
    e->sigma = 0;
    e->epsilon = 0;
    if (s->mol[i] == s->mol[j]) { /* pairs.c:61-64 */
        e->rd_excluded = 1;
        e->es_excluded = 1;
    } else {
        e->rd_excluded = (ei == 0.0 || si == 0.0 || ej == 0.0 || sj == 0.0); /* pairs.c:68-71, c6=c8=c10=0 */
        e->es_excluded = (s->q[i] == 0.0 || s->q[j] == 0.0);                 /* pairs.c:74-77 */
    }
    e->frozen = s->frozen[i] && s->frozen[j]; /* pairs.c:81 */
    switch (s->p->mixing) {
        case MPMC_MIX_WALDMAN_HAGLER: { /* pairs.c:85-103 */
            double si3 = si * si * si, si6 = si3 * si3, sj3 = sj * sj * sj, sj6 = sj3 * sj3;
            if (si < 0.0 || sj < 0.0) {
                e->attractive_only = 1;
                e->sigma = pow(0.5 * (si6 + sj6), 1. / 6.); /* epsilon left at its calloc'd 0 */
            } else if (si == 0 || sj == 0) {
                e->sigma = 0;
                e->epsilon = sqrt(ei * ej);
            } else {
                e->sigma = pow(0.5 * (si6 + sj6), 1. / 6.);
                e->epsilon = sqrt(ei * ej) * 2.0 * si3 * sj3 / (si6 + sj6);
            }
        } break;
        case MPMC_MIX_HALGREN: /* pairs.c:104-114 */
            if (si > 0.0 && sj > 0.0)
                e->sigma = (si * si * si + sj * sj * sj) / (si * si + sj * sj);
            else
                e->sigma = 0;
            if (ei > 0.0 && ej > 0.0)
                e->epsilon = 4 * ei * ej / pow(sqrt(ei) + sqrt(ej), 2);
            else
                e->epsilon = 0;
            break;
        case MPMC_MIX_C6: /* pairs.c:194-199 */
            e->sigma = 0.5 * (si + sj);
            if (e->sigma != 0.0)
                e->epsilon = 64.0 * sqrt(ei * ej) * pow(si, 3.0) * pow(sj, 3.0) / pow(si + sj, 6.0);
            else
                e->epsilon = 0.0;
            break;
        default: /* Lorentz-Berthelot, pairs.c:200-211 */
            if (si < 0.0 || sj < 0.0) {
                e->attractive_only = 1;
                e->sigma = 0.5 * (fabs(si) + fabs(sj)); /* epsilon left at its calloc'd 0 */
            } else if (si == 0 || sj == 0) {
                e->sigma = 0;
                e->epsilon = sqrt(ei * ej);
            } else {
                e->sigma = 0.5 * (si + sj);
                e->epsilon = sqrt(ei * ej);
            }
    }
}

/* lj_fh_corr(): src/energy/lj.c:11-54 */
static double o_lj_fh(const osys_t *s, int i, int j, double rimg, double epsilon, double term12, double term6) {
    int order = s->p->feynman_hibbs_order;
    double T = s->p->temperature;
    double ir = 1.0 / rimg, ir2 = ir * ir, ir3 = ir2 * ir, ir4 = ir3 * ir;
    if (order != 2 && order != 4) return NAN;
    double reduced_mass = O_AMU2KG * s->molmass[i] * s->molmass[j] / (s->molmass[i] + s->molmass[j]);
    double dE = -24.0 * epsilon * (2.0 * term12 - term6) * ir;
    double d2E = 24.0 * epsilon * (26.0 * term12 - 7.0 * term6) * ir2;
    double corr = O_M2A2 * (O_HBAR2 / (24.0 * O_KB * T * reduced_mass)) * (d2E + 2.0 * dE / rimg);
    if (order >= 4) {
        double d3E = -1344.0 * epsilon * (6.0 * term12 - term6) * ir3;
        double d4E = 12096.0 * epsilon * (10.0 * term12 - term6) * ir4;
        corr += O_M2A4 * (O_HBAR4 / (1152.0 * O_KB2 * T * T * reduced_mass * reduced_mass)) *
                (15.0 * dE * ir3 + 4.0 * d3E * ir + d4E);
    }
    return corr;
}

/* lj_lrc_corr(): src/energy/lj.c:56-83 (polarvdw / cdvdw off) */
static double o_lrc_pair(const osys_t *s, const opair_t *e, double cutoff) {
    if (e->epsilon != 0 && e->sigma != 0 && !e->frozen) {
        double sig_cut = fabs(e->sigma) / cutoff;
        double sig3 = fabs(e->sigma);
        sig3 *= sig3 * sig3;
        double sig_cut3 = sig_cut * sig_cut * sig_cut;
        double sig_cut9 = sig_cut3 * sig_cut3 * sig_cut3;
        return ((16.0 / 3.0) * M_PI * e->epsilon * sig3) * ((1.0 / 3.0) * sig_cut9 - sig_cut3) / s->p->volume;
    }
    return 0.0; /* pair->lrc stays at its calloc'd 0 */
}

/* lj_lrc_self(): src/energy/lj.c:85-107 */
static double o_lrc_self(const osys_t *s, int i, double cutoff) {
    if (s->sig[i] != 0 && s->eps[i] != 0 && !s->frozen[i]) {
        double sig_cut = fabs(s->sig[i]) / cutoff;
        double sig3 = fabs(s->sig[i]);
        sig3 *= sig3 * sig3;
        double sig_cut3 = sig_cut * sig_cut * sig_cut;
        double sig_cut9 = sig_cut3 * sig_cut3 * sig_cut3;
        return ((16.0 / 3.0) * M_PI * s->eps[i] * sig3) * ((1.0 / 3.0) * sig_cut9 - sig_cut3) / s->p->volume;
    }
    return 0;
}

/* coulombic_real_FH(): src/energy/coulombic.c:115-146 */
static double o_es_fh(const osys_t *s, int i, int j, double r, double gaussian_term, double erfc_term) {
    double rr = r * r, ir = 1.0 / r, ir2 = ir * ir, ir3 = ir * ir2, ir4 = ir2 * ir2;
    double order = s->p->feynman_hibbs_order, alpha = s->p->ewald_alpha, T = s->p->temperature;
    double a2 = alpha * alpha, a3 = a2 * alpha, a4 = a3 * alpha;
    double reduced_mass = O_AMU2KG * s->molmass[i] * s->molmass[j] / (s->molmass[i] + s->molmass[j]);
    double du = -2.0 * alpha * gaussian_term / (r * sqrt(M_PI)) - erfc_term * ir2;
    double d2u = (4.0 / sqrt(M_PI)) * gaussian_term * (a3 + 1.0 * ir2) + 2.0 * erfc_term * ir3;
    double fh2 = (O_M2A2) * (O_HBAR2 / (24.0 * O_KB * T * reduced_mass)) * (d2u + 2.0 * du / r);
    double fh4 = 0.0;
    if (order >= 4) {
        double d3u = (gaussian_term / sqrt(M_PI)) * (-8.0 * (a3 * a2) * r - 8.0 * (a3) / r - 12.0 * alpha * ir3) -
                     6.0 * erfc(alpha * r) * ir4;
        double d4u = (gaussian_term / sqrt(M_PI)) * (8.0 * a3 * a2 + 16.0 * a3 * a4 * rr + 32.0 * a3 * ir2 + 48.0 * ir4) +
                     24.0 * erfc_term * (ir4 * ir);
        fh4 = O_M2A4 * (O_HBAR4 / (1152.0 * (O_KB * O_KB * T * T * reduced_mass * reduced_mass))) *
              (15.0 * du * ir3 + 4.0 * d3u / r + d4u);
    }
    return fh2 + fh4;
}

/* lj(): src/energy/lj.c:175-178 -- with rd_crystal the LJ cutoff (pair test, image sums, LRC) is wider */
static double o_rd_cutoff(const mpmc_b200_params *P) {
    if (P->rd_crystal) return 2.0 * P->cutoff * ((double)P->rd_crystal_order - 0.5);
    return P->cutoff;
}

/* the image sums of lj() with rd_crystal on: src/energy/lj.c:196-216.  d = pos_i - pos_j, NOT imaged. */
static void o_crystal_sums(const osys_t *s, int i, int j, const opair_t *e, double cutoff, double *s6, double *s12) {
    const mpmc_b200_params *P = s->p;
    int l[3], p, q, m = P->rd_crystal_order - 1;
    double a[3], r, sor;
    *s6 = 0;
    *s12 = 0;
    for (l[0] = -m; l[0] <= m; l[0]++)
        for (l[1] = -m; l[1] <= m; l[1]++)
            for (l[2] = -m; l[2] <= m; l[2]++) {
                if (!l[0] && !l[1] && !l[2] && e->rd_excluded) continue; /* no home image for excluded pairs */
                for (p = 0; p < 3; p++) {
                    a[p] = 0;
                    for (q = 0; q < 3; q++) a[p] += P->basis[q][p] * l[q];
                    a[p] += s->pos[3 * i + p] - s->pos[3 * j + p];
                }
                r = sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]);
                if (r > cutoff) continue;
                sor = fabs(e->sigma) / r;
                *s6 += pow(sor, 6);
                *s12 += pow(sor, 12);
            }
}

/* rd_crystal_self(): src/energy/lj.c:109-162 (spectre / polarvdw / cdvdw off): an atom with its own images */
static double o_crystal_self(const osys_t *s, int i, double cutoff) {
    const mpmc_b200_params *P = s->p;
    int l[3], p, q, m = P->rd_crystal_order - 1;
    double a[3], r, sor, s6 = 0, s12 = 0;
    if (s->sig[i] == 0 && s->eps[i] == 0) return 0;
    for (l[0] = -m; l[0] <= m; l[0]++)
        for (l[1] = -m; l[1] <= m; l[1]++)
            for (l[2] = -m; l[2] <= m; l[2]++) {
                if (!l[0] && !l[1] && !l[2]) continue;
                for (p = 0; p < 3; p++) {
                    a[p] = 0;
                    for (q = 0; q < 3; q++) a[p] += P->basis[q][p] * l[q];
                }
                r = sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]);
                if (r > cutoff) continue;
                sor = fabs(s->sig[i]) / r;
                s6 += 0.5 * pow(sor, 6);
                s12 += 0.5 * pow(sor, 12);
            }
    return 4.0 * s->eps[i] * ((s->sig[i] < 0.0 ? 0.0 : s12) - s6);
}

typedef struct {
    double rd_pair, lrc_pair, es_real, min_r;
    double n_tested, n_lj, n_es;
    double rd_running; /* lj.c:256 as written: ONE running sum of (rd_energy + lrc) over the pairs in list order */
} orow_t;

/* One row i of the pair loops of lj() (src/energy/lj.c:181-260) and
 * coulombic_real() (src/energy/coulombic.c:159-191), plus the
 * cavity_absolute_check() minimum (src/energy/energy.c:49-64).
 * When seq != NULL the terms are added to seq[] in reference order instead. */
static void o_pair_row(const osys_t *s, int i, orow_t *acc) {
    const mpmc_b200_params *P = s->p;
    double cutoff = P->cutoff, alpha = P->ewald_alpha, rd_cutoff = o_rd_cutoff(P);
    int j;
    for (j = i + 1; j < s->n; j++) {
        opair_t e;
        ogeom_t g;
        o_excl(s, i, j, &e);
        if (e.frozen && !P->polarization) continue; /* pairs.c:324: no geometry; all terms below skip frozen pairs */
        o_geom(s, i, j, &g);
        acc->n_tested += 1;
        if (s->mol[i] != s->mol[j] && g.rimg < acc->min_r) acc->min_r = g.rimg;
        /* ---- lj() ---- */
        {
            double rd = 0, lrc = 0;
            if (P->rd_lrc) lrc = o_lrc_pair(s, &e, rd_cutoff);
            if ((g.rimg - O_SMALL_dR < rd_cutoff) && (!e.rd_excluded || P->rd_crystal) && !e.frozen) { /* lj.c:191-193 */
                double sor = fabs(e.sigma) / g.rimg;
                double s6 = sor * sor * sor;
                s6 *= s6;
                double s12 = s6 * s6;
                if (P->rd_crystal) o_crystal_sums(s, i, j, &e, rd_cutoff, &s6, &s12);
                double term6 = s6, term12 = e.attractive_only ? 0 : s12;
                rd += 4.0 * e.epsilon * (term12 - term6);
                if (P->feynman_hibbs) rd += o_lj_fh(s, i, j, g.rimg, e.epsilon, term12, term6);
                acc->n_lj += 1;
            }
            acc->rd_pair += rd;
            acc->lrc_pair += lrc;
            acc->rd_running += rd + lrc;
        }
        /* ---- coulombic_wolf(): coulombic.c:269-308 (option `wolf`) ---- */
        if (!P->rd_only && P->es_wolf) {
            double R = cutoff, iR = 1.0 / R, erfaRoverR = erf(alpha * R) / R;
            double r = g.rimg, ir = 1.0 / r;
            if (!e.frozen && !e.es_excluded && (r < R)) {
                acc->es_real += s->q[i] * s->q[j] * (ir - erfaRoverR - iR * iR * (R - r));
                acc->n_es += 1;
            }
        } else
        /* ---- coulombic_real() ---- */
        if (!P->rd_only && !e.frozen) {
            double r = g.rimg, es = 0, intra = 0;
            if (!((r > cutoff) || e.es_excluded)) {
                double erfc_term = erfc(alpha * r);
                double gaussian_term = exp(-alpha * alpha * r * r);
                es += s->q[i] * s->q[j] * erfc_term / r;
                if (P->feynman_hibbs) es += o_es_fh(s, i, j, r, gaussian_term, erfc_term);
                acc->n_es += 1;
            } else if (e.es_excluded) {
                intra = s->q[i] * s->q[j] * erf(alpha * g.r) / g.r; /* un-imaged r: coulombic.c:180 */
            }
            acc->es_real += es - intra;
        }
    }
}

/* coulombic_reciprocal(): src/energy/coulombic.c:42-95 */
static double o_recip(const osys_t *s, double *nk_out, double *natoms_out) {
    const mpmc_b200_params *P = s->p;
    int kmax = P->ewald_kmax, l[3], p, q, i, nk = 0, nq = 0;
    double alpha = P->ewald_alpha, potential = 0;
    int *idx = (int *)malloc(sizeof(int) * (s->n > 0 ? s->n : 1));
    for (i = 0; i < s->n; i++)
        if (!s->frozen[i] && s->q[i] != 0.0) idx[nq++] = i;
    /* enumerate the hemisphere first so the k loop can be threaded */
    int cap = (kmax + 1) * (2 * kmax + 1) * (2 * kmax + 1);
    int *ls = (int *)malloc(sizeof(int) * 3 * (cap > 0 ? cap : 1));
    for (l[0] = 0; l[0] <= kmax; l[0]++)
        for (l[1] = (!l[0] ? 0 : -kmax); l[1] <= kmax; l[1]++)
            for (l[2] = ((!l[0] && !l[1]) ? 1 : -kmax); l[2] <= kmax; l[2]++) {
                if (l[0] * l[0] + l[1] * l[1] + l[2] * l[2] > kmax * kmax) continue;
                ls[3 * nk] = l[0];
                ls[3 * nk + 1] = l[1];
                ls[3 * nk + 2] = l[2];
                nk++;
            }
    double *term = (double *)malloc(sizeof(double) * (nk > 0 ? nk : 1));
    int ik;
#pragma omp parallel for schedule(static) private(p, q, i)
    for (ik = 0; ik < nk; ik++) {
        double k[3], k_squared, SF_re = 0, SF_im = 0;
        for (p = 0; p < 3; p++) {
            k[p] = 0;
            for (q = 0; q < 3; q++) k[p] += 2.0 * M_PI * P->reciprocal_basis[p][q] * ls[3 * ik + q];
        }
        k_squared = k[0] * k[0] + k[1] * k[1] + k[2] * k[2];
        for (i = 0; i < nq; i++) {
            const double *r = s->pos + 3 * idx[i];
            double pp = k[0] * r[0] + k[1] * r[1] + k[2] * r[2];
            SF_re += s->q[idx[i]] * cos(pp);
            SF_im += s->q[idx[i]] * sin(pp);
        }
        term[ik] = exp(-k_squared / (4.0 * alpha * alpha)) / k_squared * (SF_re * SF_re + SF_im * SF_im);
    }
    for (ik = 0; ik < nk; ik++) potential += term[ik];
    potential *= 4.0 * M_PI / P->volume;
    free(term);
    free(ls);
    free(idx);
    if (nk_out) *nk_out = nk;
    if (natoms_out) *natoms_out = nq;
    return potential;
}

/* coulombic_self(): src/energy/coulombic.c:97-112 */
static double o_self(const osys_t *s) {
    double alpha = s->p->ewald_alpha, self_potential = 0.0;
    int i;
    for (i = 0; i < s->n; i++) {
        if (s->frozen[i]) continue;
        self_potential -= alpha * s->q[i] * s->q[i] / sqrt(M_PI);
    }
    return self_potential;
}

/* ------------------------------------------------------------------ */
/* static field                                                        */
/* ------------------------------------------------------------------ */
#define O_OneOverSqrtPi_field 0.56418958354 /* src/polarization/thole_field.c:10 */

/* Contribution of the unordered pair (lo,hi), lo<hi, to row `row` (== lo or hi):
 * thole_field_nopbc() src/polarization/thole_field.c:39-68,
 * thole_field_wolf()  src/polarization/thole_field.c:71-124 */
static void o_field_pair(const osys_t *s, int lo, int hi, int row, const double *wolf_table, int wolf_table_max,
                         double *E, double *count) {
    const mpmc_b200_params *P = s->p;
    ogeom_t g;
    int p;
    if (s->frozen[lo] && s->frozen[hi]) return;
    if (s->mol[lo] == s->mol[hi]) return;
    o_geom(s, lo, hi, &g);
    double r = g.rimg;
    if (!((r - O_SMALL_dR < P->cutoff) && (r != 0.))) return;
    if (count) *count += 1;
    if (P->polar_wolf || P->polar_wolf_full) {
        double R = P->cutoff, rR = 1. / R, a = P->polar_wolf_alpha, rr = 1. / r;
        double erR = erfc(a * R);
        double cutoffterm = (erR * rR * rR + 2.0 * a * O_OneOverSqrtPi_field * exp(-a * a * R * R) * rR);
        double bigmess = 0;
        if ((a != 0) & P->polar_wolf_alpha_lookup) { /* polar_wolf_alpha_getval(): polar_wolf_lookup.c:28-32 */
            int k = (int)(r * 1000);
            bigmess = (k >= wolf_table_max) ? 0.0 : wolf_table[k];
        } else if (a != 0)
            bigmess = (erfc(a * r) * rr * rr + 2.0 * a * O_OneOverSqrtPi_field * exp(-a * a * r * r) * rr);
        for (p = 0; p < 3; p++) {
            if (a == 0) {
                if (row == lo)
                    E[p] += (s->q[hi]) * (rr * rr - rR * rR) * g.dimg[p] * rr;
                else
                    E[p] -= (s->q[lo]) * (rr * rr - rR * rR) * g.dimg[p] * rr;
            } else {
                if (row == lo)
                    E[p] += s->q[hi] * (bigmess - cutoffterm) * g.dimg[p] * rr;
                else
                    E[p] -= s->q[lo] * (bigmess - cutoffterm) * g.dimg[p] * rr;
            }
        }
    } else {
        for (p = 0; p < 3; p++) {
            if (row == lo)
                E[p] += s->q[hi] * g.dimg[p] / (r * r * r);
            else
                E[p] -= s->q[lo] * g.dimg[p] / (r * r * r);
        }
    }
}

/* real_term(): src/polarization/polar_ewald.c:38-74 */
#define O_OneOverSqrtPi_ewald 0.5641895835477562869480794515607725858440506293289988
static void o_ewald_real_pair(const osys_t *s, int lo, int hi, int row, double *E, double *count) {
    const mpmc_b200_params *P = s->p;
    ogeom_t g;
    opair_t e;
    int p;
    double a = P->polar_ewald_alpha, factor;
    if (s->frozen[lo] && s->frozen[hi]) return;
    o_geom(s, lo, hi, &g);
    double r = g.rimg;
    if ((r > P->cutoff) || (r == 0.0)) return;
    if (count) *count += 1;
    o_excl(s, lo, hi, &e);
    double r2 = r * r;
    if (e.es_excluded)
        factor = (2.0 * a * O_OneOverSqrtPi_ewald * exp(-a * a * r2) * r - erf(a * r)) / (r * r2);
    else
        factor = (2.0 * a * O_OneOverSqrtPi_ewald * exp(-a * a * r2) * r + erfc(a * r)) / (r2 * r);
    for (p = 0; p < 3; p++) {
        if (row == lo)
            E[p] += factor * s->q[hi] * g.dimg[p];
        else
            E[p] -= factor * s->q[lo] * g.dimg[p];
    }
}

/* recip_term(): src/polarization/polar_ewald.c:78-132 (all atoms, frozen included) */
static void o_ewald_recip_field(const osys_t *s, double *ef) {
    const mpmc_b200_params *P = s->p;
    int kmax = P->ewald_kmax, l[3], p, q, i;
    double ea = P->polar_ewald_alpha;
    for (l[0] = 0; l[0] <= kmax; l[0]++)
        for (l[1] = (!l[0] ? 0 : -kmax); l[1] <= kmax; l[1]++)
            for (l[2] = ((!l[0] && !l[1]) ? 1 : -kmax); l[2] <= kmax; l[2]++) {
                double k[3], k2, kweight[3], float1 = 0, float2 = 0;
                if (l[0] * l[0] + l[1] * l[1] + l[2] * l[2] > kmax * kmax) continue;
                for (p = 0; p < 3; p++) {
                    k[p] = 0;
                    for (q = 0; q < 3; q++) k[p] += 2.0 * M_PI * P->reciprocal_basis[p][q] * l[q];
                }
                k2 = k[0] * k[0] + k[1] * k[1] + k[2] * k[2];
                for (p = 0; p < 3; p++) kweight[p] = k[p] / k2 * exp(-k2 / (4.0 * ea * ea));
                for (i = 0; i < s->n; i++) {
                    const double *r = s->pos + 3 * i;
                    double pp = k[0] * r[0] + k[1] * r[1] + k[2] * r[2];
                    float1 += s->q[i] * cos(pp);
                    float2 += s->q[i] * sin(pp);
                }
#pragma omp parallel for schedule(static) private(p)
                for (i = 0; i < s->n; i++) {
                    const double *r = s->pos + 3 * i;
                    double pp = k[0] * r[0] + k[1] * r[1] + k[2] * r[2];
                    for (p = 0; p < 3; p++) {
                        ef[3 * i + p] += kweight[p] * sin(pp) * float1;
                        ef[3 * i + p] -= kweight[p] * cos(pp) * float2;
                    }
                }
            }
    for (i = 0; i < s->n; i++)
        for (p = 0; p < 3; p++) ef[3 * i + p] *= 8.0 * M_PI / P->volume;
}

/* thole_field(): src/polarization/thole_field.c:14-36 */
static void o_static_field(const osys_t *s, double *ef, double *pairs_field) {
    const mpmc_b200_params *P = s->p;
    int n = s->n, i;
    double *wolf_table = NULL;
    int wolf_table_max = 0;
    double total = 0;
    memset(ef, 0, sizeof(double) * 3 * n);
    if (P->polar_ewald) {
        o_ewald_recip_field(s, ef); /* ewald_estatic(): polar_ewald.c:166-174: recip then real */
    } else if ((P->polar_wolf || P->polar_wolf_full) && P->polar_wolf_alpha_lookup) {
        /* polar_wolf_alpha_lookup_init(): src/polarization/polar_wolf_lookup.c:8-26 */
        double a = P->polar_wolf_alpha;
        wolf_table_max = (int)(ceil(P->polar_wolf_alpha_lookup_cutoff) * 1000);
        wolf_table = (double *)malloc(sizeof(double) * (wolf_table_max > 0 ? wolf_table_max : 1));
        for (i = 1; i < wolf_table_max; i++) {
            double r = (double)i / 1000., rr = 1.0 / r;
            wolf_table[i] = erfc(a * r) * rr * rr + 2.0 * a * O_OneOverSqrtPi_field * exp(-a * a * r * r) * rr;
        }
        if (wolf_table_max > 1) wolf_table[0] = wolf_table[1];
    }
    /* Row-owned accumulation.  For row a the reference's updates arrive in the
     * order: pairs (i,a) for i = 0..a-1 (while the outer loop is at i), then pairs
     * (a,j) for j = a+1.. -- exactly the j = 0..n-1 order used here, so every
     * ef_static component is accumulated in the reference's order. */
#pragma omp parallel for schedule(dynamic, 16) reduction(+ : total)
    for (i = 0; i < n; i++) {
        double E[3] = {ef[3 * i], ef[3 * i + 1], ef[3 * i + 2]}, cnt = 0;
        int j;
        for (j = 0; j < n; j++) {
            if (j == i) continue;
            int lo = i < j ? i : j, hi = i < j ? j : i;
            if (P->polar_ewald)
                o_ewald_real_pair(s, lo, hi, i, E, &cnt);
            else
                o_field_pair(s, lo, hi, i, wolf_table, wolf_table_max, E, &cnt);
        }
        ef[3 * i] = E[0];
        ef[3 * i + 1] = E[1];
        ef[3 * i + 2] = E[2];
        total += cnt;
    }
    if (pairs_field) *pairs_field = total / 2;
    free(wolf_table);
}

/* ------------------------------------------------------------------ */
/* dipole field tensor, matrix-free                                    */
/* ------------------------------------------------------------------ */
/* One 3x3 block T_ij of thole_amatrix(): src/polarization/thole_matrix.c:73-143.
 * The reference fills the block for i<j from pair (i,j) and mirrors it; geometry
 * is even in d, so the block is evaluated from (min,max). */
static void o_tensor(const osys_t *s, int i, int j, double T[3][3]) {
    const mpmc_b200_params *P = s->p;
    int lo = i < j ? i : j, hi = i < j ? j : i, p, q;
    ogeom_t g;
    double damp1 = 0, damp2 = 0, wdamp1 = 0, wdamp2 = 0, ir = 0, ir3, ir5;
    double rcut = P->cutoff, rcut2 = rcut * rcut, rcut3 = rcut2 * rcut;
    double l = P->polar_damp, l2 = l * l, l3 = l2 * l;
    o_geom(s, lo, hi, &g);
    double r = g.rimg, r2 = r * r;
    if (g.rimg == 0.)
        ir3 = ir5 = O_MAXVALUE;
    else {
        ir = 1.0 / r;
        ir3 = ir * ir * ir;
        ir5 = ir3 * ir * ir;
    }
    switch (P->damp_type) {
        case MPMC_DAMP_OFF: { /* thole_matrix.c:92-97: es_excluded pairs are switched off */
            opair_t e;
            o_excl(s, lo, hi, &e);
            if (e.es_excluded)
                damp1 = damp2 = wdamp1 = wdamp2 = 0.0;
            else
                damp1 = damp2 = wdamp1 = wdamp2 = 1.0;
        } break;
        case MPMC_DAMP_LINEAR: { /* thole_matrix.c:98-107 */
            double sl = l * pow(s->alpha[lo] * s->alpha[hi], 1.0 / 6.0);
            double v = r / sl;
            if (r < sl) {
                damp1 = (4.0 - 3.0 * v) * v * v * v;
                damp2 = v * v * v * v;
            } else
                damp1 = damp2 = 1.0;
        } break;
        default: { /* exponential, thole_matrix.c:108-116 */
            double explr = exp(-l * r);
            damp1 = 1.0 - explr * (0.5 * l2 * r2 + l * r + 1.0);
            damp2 = damp1 - explr * (l3 * r2 * r / 6.0);
            if (P->polar_wolf_full) {
                double explrcut = exp(-l * rcut);
                wdamp1 = 1.0 - explrcut * (0.5 * l2 * rcut2 + l * rcut + 1.0);
                wdamp2 = wdamp1 - explrcut * (l3 * rcut3 / 6.0);
            }
        }
    }
    for (p = 0; p < 3; p++)
        for (q = 0; q < 3; q++) {
            double t = -3.0 * g.dimg[p] * g.dimg[q] * damp2 * ir5;
            if (P->polar_wolf_full) t -= -3.0 * g.dimg[p] * g.dimg[q] * wdamp2 * ir * ir / rcut3;
            if (p == q) {
                t += damp1 * ir3;
                if (P->polar_wolf_full) t -= wdamp1 / (rcut3);
            }
            T[p][q] = t;
        }
}

/* sum_j!=i T_ij mu_j over polarizable j, j ascending (the reference also adds the
 * exact zeros of non-polarizable j, whose mu is 0: thole_iterative.c:40-45). */
static void o_row_contract(const osys_t *s, int i, const int *pol, int npol, const double *mu, double *acc) {
    int jj, p;
    for (jj = 0; jj < npol; jj++) {
        int j = pol[jj];
        double T[3][3];
        if (j == i) continue;
        o_tensor(s, i, j, T);
        for (p = 0; p < 3; p++) acc[p] -= T[p][0] * mu[3 * j] + T[p][1] * mu[3 * j + 1] + T[p][2] * mu[3 * j + 2];
    }
}

/* rank metric of pairs(): src/energy/pairs.c:337-360 */
static void o_rank_metric(const osys_t *s, const int *pol, int npol, double *rank) {
    int n = s->n, a, b, i;
    double rmin = O_MAXVALUE;
    for (i = 0; i < n; i++) rank[i] = 0;
#pragma omp parallel for schedule(dynamic, 16) reduction(min : rmin)
    for (a = 0; a < npol; a++) {
        int bb;
        for (bb = a + 1; bb < npol; bb++) {
            ogeom_t g;
            o_geom(s, pol[a], pol[bb], &g);
            if (g.rimg < rmin) rmin = g.rimg;
        }
    }
#pragma omp parallel for schedule(dynamic, 16)
    for (a = 0; a < npol; a++) {
        double c = 0;
        for (b = 0; b < npol; b++) {
            ogeom_t g;
            if (b == a) continue;
            o_geom(s, pol[a] < pol[b] ? pol[a] : pol[b], pol[a] < pol[b] ? pol[b] : pol[a], &g);
            if (g.r <= rmin * 1.5) c += 1.0; /* un-imaged r against imaged rmin: pairs.c:354 */
        }
        rank[pol[a]] = c;
    }
}

/* thole_iterative(): src/polarization/thole_iterative.c:168-259, with
 * init_dipoles :13-25, contract_dipoles :27-59, calc_dipole_rrms :61-92,
 * are_we_done_yet :94-117, palmo_contraction :119-141, update_ranking :143-164 */
static int o_thole_iterative(const osys_t *s, const double *ef_static, const double *rank_metric, double *mu,
                             double *ef_induced, double *ef_induced_change, double *old_mu, double *new_mu,
                             double *rrms, int *iter_success, double *contractions) {
    const mpmc_b200_params *P = s->p;
    int n = s->n, i, p, npol = 0, iteration_counter = 0, keep_iterating = 1;
    int *ranked = (int *)malloc(sizeof(int) * (n > 0 ? n : 1));
    int *pol = (int *)malloc(sizeof(int) * (n > 0 ? n : 1));
    int gs = P->polar_gs || P->polar_gs_ranked;
    for (i = 0; i < n; i++) {
        ranked[i] = i;
        if (s->alpha[i] != 0.0) pol[npol++] = i;
    }
    for (i = 0; i < n; i++)
        for (p = 0; p < 3; p++) {
            mu[3 * i + p] = s->alpha[i] * (ef_static[3 * i + p] + 0.0);
            if (!P->polar_sor && !P->polar_esor) mu[3 * i + p] *= P->polar_gamma;
        }
    if (P->polar_zodid) {
        free(ranked);
        free(pol);
        return 0;
    }
    while (keep_iterating) {
        iteration_counter++;
        if (iteration_counter >= O_MAX_ITERATION_COUNT && P->polar_precision) { /* :199-210 */
            for (i = 0; i < n; i++)
                for (p = 0; p < 3; p++) {
                    mu[3 * i + p] = s->alpha[i] * (ef_static[3 * i + p] + 0.0);
                    ef_induced_change[3 * i + p] = 0.0;
                }
            *iter_success = 1;
            free(ranked);
            free(pol);
            return iteration_counter;
        }
        memset(ef_induced, 0, sizeof(double) * 3 * n);
        if (P->polar_rrms || P->polar_precision > 0 || P->polar_sor || P->polar_esor)
            memcpy(old_mu, mu, sizeof(double) * 3 * n);

        /* contract_dipoles */
        if (gs) {
            for (i = 0; i < n; i++) {
                int index = ranked[i];
                if (s->alpha[index] == 0) {
                    for (p = 0; p < 3; p++) new_mu[3 * index + p] = mu[3 * index + p] = 0;
                    continue;
                }
                o_row_contract(s, index, pol, npol, mu, ef_induced + 3 * index);
                for (p = 0; p < 3; p++) {
                    new_mu[3 * index + p] = s->alpha[index] * (ef_static[3 * index + p] + 0.0 + ef_induced[3 * index + p]);
                    mu[3 * index + p] = new_mu[3 * index + p];
                }
            }
        } else {
#pragma omp parallel for schedule(dynamic, 8) private(p)
            for (i = 0; i < n; i++) {
                if (s->alpha[i] == 0) {
                    for (p = 0; p < 3; p++) new_mu[3 * i + p] = 0;
                    continue;
                }
                o_row_contract(s, i, pol, npol, mu, ef_induced + 3 * i);
                for (p = 0; p < 3; p++)
                    new_mu[3 * i + p] = s->alpha[i] * (ef_static[3 * i + p] + 0.0 + ef_induced[3 * i + p]);
            }
            for (i = 0; i < n; i++)
                if (s->alpha[i] == 0)
                    for (p = 0; p < 3; p++) mu[3 * i + p] = 0;
        }
        *contractions += 1;

        if (P->polar_rrms || P->polar_precision > 0) { /* calc_dipole_rrms */
            for (i = 0; i < n; i++) {
                double acc = 0;
                for (p = 0; p < 3; p++) {
                    double carry = new_mu[3 * i + p] - old_mu[3 * i + p];
                    acc += carry * carry;
                }
                acc /= new_mu[3 * i] * new_mu[3 * i] + new_mu[3 * i + 1] * new_mu[3 * i + 1] +
                       new_mu[3 * i + 2] * new_mu[3 * i + 2];
                acc = sqrt(acc);
                if (!isfinite(acc)) acc = 0;
                rrms[i] = acc;
            }
        }

        /* are_we_done_yet */
        keep_iterating = 0;
        if (P->polar_precision == 0.0) {
            if (iteration_counter != P->polar_max_iter) keep_iterating = 1;
        } else {
            double allowed = P->polar_precision * P->polar_precision * O_DEBYE2SKA * O_DEBYE2SKA;
            for (i = 0; i < n && !keep_iterating; i++)
                for (p = 0; p < 3; p++) {
                    double err = new_mu[3 * i + p] - old_mu[3 * i + p];
                    if (err * err > allowed) {
                        keep_iterating = 1;
                        break;
                    }
                }
        }

        if (P->polar_palmo && !keep_iterating) { /* palmo_contraction, uses the current mu[] */
#pragma omp parallel for schedule(dynamic, 8) private(p)
            for (i = 0; i < n; i++) {
                for (p = 0; p < 3; p++) ef_induced_change[3 * i + p] = -ef_induced[3 * i + p];
                o_row_contract(s, i, pol, npol, mu, ef_induced_change + 3 * i);
            }
            *contractions += 1;
        }

        if (P->polar_gs_ranked && keep_iterating) { /* update_ranking: stable descending bubble sort */
            int a, b, sorted;
            for (a = 0; a < n; a++) {
                for (b = 0, sorted = 1; b < n - 1; b++)
                    if (rank_metric[ranked[b]] < rank_metric[ranked[b + 1]]) {
                        int tmp = ranked[b];
                        ranked[b] = ranked[b + 1];
                        ranked[b + 1] = tmp;
                        sorted = 0;
                    }
                if (sorted) break;
            }
        }

        for (i = 0; i < n; i++)
            for (p = 0; p < 3; p++) {
                if (P->polar_sor)
                    mu[3 * i + p] = P->polar_gamma * new_mu[3 * i + p] + (1.0 - P->polar_gamma) * old_mu[3 * i + p];
                else if (P->polar_esor)
                    mu[3 * i + p] = (1.0 - exp(-P->polar_gamma * iteration_counter)) * new_mu[3 * i + p] +
                                    exp(-P->polar_gamma * iteration_counter) * old_mu[3 * i + p];
                else
                    mu[3 * i + p] = new_mu[3 * i + p];
            }
    }
    free(ranked);
    free(pol);
    return iteration_counter;
}


/* ------------------------------------------------------------------ */
/* full-Ewald polarization                                             */
/* ------------------------------------------------------------------ */
/* damp_factor(): src/polarization/polar_ewald.c:28-36 */
static double o_damp_factor(double t, int i) {
    double poo = 1.0 + t + 0.5 * t * t;
    if (i == 3) poo += t * t * t / 6.0;
    return poo * exp(-t);
}

/* One row of induced_real_term(): src/polarization/polar_ewald.c:176-230.  Both atoms polarizable,
 * rimg <= cutoff; frozen-frozen and intramolecular pairs are NOT excluded.  The reference visits
 * pair (lo,hi) once and updates both rows; row i receives its updates in ascending partner order,
 * component c in the order q = 0,1,2 (see the p,q loop at :206-226). */
static void o_ewald_induced_real_row(const osys_t *s, int i, const int *pol, int npol, const double *mu, double *acc) {
    const mpmc_b200_params *P = s->p;
    double a = P->polar_ewald_alpha, l = P->polar_damp;
    int jj, p, q;
    for (jj = 0; jj < npol; jj++) {
        int j = pol[jj];
        ogeom_t g;
        if (j == i) continue;
        o_geom(s, i < j ? i : j, i < j ? j : i, &g);
        if (g.rimg > P->cutoff) continue;
        double r = g.rimg, ir = 1.0 / r, ir3 = ir * ir * ir, ir5 = ir * ir * ir3;
        double erfcar = erfc(a * r), expa2r2 = exp(-a * a * r * r);
        double s2 = erfcar + 2.0 * a * r * O_OneOverSqrtPi_ewald * expa2r2 +
                    4.0 * a * a * a * r * r * r / 3.0 * O_OneOverSqrtPi_ewald * expa2r2 - o_damp_factor(l * r, 3);
        double s1d = erfcar + 2.0 * a * r * O_OneOverSqrtPi_ewald * expa2r2 - o_damp_factor(l * r, 2);
        for (p = 0; p < 3; p++)
            for (q = 0; q < 3; q++) {
                double s1 = (p == q) ? s1d : 0;
                double T = 3.0 * g.dimg[p] * g.dimg[q] * s2 * ir5 - s1 * ir3;
                acc[p] += T * mu[3 * j + q];
            }
    }
}

/* induced_recip_term(): src/polarization/polar_ewald.c:232-294, all atoms.  The loop at :265-266
 * overwrites kweight for p = 0,1,2, so only k[2] survives and the same scalar is added to all three
 * field components (:282-286) -- reproduced as is. */
static void o_ewald_induced_recip(const osys_t *s, const double *mu, double *ef) {
    const mpmc_b200_params *P = s->p;
    int kmax = P->ewald_kmax, l[3], p, q, i, n = s->n;
    double a = P->polar_ewald_alpha;
    for (l[0] = 0; l[0] <= kmax; l[0]++)
        for (l[1] = (!l[0] ? 0 : -kmax); l[1] <= kmax; l[1]++)
            for (l[2] = ((!l[0] && !l[1]) ? 1 : -kmax); l[2] <= kmax; l[2]++) {
                double k[3], k2, kweight = 0, Pcos = 0, Psin = 0;
                if (l[0] * l[0] + l[1] * l[1] + l[2] * l[2] > kmax * kmax) continue;
                for (p = 0; p < 3; p++) {
                    k[p] = 0;
                    for (q = 0; q < 3; q++) k[p] += 2.0 * M_PI * P->reciprocal_basis[p][q] * l[q];
                }
                k2 = k[0] * k[0] + k[1] * k[1] + k[2] * k[2];
                for (p = 0; p < 3; p++) kweight = 8.0 * M_PI / P->volume * exp(-k2 / (4.0 * a * a)) / k2 * k[p];
                for (i = 0; i < n; i++) {
                    const double *r = s->pos + 3 * i, *m = mu + 3 * i;
                    double d1 = k[0] * m[0] + k[1] * m[1] + k[2] * m[2];
                    double d2 = k[0] * r[0] + k[1] * r[1] + k[2] * r[2];
                    Pcos += d1 * cos(d2);
                    Psin += d1 * sin(d2);
                }
#pragma omp parallel for schedule(static) private(p)
                for (i = 0; i < n; i++) {
                    const double *r = s->pos + 3 * i;
                    for (p = 0; p < 3; p++) {
                        double d1 = k[0] * r[0] + k[1] * r[1] + k[2] * r[2];
                        ef[3 * i + p] += kweight * (-sin(d1) * Psin - cos(d1) * Pcos);
                    }
                }
            }
}

/* clear_ef_induced + induced_real_term + induced_recip_term + induced_corr_term
 * (polar_ewald.c:151-163,176-317) for the current mu */
#define O_SqrtPi_ewald 1.77245385091 /* polar_ewald.c:7, the truncated constant is the reference's */
static void o_ewald_induced_field(const osys_t *s, const int *pol, int npol, const double *mu, double *ef) {
    const mpmc_b200_params *P = s->p;
    int n = s->n, i, p;
    double a = P->polar_ewald_alpha, totalmu[3] = {0, 0, 0};
    memset(ef, 0, sizeof(double) * 3 * n);
#pragma omp parallel for schedule(dynamic, 8)
    for (i = 0; i < n; i++)
        if (s->alpha[i] != 0) o_ewald_induced_real_row(s, i, pol, npol, mu, ef + 3 * i);
    o_ewald_induced_recip(s, mu, ef);
    for (i = 0; i < n; i++)
        for (p = 0; p < 3; p++) totalmu[p] += mu[3 * i + p];
    for (i = 0; i < n; i++)
        for (p = 0; p < 3; p++)
            ef[3 * i + p] += -4.0 * M_PI / (3.0 * P->volume) * totalmu[p] + 4.0 * a * a * a / (3.0 * O_SqrtPi_ewald) * mu[3 * i + p];
}

/* ewald_full(): src/polarization/polar_ewald.c:375-418 with init_dipoles_ewald :135-148,
 * new_dipoles :319-345, ewald_palmo_contraction :347-371; calc_dipole_rrms / are_we_done_yet are the
 * ones of thole_iterative.c:61-117.  The iteration counter is 0-based here (are_we_done_yet sees
 * 0,1,2,...: polar_max_iter = m gives m+1 passes) and the rrms is taken BEFORE new_dipoles, i.e.
 * from the previous pass' new_mu / old_mu -- both as in the reference.  Returns the number of
 * passes (the reference does not report it: nodestats->polarization_iterations is left alone). */
static int o_ewald_full(const osys_t *s, const double *ef_static, double *mu, double *ef_induced, double *eic,
                        double *old_mu, double *new_mu, double *rrms, int *iter_success, double *contractions) {
    const mpmc_b200_params *P = s->p;
    int n = s->n, i, p, npol = 0, count = 0, keep = 1;
    int *pol = (int *)malloc(sizeof(int) * (n > 0 ? n : 1));
    for (i = 0; i < n; i++)
        if (s->alpha[i] != 0.0) pol[npol++] = i;
    for (i = 0; i < n; i++)
        for (p = 0; p < 3; p++) {
            old_mu[3 * i + p] = 0;
            new_mu[3 * i + p] = mu[3 * i + p] = s->alpha[i] * ef_static[3 * i + p];
        }
    while (keep) {
        if (count >= O_MAX_ITERATION_COUNT && P->polar_precision) {
            *iter_success = 1;
            break;
        }
        o_ewald_induced_field(s, pol, npol, mu, ef_induced);
        *contractions += 1;
        if (P->polar_rrms || P->polar_precision > 0)
            for (i = 0; i < n; i++) {
                double acc = 0;
                for (p = 0; p < 3; p++) {
                    double carry = new_mu[3 * i + p] - old_mu[3 * i + p];
                    acc += carry * carry;
                }
                acc /= new_mu[3 * i] * new_mu[3 * i] + new_mu[3 * i + 1] * new_mu[3 * i + 1] + new_mu[3 * i + 2] * new_mu[3 * i + 2];
                acc = sqrt(acc);
                if (!isfinite(acc)) acc = 0;
                rrms[i] = acc;
            }
        for (i = 0; i < n; i++)
            for (p = 0; p < 3; p++) {
                int k = 3 * i + p;
                old_mu[k] = mu[k];
                new_mu[k] = s->alpha[i] * (ef_static[k] + ef_induced[k]);
                if (P->polar_sor)
                    new_mu[k] = mu[k] = P->polar_gamma * new_mu[k] + (1.0 - P->polar_gamma) * old_mu[k];
                else if (P->polar_esor)
                    new_mu[k] = mu[k] = (1.0 - exp(-P->polar_gamma * (count + 1))) * new_mu[k] +
                                        exp(-P->polar_gamma * (count + 1)) * old_mu[k];
                else
                    mu[k] = new_mu[k];
            }
        keep = 0;
        if (P->polar_precision == 0.0) {
            if (count != P->polar_max_iter) keep = 1;
        } else {
            double allowed = P->polar_precision * P->polar_precision * O_DEBYE2SKA * O_DEBYE2SKA;
            for (i = 0; i < n && !keep; i++)
                for (p = 0; p < 3; p++) {
                    double err = new_mu[3 * i + p] - old_mu[3 * i + p];
                    if (err * err > allowed) {
                        keep = 1;
                        break;
                    }
                }
        }
        if (P->polar_palmo && !keep) {
            o_ewald_induced_field(s, pol, npol, mu, ef_induced);
            *contractions += 1;
            for (i = 0; i < n; i++) {
                if (s->alpha[i] == 0) continue;
                for (p = 0; p < 3; p++)
                    eic[3 * i + p] = ef_induced[3 * i + p] - (new_mu[3 * i + p] / s->alpha[i] - ef_static[3 * i + p]);
            }
        }
        count++;
    }
    free(pol);
    return count;
}

/* ------------------------------------------------------------------ */
/* entry points                                                        */
/* ------------------------------------------------------------------ */
static int o_check(const mpmc_b200_params *P) {
    if (P->polarization && !P->polar_iterative && !P->polar_zodid && !P->polar_ewald_full)
        return 3; /* LU inversion path: out of scope */
    if (P->es_wolf && !P->rd_only && P->feynman_hibbs) return 3; /* coulombic.c:293-297: the reference dies */
    if (P->rd_crystal && P->rd_crystal_order <= 0) return 3;      /* check_input.c:1258-1262 */
    return 0;
}

/* polar(): src/energy/polar.c:31-135.  Arrays may be NULL (scratch is used). */
int mpmc_oracle_polar(const mpmc_b200_params *P, int n, const double *pos, const double *charge,
                      const double *polarizability, const double *epsilon, const double *sigma,
                      const double *mol_mass, const int *mol_id, const int *frozen, mpmc_b200_energies *out,
                      mpmc_b200_counters *cnt, double *mu, double *ef_static, double *ef_induced,
                      double *ef_induced_change, double *dipole_rrms, double *rank_metric, int nthreads) {
    osys_t s = {P, n, pos, charge, polarizability, epsilon, sigma, mol_mass, mol_id, frozen};
    int rc = o_check(P), i, p, npol = 0;
    if (rc) return rc;
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
    size_t sz = sizeof(double) * 3 * (n > 0 ? n : 1);
    double *b_mu = mu ? mu : (double *)malloc(sz), *b_es = ef_static ? ef_static : (double *)malloc(sz);
    double *b_ei = ef_induced ? ef_induced : (double *)malloc(sz);
    double *b_eic = ef_induced_change ? ef_induced_change : (double *)malloc(sz);
    double *b_rr = dipole_rrms ? dipole_rrms : (double *)malloc(sz), *b_rk = rank_metric ? rank_metric : (double *)malloc(sz);
    double *old_mu = (double *)calloc(3 * (n > 0 ? n : 1), sizeof(double));
    double *new_mu = (double *)calloc(3 * (n > 0 ? n : 1), sizeof(double));
    int *pol = (int *)malloc(sizeof(int) * (n > 0 ? n : 1));
    double pairs_field = 0, contractions = 0, potential = 0, rr = 0;
    memset(b_ei, 0, sz);
    memset(b_eic, 0, sz);
    memset(b_rr, 0, sizeof(double) * (n > 0 ? n : 1));
    memset(b_rk, 0, sizeof(double) * (n > 0 ? n : 1));
    for (i = 0; i < n; i++)
        if (polarizability[i] != 0.0) pol[npol++] = i;
    if (P->polar_iterative && P->polar_gs_ranked) o_rank_metric(&s, pol, npol, b_rk);
    out->iter_success = 0;
    if (P->polar_ewald_full) { /* polar.c:54-57: ewald_full() does its own Ewald static field (zero_out, recip, real) */
        mpmc_b200_params P2 = *P;
        osys_t s2 = s;
        P2.polar_ewald = 1;
        s2.p = &P2;
        o_static_field(&s2, b_es, &pairs_field);
        o_ewald_full(&s, b_es, b_mu, b_ei, b_eic, old_mu, new_mu, b_rr, &out->iter_success, &contractions);
        out->polar_iterations = 0; /* not reported by the reference on this branch */
        out->dipole_rrms = 0;
    } else {
        o_static_field(&s, b_es, &pairs_field);
        out->polar_iterations = o_thole_iterative(&s, b_es, b_rk, b_mu, b_ei, b_eic, old_mu, new_mu, b_rr,
                                                  &out->iter_success, &contractions);
        for (i = 0; i < n; i++) { /* get_dipole_rrms(): polar.c:13-28 */
            if (isfinite(b_rr[i])) rr += b_rr[i];
        }
        out->dipole_rrms = n ? rr / n : 0;
    }
    for (i = 0; i < n; i++) { /* polar.c:108-116 */
        potential += b_mu[3 * i] * b_es[3 * i] + b_mu[3 * i + 1] * b_es[3 * i + 1] + b_mu[3 * i + 2] * b_es[3 * i + 2];
        if (P->polar_palmo)
            potential +=
                b_mu[3 * i] * b_eic[3 * i] + b_mu[3 * i + 1] * b_eic[3 * i + 1] + b_mu[3 * i + 2] * b_eic[3 * i + 2];
    }
    potential *= -0.5;
    out->polar = potential;
    if (cnt) {
        cnt->pairs_field = pairs_field;
        cnt->pairs_dipole = (double)npol * (npol > 0 ? npol - 1 : 0);
        cnt->contractions = contractions;
        cnt->n_polarizable = npol;
    }
    (void)p;
    if (!mu) free(b_mu);
    if (!ef_static) free(b_es);
    if (!ef_induced) free(b_ei);
    if (!ef_induced_change) free(b_eic);
    if (!dipole_rrms) free(b_rr);
    if (!rank_metric) free(b_rk);
    free(old_mu);
    free(new_mu);
    free(pol);
    return 0;
}

/* energy(): src/energy/energy.c:67-226 */
int mpmc_oracle_energy(const mpmc_b200_params *P, int n, const double *pos, const double *charge,
                       const double *polarizability, const double *epsilon, const double *sigma,
                       const double *mol_mass, const int *mol_id, const int *frozen, mpmc_b200_energies *out,
                       mpmc_b200_counters *cnt, double *mu, double *ef_static, double *ef_induced,
                       double *ef_induced_change, double *dipole_rrms, double *rank_metric, int nthreads) {
    osys_t s = {P, n, pos, charge, polarizability, epsilon, sigma, mol_mass, mol_id, frozen};
    int rc = o_check(P), i;
    if (rc) return rc;
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
    memset(out, 0, sizeof(*out));
    if (cnt) memset(cnt, 0, sizeof(*cnt));
    if (cnt) cnt->n_atoms = n;

    /* pair loops: per-row partials (added in row order), rows threaded */
    orow_t *rows = (orow_t *)calloc(n > 0 ? n : 1, sizeof(orow_t));
    orow_t tot;
    memset(&tot, 0, sizeof(tot));
    tot.min_r = O_MAXVALUE;
    if (nthreads == 1) {
        /* exact reference accumulation order: one running sum over all pairs */
        for (i = 0; i < n - 1; i++) o_pair_row(&s, i, &tot);
    } else {
#pragma omp parallel for schedule(dynamic, 8)
        for (i = 0; i < n - 1; i++) {
            rows[i].min_r = O_MAXVALUE;
            o_pair_row(&s, i, &rows[i]);
        }
        for (i = 0; i < n - 1; i++) {
            tot.rd_pair += rows[i].rd_pair;
            tot.lrc_pair += rows[i].lrc_pair;
            tot.es_real += rows[i].es_real;
            tot.n_tested += rows[i].n_tested;
            tot.n_lj += rows[i].n_lj;
            tot.n_es += rows[i].n_es;
            if (rows[i].min_r < tot.min_r) tot.min_r = rows[i].min_r;
        }
    }
    free(rows);
    out->min_intermolecular_r = tot.min_r;
    if (cnt) {
        cnt->pairs_tested = tot.n_tested;
        cnt->pairs_lj = tot.n_lj;
        cnt->pairs_es = tot.n_es;
    }

    if (P->cavity_autoreject_absolute && tot.min_r < P->cavity_autoreject_scale) { /* energy.c:99-102,187-191 */
        out->autorejected = 1;
        out->total = O_MAXVALUE;
        return 0;
    }

    if (!P->rd_only && P->polarization) { /* energy.c:108-129 */
        mpmc_b200_counters c2;
        memset(&c2, 0, sizeof(c2));
        rc = mpmc_oracle_polar(P, n, pos, charge, polarizability, epsilon, sigma, mol_mass, mol_id, frozen, out, &c2, mu,
                               ef_static, ef_induced, ef_induced_change, dipole_rrms, rank_metric, nthreads);
        if (rc) return rc;
        if (cnt) {
            cnt->pairs_field = c2.pairs_field;
            cnt->pairs_dipole = c2.pairs_dipole;
            cnt->contractions = c2.contractions;
            cnt->n_polarizable = c2.n_polarizable;
        }
    }

    { /* lj(): lj.c:256,270-273 */
        double lrc_self = 0;
        double crystal_self = 0; /* lj.c:264-267: every atom, frozen ones included; kept in rd_pair */
        if (P->rd_crystal)
            for (i = 0; i < n; i++) crystal_self += o_crystal_self(&s, i, o_rd_cutoff(P));
        tot.rd_pair += crystal_self;
        if (P->rd_lrc)
            for (i = 0; i < n; i++) lrc_self += o_lrc_self(&s, i, o_rd_cutoff(P));
        out->rd_pair = tot.rd_pair;
        out->rd_lrc = tot.lrc_pair + lrc_self;
        out->rd = (tot.rd_pair + tot.lrc_pair) + lrc_self;
        if (nthreads == 1) {
            /* The reference's own association (lj.c:256,264-273): every pair's (rd_energy + lrc), then every atom's
             * image self term, then every atom's LRC self term, all added one by one to the SAME running sum.  At
             * 10^4 atoms that sum carries ~1e-10 of rounding that the separated sums above do not (a constant ~1e-4 K
             * LRC term added 7.5e7 times to a 1e8 K sum rounds the same way every time); reproducing it bit for bit
             * is what ties the large reference-made pin (tests/golden/pin_co2_fluid_12k.npz) to this restatement. */
            double potential = tot.rd_running;
            if (P->rd_crystal)
                for (i = 0; i < n; i++) potential += o_crystal_self(&s, i, o_rd_cutoff(P));
            if (P->rd_lrc)
                for (i = 0; i < n; i++) potential += o_lrc_self(&s, i, o_rd_cutoff(P));
            out->rd = potential;
        }
    }
    if (!P->rd_only) { /* coulombic(): coulombic.c:22-39 */
        double nk = 0, nq = 0;
        out->es_real = tot.es_real;
        if (!P->es_wolf) { /* coulombic.c:27-28: wolf replaces real + reciprocal + self */
            out->es_recip = o_recip(&s, &nk, &nq);
            out->es_self = o_self(&s);
        }
        out->coulombic = out->es_real + out->es_recip + out->es_self;
        if (cnt) {
            cnt->kvectors = nk;
            cnt->recip_atoms = nq;
        }
    }
    out->total = out->rd + out->coulombic + out->polar; /* energy.c:194 */
    return 0;
}

int mpmc_oracle_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
