// pair_kernel.cu -- K1: fused cutoff Lennard-Jones (+Feynman-Hibbs) and Ewald real-space
// (+Feynman-Hibbs) pair-tile kernel, with the intramolecular erf term and the minimum
// intermolecular distance (cavity_autoreject_absolute) folded in.
//
// One CTA = one (I,J) tile pair, J >= I, of the upper triangle: thread t keeps atom I*128+t in
// registers and walks the 128 atoms of tile J staged in shared memory (every lane reads the same
// j => shared-memory broadcast).  No pair list exists; exclusions come from the molecule index
// and flag bits, mixed LJ parameters are formed in registers.
//
// Reference loops replaced (relative to the reference tree):
//   pair_exclusions()  src/energy/pairs.c:56-227     minimum_image()  src/energy/pairs.c:230-290
//   lj()               src/energy/lj.c:181-260       lj_fh_corr()     src/energy/lj.c:11-54
//   coulombic_real()   src/energy/coulombic.c:159-191  coulombic_real_FH()  src/energy/coulombic.c:115-146
//   coulombic_wolf()   src/energy/coulombic.c:269-308 (option `wolf`; replaces real + reciprocal + self)
//   lj() with rd_crystal   src/energy/lj.c:175-216 (image sums; crystal_lj_kernel below)
//   cavity_absolute_check()  src/energy/energy.c:49-64
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "kernels.h"

namespace mpmc {

struct PairConsts {
    double cutoff, alpha, a2, a3, a4;
    double fh2_pref, fh4_pref;  // M2A2*HBAR2/(24 KB T) and M2A4*HBAR4/(1152 KB2 T^2), before 1/mu_red powers
    double fh4_pref_es;         // coulombic.c:141 spells KB*KB instead of the KB2 constant (differs at 4.7e-10)
    double inv_sqrt_pi;
    // range tests on r^2: below in2 = (cutoff - 1e-9)^2 a pair is inside, above out2 = (cutoff + 1e-9)^2 outside; only in
    // the shell between them is the square root taken and the reference's own comparison evaluated
    double in2, out2;
    double inv_amu;  // 1 / AMU2KG: 1 / mu_red = (1/m_i + 1/m_j) / AMU2KG
    double one_m_alpha, a5x24, a7x16, c_ir4;  // 1 - a, 24 a^5, 16 a^7, 48 - 78 a: the collapsed sums of coulombic_real_FH()
    double ec[12];   // 1/q!: exp_neg_fast() (common.cuh) for the Gaussian of coulombic_real_FH() and of erfc_fast()
    double pc[17];   // erfc_fast(): exp(x^2) erfc(x) as a polynomial in (x - 4) / (x + 4), profiles/k1_erfc_coeffs.py
    // coulombic_wolf(): erf(alpha R)/R and 1/R^2 (coulombic.c:275-276,290); only read by the FH == 0 kernels
    double wolf_shift, wolf_ir2;
    int wolf;
    int no_lj;  // rd_crystal: the LJ term comes from crystal_lj_kernel, K1 keeps electrostatics + minimum distance
};

// Per-atom quantities of the pair terms that the reference re-derives for every pair: sqrt(epsilon) (Lorentz-Berthelot
// eps_ij = sqrt(eps_i eps_j), pairs.c:200-211, here se_i * se_j: one rounding more, no square root per pair) and the
// inverse molecular mass (Feynman-Hibbs reduced mass, lj.c:22-23: 1 / mu_red = (1/m_i + 1/m_j) / AMU2KG, no division per pair).
struct PairAtom {
    double eps, sig, se, imm;
};
template <int FH>
__device__ __forceinline__ PairAtom pair_atom(double eps, double sig, double mm) {
    PairAtom a;
    a.eps = eps;
    a.sig = sig;
    a.se = sqrt(fabs(eps));
    a.imm = FH ? 1.0 / mm : 0.0;
    return a;
}

__device__ __forceinline__ void mix_lj(int mixing, double ei, double si, double sei, double ej, double sj, double sej,
                                       double& eps, double& sigma, bool& attractive_only) {
    attractive_only = false;
    eps = 0.0;
    sigma = 0.0;
    switch (mixing) {
        case MPMC_MIX_WALDMAN_HAGLER: {  // pairs.c:85-103
            double si3 = si * si * si, si6 = si3 * si3, sj3 = sj * sj * sj, sj6 = sj3 * sj3;
            if (si < 0.0 || sj < 0.0) {
                attractive_only = true;
                sigma = pow(0.5 * (si6 + sj6), 1. / 6.);
            } else if (si == 0 || sj == 0) {
                eps = sei * sej;
            } else {
                sigma = pow(0.5 * (si6 + sj6), 1. / 6.);
                eps = sei * sej * 2.0 * si3 * sj3 / (si6 + sj6);
            }
        } break;
        case MPMC_MIX_HALGREN:  // pairs.c:104-114
            if (si > 0.0 && sj > 0.0) sigma = (si * si * si + sj * sj * sj) / (si * si + sj * sj);
            if (ei > 0.0 && ej > 0.0) {
                double t = sei + sej;
                eps = 4 * ei * ej / (t * t);
            }
            break;
        case MPMC_MIX_C6:  // pairs.c:194-199
            sigma = 0.5 * (si + sj);
            if (sigma != 0.0) {
                double sp = si + sj, sp3 = sp * sp * sp;
                eps = 64.0 * (sei * sej) * (si * si * si) * (sj * sj * sj) / (sp3 * sp3);
            }
            break;
        default:  // Lorentz-Berthelot, pairs.c:200-211
            if (si < 0.0 || sj < 0.0) {
                attractive_only = true;
                sigma = 0.5 * (fabs(si) + fabs(sj));
            } else if (si == 0 || sj == 0) {
                eps = sei * sej;
            } else {
                sigma = 0.5 * (si + sj);
                eps = sei * sej;
            }
    }
}

struct PairAcc {
    double rd, es, intra, minr2;  // minr2: smallest squared intermolecular distance (sqrt is monotone: taken once at the end)
    int n_tested, n_lj, n_es;
};
constexpr double kMinR2Init = 1.0e80;
__device__ __forceinline__ double min_r_from(double minr2) { return minr2 >= 1.0e79 ? kMAXVALUE : sqrt(minr2); }

// What is known about a pair before any expensive term: the minimum-image separation and which terms apply.
struct PairScreen {
    double r2, r, ir, qq;
    bool do_lj, do_es;
};

// pair_exclusions + minimum_image + the range tests of lj() and coulombic_real(): false when the pair adds nothing
// beyond what is accumulated here (the intramolecular erf term, the minimum intermolecular distance, the count)
template <bool ORTHO, int FH>
__device__ __forceinline__ bool pair_screen(const Cell& c, const PairConsts& k, int rd_only, double xi, double yi, double zi,
                                            double qi, int moli, int fli, bool lj_i, double xj, double yj, double zj,
                                            double qj, double ej, double sj, int molj, int fj, PairAcc& A, PairScreen& g) {
    // skip padding and frozen-frozen pairs (pairs.c:81, lj.c:193, coulombic.c:165)
    if (!((fli & fj) & F_VALID) || ((fli & fj) & F_FROZEN)) return false;
    double dx = xi - xj, dy = yi - yj, dz = zi - zj;
    g.qq = qi * qj;
    if (moli == molj) {
        // same molecule: rd and es excluded (pairs.c:61-64); charge-to-screen term with the
        // UN-imaged separation (coulombic.c:179-180)
        if (!rd_only && g.qq != 0.0 && !(FH == 0 && k.wolf)) {
            double r = sqrt(dx * dx + dy * dy + dz * dz);
            A.intra += g.qq * erf(k.alpha * r) / r;
        }
        return false;
    }
    min_image<ORTHO>(c, dx, dy, dz);
    g.r2 = dx * dx + dy * dy + dz * dz;
    A.n_tested++;
    A.minr2 = fmin(A.minr2, g.r2);
    // lj.c:191 keeps r - SMALL_dR < cutoff; coulombic.c:167 skips r > cutoff; coulombic_wolf() keeps r < R only
    // (coulombic.c:288).  Decided on r^2; the reference's comparisons on sqrt(r^2) only inside the 1e-9 A shell.
    bool in_lj, in_es;
    if (g.r2 < k.in2) {
        in_lj = in_es = true;
    } else if (g.r2 > k.out2) {
        in_lj = in_es = false;
    } else {
        const double rs = sqrt(g.r2);
        in_lj = rs - kSMALL_dR < k.cutoff;
        in_es = (FH == 0 && k.wolf) ? (rs < k.cutoff) : !(rs > k.cutoff);
    }
    g.do_lj = in_lj && lj_i && (ej != 0.0) && (sj != 0.0);
    g.do_es = !rd_only && in_es && (g.qq != 0.0);
    if (!(g.do_lj || g.do_es)) return false;
    // 1/r from the rsqrt seed + one third-order step (common.cuh), r = r^2 / r: no square root, no division
    g.ir = rsqrt_fast(g.r2);
    g.r = g.r2 * g.ir;
    if (g.r2 == 0.0) {  // coincident atoms of two molecules: 1/r = inf, and the terms follow as in the reference
        g.ir = 1.0 / g.r2;
        g.r = 0.0;
    }
    return true;
}

// erfc(x) and exp(-x^2) for 0 <= x <= 6 without the library's range splits: erfc(x) = exp(-x^2) P((x - 4) / (x + 4)) with a
// degree-16 polynomial (the mapped variable makes exp(x^2) erfc(x) nearly polynomial on the whole range; coefficients sum
// to 1 in magnitude, so Horner is well conditioned), the Gaussian from exp_neg_fast() -- which the Feynman-Hibbs terms
// need anyway -- and 1 / (x + 4) from the reciprocal seed + two Newton steps.  ~45 fp64 instructions, no branch; 7e-15
// relative against a 50-digit reference.  The library erfc() remains for x > 6 (ewald_alpha * cutoff is 3.5 by default).
__device__ __forceinline__ double erfc_fast(const PairConsts& k, double x, double& gauss) {
    gauss = exp_neg_fast(k, x * x);
    const double d = x + 4.0;
    double rc;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(rc) : "d"(d));
    double e = fma(-d, rc, 1.0);
    rc = fma(rc, e, rc);
    e = fma(-d, rc, 1.0);
    rc = fma(rc, e, rc);
    const double y = (x - 4.0) * rc;
    double p = k.pc[16];
#pragma unroll
    for (int q = 15; q >= 0; q--) p = fma(p, y, k.pc[q]);
    return gauss * p;
}

// lj() + lj_fh_corr() and coulombic_real() + coulombic_real_FH() / coulombic_wolf() of a screened pair
template <int FH>
__device__ __forceinline__ void pair_body(const PairConsts& k, int mixing, const PairAtom& pi, const PairAtom& pj,
                                          const PairScreen& g, PairAcc& A) {
    const double r = g.r, r2 = g.r2, qq = g.qq;
    const bool do_lj = g.do_lj, do_es = g.do_es;
    const double ir = g.ir;
    double inv_mured = 0.0;  // 1 / (AMU2KG m_i m_j / (m_i + m_j)), lj.c:22-23
    if (FH) inv_mured = (pi.imm + pj.imm) * k.inv_amu;

    if (do_lj) {
        double eps, sigma;
        bool att;
        mix_lj(mixing, pi.eps, pi.sig, pi.se, pj.eps, pj.sig, pj.se, eps, sigma, att);
        const double sor = fabs(sigma) * ir;
        double s6 = sor * sor * sor;
        s6 *= s6;
        // with A = eps (sigma/r)^12 and B = eps (sigma/r)^6 the derivative sums of lj_fh_corr() (lj.c:29-50) collapse:
        //   d2E + 2 dE / r                   = (528 A - 120 B) / r^2
        //   15 dE / r^3 + 4 d3E / r + d4E    = (87984 A - 6360 B) / r^4
        // (dE = -24 (2A - B)/r, d2E = 24 (26A - 7B)/r^2, d3E = -1344 (6A - B)/r^3, d4E = 12096 (10A - B)/r^4)
        const double B6 = eps * s6, A12 = att ? 0.0 : B6 * s6;
        double u = 4.0 * (A12 - B6);
        if (FH) {
            const double ir2 = ir * ir;
            u = fma((k.fh2_pref * inv_mured) * ir2, fma(528.0, A12, -120.0 * B6), u);
            if (FH >= 4)
                u = fma((k.fh4_pref * (inv_mured * inv_mured)) * (ir2 * ir2), fma(87984.0, A12, -6360.0 * B6), u);
        }
        A.rd += u;
        A.n_lj++;
    }
    if (FH == 0 && k.wolf) {
        if (do_es) {  // coulombic.c:289-290
            A.es += qq * (ir - k.wolf_shift - k.wolf_ir2 * (k.cutoff - r));
            A.n_es++;
        }
    } else if (do_es) {
        const double x = k.alpha * r;
        double erfc_term, gaussian_term = 0.0;
        if (x <= 6.0) {
            erfc_term = erfc_fast(k, x, gaussian_term);
        } else {
            erfc_term = erfc(x);
            if (FH) gaussian_term = x < 26.0 ? exp_neg_fast(k, x * x) : 0.0;  // exp(-alpha^2 r^2), valid below 700
        }
        double u = qq * erfc_term * ir;
        if (FH) {
            // coulombic_real_FH() (coulombic.c:132-141) with G = exp(-a^2 r^2) / sqrt(pi), E = erfc(a r); its derivative sums
            // collapse (the erfc terms of the second-order sum cancel identically, those of the fourth-order sum to -15 E / r^5):
            //   d2u + 2 du / r                 = 4 G (a^3 + (1 - a) / r^2)
            //   15 du / r^3 + 4 d3u / r + d4u  = G (16 a^7 r^2 - 24 a^5 + (48 - 78 a) / r^4) - 15 E / r^5
            // (the reference's own mix of powers of alpha is kept; the corrections carry no charge product, as there)
            const double ir2 = ir * ir;
            const double gg = gaussian_term * k.inv_sqrt_pi;
            u = fma(k.fh2_pref * inv_mured, 4.0 * gg * fma(k.one_m_alpha, ir2, k.a3), u);
            if (FH >= 4) {
                const double ir4 = ir2 * ir2;
                const double poly = fma(k.a7x16, r2, fma(k.c_ir4, ir4, -k.a5x24));
                u = fma(k.fh4_pref_es * (inv_mured * inv_mured), fma(gg, poly, -15.0 * erfc_term * (ir4 * ir)), u);
            }
        }
        A.es += u;
        A.n_es++;
    }
}

// all terms of one pair (i in registers, j given): pair_exclusions + minimum_image + lj + coulombic_real
template <bool ORTHO, int FH>
__device__ __forceinline__ void pair_terms(const Cell& c, const PairConsts& k, int mixing, int rd_only, double xi, double yi,
                                           double zi, double qi, const PairAtom& pi, int moli, int fli, bool lj_i, double xj,
                                           double yj, double zj, double qj, const PairAtom& pj, int molj, int fj, PairAcc& A) {
    PairScreen g;
    if (pair_screen<ORTHO, FH>(c, k, rd_only, xi, yi, zi, qi, moli, fli, lj_i, xj, yj, zj, qj, pj.eps, pj.sig, molj, fj, A, g))
        pair_body<FH>(k, mixing, pi, pj, g, A);
}

template <bool ORTHO, int FH>
__global__ void __launch_bounds__(TILE) pair_energy_kernel(Cell c, AtomsDev a, PairConsts k, int mixing, int rd_only,
                                                           const int2* __restrict__ tiles, double* __restrict__ partial,
                                                           int stride) {
    __shared__ double sx[TILE], sy[TILE], sz[TILE], sq[TILE];
    __shared__ PairAtom sp[TILE];
    __shared__ int smol[TILE], sfl[TILE];
    __shared__ double red[32];

    const int2 t = tiles[blockIdx.x];
    const int tid = threadIdx.x;
    const int i = t.x * TILE + tid;
    {
        const int j = t.y * TILE + tid;
        sx[tid] = a.x[j];
        sy[tid] = a.y[j];
        sz[tid] = a.z[j];
        sq[tid] = a.q[j];
        sp[tid] = pair_atom<FH>(a.eps[j], a.sig[j], a.molmass[j]);
        smol[tid] = a.mol[j];
        sfl[tid] = a.flags[j];
    }
    const double xi = a.x[i], yi = a.y[i], zi = a.z[i], qi = a.q[i], ei = a.eps[i], si = a.sig[i], mi = a.molmass[i];
    const int moli = a.mol[i], fli = a.flags[i];
    __syncthreads();

    PairAcc A = {0.0, 0.0, 0.0, kMinR2Init, 0, 0, 0};
    const bool lj_i = (ei != 0.0) && (si != 0.0) && !k.no_lj;
    const PairAtom pi = pair_atom<FH>(ei, si, mi);
    const int j0 = (t.x == t.y) ? tid + 1 : 0;

    for (int jj = j0; jj < TILE; jj++) {
        pair_terms<ORTHO, FH>(c, k, mixing, rd_only, xi, yi, zi, qi, pi, moli, fli, lj_i, sx[jj], sy[jj], sz[jj], sq[jj], sp[jj],
                              smol[jj], sfl[jj], A);
    }
    const double rd = A.rd, es = A.es, intra = A.intra, minr2 = A.minr2;
    const int n_tested = A.n_tested, n_lj = A.n_lj, n_es = A.n_es;

    // fixed-order block reduction, one partial row per tile pair
    double v;
    v = block_reduce<0>(rd, red);
    if (tid == 0) partial[0 * stride + blockIdx.x] = v;
    v = block_reduce<0>(es, red);
    if (tid == 0) partial[1 * stride + blockIdx.x] = v;
    v = block_reduce<0>(intra, red);
    if (tid == 0) partial[2 * stride + blockIdx.x] = v;
    v = block_reduce<1>(minr2, red);
    if (tid == 0) partial[3 * stride + blockIdx.x] = min_r_from(v);
    v = block_reduce<0>((double)n_tested, red);
    if (tid == 0) partial[4 * stride + blockIdx.x] = v;
    v = block_reduce<0>((double)n_lj, red);
    if (tid == 0) partial[5 * stride + blockIdx.x] = v;
    v = block_reduce<0>((double)n_es, red);
    if (tid == 0) partial[6 * stride + blockIdx.x] = v;
}

// ------------------------------------------------------------------------------------------------
// Cluster form of K1 (the default): atoms in spatial order, one WARP per 32-atom i-cluster.  Lane l keeps
// atom ci*32+l in registers.  Candidate j-clusters are (ci + k) mod nclus, k = 0..nclus/2, so every unordered
// cluster pair is met exactly once and every warp has the same number of candidates; a batch of 32 candidates
// is tested against the cutoff in one step (lane l tests candidate l: bounding-box gap, cluster_gap2) and only
// the survivors are staged (per-warp shared memory, 64-byte records) and walked.  With an 8 A cutoff in a
// 128 x 128 x 43 A box this visits ~3 % of the pairs the reference tests -- the rest are proven to be
// outside the cutoff, where lj() and coulombic_real() add nothing (lj.c:191, coulombic.c:167).
// No CTA-level synchronisation; one partial column per (slice, cluster), reduced in fixed order.
// ------------------------------------------------------------------------------------------------
struct __align__(16) PairRec {
    double x, y, z, q;
    PairAtom p;  // eps, sig, sqrt(eps), 1 / molecule mass: the last two formed once per staged atom, not once per pair
    int mol, fl;
};

// COMPACT (the default): a cluster pair is screened first -- every lane tests its atom against the 32 staged ones
// (geometry, range, exclusions: ~35 instructions) and keeps a bit per pair that has an LJ or erfc term -- then the
// surviving pairs of the whole cluster pair are dealt out densely, 32 at a time, for the expensive part (erfc, exp,
// the Feynman-Hibbs polynomials: ~200 instructions).  Walking j for all lanes at once kept 16 of 32 lanes busy (ncu,
// profiles/r01_k1_cluster_ncu_details.csv): a lane whose pair is outside the cutoff waited for its neighbours' bodies.
// K1 only produces sums, so it does not matter which lane evaluates which pair; the order stays a pure function of
// the input (bit-reproducible).
template <bool ORTHO, int FH, bool COMPACT>
__global__ void __launch_bounds__(TILE) pair_cluster_kernel(Cell c, SubsetDev a, PairConsts k, BoxLen box, int mixing,
                                                            int rd_only, int clus_begin, int clus_end, int kper, double cull2,
                                                            double* __restrict__ partial, int stride) {
    __shared__ PairRec rec[4][32];
    __shared__ PairRec irec[COMPACT ? 4 : 1][32];
    __shared__ unsigned short plist[COMPACT ? 4 : 1][1024];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int ci = clus_begin + blockIdx.x * 4 + w;
    if (ci >= clus_end) return;  // whole warp; no block-level barrier is used below
    const int nclus = a.n_pad >> 5;
    const int i = ci * 32 + lane;
    const double xi = a.x[i], yi = a.y[i], zi = a.z[i], qi = a.q[i], ei = a.eps[i], si = a.sig[i], mi = a.molmass[i];
    const int moli = a.mol[i], fli = a.flags[i];
    const bool lj_i = (ei != 0.0) && (si != 0.0) && !k.no_lj;
    const PairAtom pi = pair_atom<FH>(ei, si, mi);
    PairAcc A = {0.0, 0.0, 0.0, kMinR2Init, 0, 0, 0};
    if (COMPACT) {
        PairRec r;
        r.x = xi, r.y = yi, r.z = zi, r.q = qi, r.p = pi, r.mol = moli, r.fl = fli;
        irec[w][lane] = r;
    }

    const int K = nclus / 2;
    const int k0 = blockIdx.y * kper, k1 = min(k0 + kper, K + 1);
    for (int kb = k0; kb < k1; kb += 32) {
        const int kk = kb + lane;
        bool pass = kk < k1 && !(kk > 0 && 2 * kk == nclus && ci >= kk);  // even nclus: half-way offset for ci < nclus/2 only
        int cj = ci + kk;
        if (cj >= nclus) cj -= nclus;
        if (pass && cull2 > 0.0) pass = cluster_gap2(box, a.cbox, nclus, ci, a.cbox, nclus, cj) <= cull2;
        // a cluster pair that lies entirely inside the cutoff has nothing to screen out by range: walk it (the dense
        // form only pays where a good part of the 1,024 pairs drops out: 8 A cutoff in a MOF, not L/2 in a fluid)
        const bool inside = COMPACT && pass && cull2 > 0.0 &&
                            cluster_span2(box, a.cbox, nclus, ci, a.cbox, nclus, cj) <= k.cutoff * k.cutoff;
        unsigned mask = __ballot_sync(0xffffffffu, pass);
        const unsigned inside_mask = __ballot_sync(0xffffffffu, inside);
        while (mask) {
            const int b = __ffs(mask) - 1;
            mask &= mask - 1;
            const int kk2 = kb + b;
            int cj2 = ci + kk2;
            if (cj2 >= nclus) cj2 -= nclus;
            __syncwarp();
            {
                const int j = cj2 * 32 + lane;
                PairRec r;
                r.x = a.x[j];
                r.y = a.y[j];
                r.z = a.z[j];
                r.q = a.q[j];
                r.p = pair_atom<FH>(a.eps[j], a.sig[j], a.molmass[j]);
                r.mol = a.mol[j];
                r.fl = a.flags[j];
                rec[w][lane] = r;
            }
            __syncwarp();
            const int s0 = (kk2 == 0) ? lane + 1 : 0;  // inside one cluster: j > i
            if (!COMPACT || ((inside_mask >> b) & 1u)) {
                for (int s = s0; s < 32; s++) {
                    const PairRec& r = rec[w][s];
                    pair_terms<ORTHO, FH>(c, k, mixing, rd_only, xi, yi, zi, qi, pi, moli, fli, lj_i, r.x, r.y, r.z, r.q, r.p,
                                          r.mol, r.fl, A);
                }
            } else {
                // screen: one bit per staged atom whose pair with this lane's atom has a body
                unsigned bits = 0;
                for (int s = s0; s < 32; s++) {
                    const PairRec& r = rec[w][s];
                    PairScreen g;
                    if (pair_screen<ORTHO, FH>(c, k, rd_only, xi, yi, zi, qi, moli, fli, lj_i, r.x, r.y, r.z, r.q, r.p.eps, r.p.sig,
                                               r.mol, r.fl, A, g))
                        bits |= 1u << s;
                }
                // exclusive prefix of the lanes' counts -> where this lane's pairs go in the list
                const int cnt = __popc(bits);
                int incl = cnt;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const int v = __shfl_up_sync(0xffffffffu, incl, o);
                    if (lane >= o) incl += v;
                }
                const int total = __shfl_sync(0xffffffffu, incl, 31);
                int pos = incl - cnt;
                while (bits) {
                    const int s = __ffs(bits) - 1;
                    bits &= bits - 1;
                    plist[w][pos++] = (unsigned short)((lane << 5) | s);
                }
                __syncwarp();
                // bodies, densely: lane l takes pairs l, l + 32, ...
                PairAcc scratch = {0.0, 0.0, 0.0, kMinR2Init, 0, 0, 0};  // the screen's side sums were taken above
                for (int t = lane; t < total; t += 32) {
                    const int pr = plist[w][t];
                    const PairRec& ri = irec[w][pr >> 5];
                    const PairRec& rj = rec[w][pr & 31];
                    const bool lj_a = (ri.p.eps != 0.0) && (ri.p.sig != 0.0) && !k.no_lj;
                    PairScreen g;
                    if (pair_screen<ORTHO, FH>(c, k, rd_only, ri.x, ri.y, ri.z, ri.q, ri.mol, ri.fl, lj_a, rj.x, rj.y, rj.z, rj.q,
                                               rj.p.eps, rj.p.sig, rj.mol, rj.fl, scratch, g))
                        pair_body<FH>(k, mixing, ri.p, rj.p, g, A);
                }
            }
        }
    }
    // same-molecule pairs are handled by intra_erf_kernel (list order): drop what pair_terms collected for them
    const double rd = warp_sum(A.rd), es = warp_sum(A.es), minr = min_r_from(warp_min(A.minr2));
    const double nlj = warp_sum((double)A.n_lj), nes = warp_sum((double)A.n_es);
    if (lane == 0) {
        const int col = blockIdx.y * (clus_end - clus_begin) + (ci - clus_begin);
        partial[0 * stride + col] = rd;
        partial[1 * stride + col] = es;
        partial[2 * stride + col] = 0.0;
        partial[3 * stride + col] = minr;
        partial[4 * stride + col] = 0.0;
        partial[5 * stride + col] = nlj;
        partial[6 * stride + col] = nes;
    }
}

__global__ void __launch_bounds__(256) intra_erf_kernel(AtomsDev a, const int* __restrict__ follow, double alpha,
                                                        int row_begin, int row_end, double* __restrict__ part) {
    __shared__ double red[32];
    const int i = row_begin + blockIdx.x * blockDim.x + threadIdx.x;
    double acc = 0.0;
    if (i < row_end) {
        const int nf = follow[i];
        if (nf > 0) {
            const double xi = a.x[i], yi = a.y[i], zi = a.z[i], qi = a.q[i];
            const int fi = a.flags[i];
            for (int j = i + 1; j <= i + nf; j++) {
                if ((fi & a.flags[j]) & F_FROZEN) continue;  // coulombic.c:165
                const double qq = qi * a.q[j];
                if (qq == 0.0) continue;
                const double dx = xi - a.x[j], dy = yi - a.y[j], dz = zi - a.z[j];
                const double r = sqrt(dx * dx + dy * dy + dz * dz);  // UN-imaged (coulombic.c:180)
                acc += qq * erf(alpha * r) / r;
            }
        }
    }
    acc = block_reduce<0>(acc, red);
    if (threadIdx.x == 0) part[blockIdx.x] = acc;
}

__global__ void __launch_bounds__(1024) intra_sum_kernel(const double* __restrict__ in, int n, double* __restrict__ out) {
    __shared__ double red[32];
    double acc = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) acc += in[i];
    acc = block_reduce<0>(acc, red);
    if (threadIdx.x == 0) out[0] = acc;
}

// One CTA sums the per-tile partials in a fixed order (deterministic run to run).
__global__ void __launch_bounds__(1024) pair_reduce_kernel(const double* __restrict__ partial, int ntiles, int stride,
                                                           double* __restrict__ out) {
    __shared__ double red[32];
    for (int q = 0; q < 7; q++) {
        double acc = (q == 3) ? kMAXVALUE : 0.0;
        for (int t = threadIdx.x; t < ntiles; t += blockDim.x) {
            double v = partial[q * stride + t];
            acc = (q == 3) ? fmin(acc, v) : acc + v;
        }
        double v = (q == 3) ? block_reduce<1>(acc, red) : block_reduce<0>(acc, red);
        if (threadIdx.x == 0) out[q] = v;
    }
}

static PairConsts make_pair_consts(const Cell& c, const PairOpts& o);

void launch_pair_energy_clusters(const Cell& c, bool ortho, const SubsetDev& a, const PairOpts& o, int clus_begin,
                                 int clus_end, int kslices, double cull, double* partial, int stride, cudaStream_t s) {
    const int nown = clus_end - clus_begin;
    if (nown <= 0) return;
    const PairConsts k = make_pair_consts(c, o);
    const int nclus = a.n_pad / 32, K = nclus / 2;
    if (kslices < 1) kslices = 1;
    int kper = (K + 1 + kslices - 1) / kslices;
    if (kper > 32) kper = ((kper + 31) / 32) * 32;  // whole candidate batches per slice; small systems: finer slices
                                                    // (a warp's serial walk over its cluster pairs is the latency)
    kslices = (K + 1 + kper - 1) / kper;
    const double cull2 = (ortho && cull > 0.0) ? cull * cull : 0.0;  // the gap bound is orthorhombic only
    const BoxLen box = box_len(c);
    dim3 grid((nown + 3) / 4, kslices);
    // Screening pays where a good part of a passing cluster pair's 1,024 atom pairs drops out, i.e. where the cutoff is
    // not much larger than a cluster (8 A in a MOF: K1 0.83 -> 0.67 ms at 63k atoms); with the reference's default
    // cutoff of half the box most cluster pairs are entirely inside it and the two-phase form costs 3-13 % instead
    // (profiles/r02_k1_compaction.md).  MPMC_B200_K1=walk|compact forces either form (A/B).
    static const int forced = [] {
        const char* v = getenv("MPMC_B200_K1");
        return !v ? 0 : !strcmp(v, "walk") ? 1 : !strcmp(v, "compact") ? 2 : 0;
    }();
    const double cluster_edge = cbrt(32.0 * c.volume / (a.n > 0 ? a.n : 1));
    const bool compact = forced == 2 || (forced == 0 && cull2 > 0.0 && c.cutoff < 2.5 * cluster_edge);
#define LAUNCH(ORTHO, FH)                                                                                                  \
    if (compact)                                                                                                           \
        pair_cluster_kernel<ORTHO, FH, true><<<grid, TILE, 0, s>>>(c, a, k, box, o.mixing, o.rd_only, clus_begin, clus_end,  \
                                                                   kper, cull2, partial, stride);                          \
    else                                                                                                                   \
        pair_cluster_kernel<ORTHO, FH, false><<<grid, TILE, 0, s>>>(c, a, k, box, o.mixing, o.rd_only, clus_begin, clus_end, \
                                                                    kper, cull2, partial, stride)
    if (ortho) {
        if (o.fh_order == 0)
            LAUNCH(true, 0);
        else if (o.fh_order == 2)
            LAUNCH(true, 2);
        else
            LAUNCH(true, 4);
    } else {
        if (o.fh_order == 0)
            LAUNCH(false, 0);
        else if (o.fh_order == 2)
            LAUNCH(false, 2);
        else
            LAUNCH(false, 4);
    }
#undef LAUNCH
}

int pair_cluster_columns(int n_pad_sorted, int clus_begin, int clus_end, int kslices) {
    const int nclus = n_pad_sorted / 32, K = nclus / 2;
    if (kslices < 1) kslices = 1;
    int kper = (K + 1 + kslices - 1) / kslices;
    if (kper > 32) kper = ((kper + 31) / 32) * 32;
    kslices = (K + 1 + kper - 1) / kper;
    return kslices * (clus_end - clus_begin);
}

void launch_intra_erf(const AtomsDev& a, const int* follow, double ewald_alpha, int row_begin, int row_end, double* part,
                      double* sum_out, cudaStream_t s) {
    const int rows = row_end - row_begin;
    const int nb = (rows + 255) / 256;
    if (rows > 0) intra_erf_kernel<<<nb, 256, 0, s>>>(a, follow, ewald_alpha, row_begin, row_end, part);
    intra_sum_kernel<<<1, 1024, 0, s>>>(part, rows > 0 ? nb : 0, sum_out);
}

static PairConsts make_pair_consts(const Cell& c, const PairOpts& o) {
    PairConsts k;
    k.cutoff = c.cutoff;
    k.in2 = (c.cutoff - 1e-9) * (c.cutoff - 1e-9);
    k.out2 = (c.cutoff + 1e-9) * (c.cutoff + 1e-9);
    k.inv_amu = 1.0 / kAMU2KG;
    fill_exp_coeffs(k.ec);
    static const double erfc_poly[17] = {  // profiles/k1_erfc_coeffs.py: max relative error 3.5e-15 on 0 <= x <= 6
        0.13699945762506133,     -0.25906804876015949,    0.21871967891831384,    -0.16425781670012546,
        0.10896317738612751,     -0.06310781547805365,    0.03129905758162637,    -0.012843947793199894,
        0.004060233939324147,    -0.00079919577948827682, -2.1624460041698876e-05, 8.0102508503569764e-05,
        -1.9647828005712331e-05, 2.3461777077804792e-06,  6.6863398646632111e-06,  1.4399022917082169e-06,
        4.4575205502236814e-08};
    for (int q = 0; q < 17; q++) k.pc[q] = erfc_poly[q];
    k.alpha = o.ewald_alpha;
    k.a2 = k.alpha * k.alpha;
    k.a3 = k.a2 * k.alpha;
    k.a4 = k.a3 * k.alpha;
    k.one_m_alpha = 1.0 - k.alpha;
    k.a5x24 = 24.0 * (k.a3 * k.a2);
    k.a7x16 = 16.0 * k.a3 * k.a4;
    k.c_ir4 = 48.0 - 78.0 * k.alpha;
    k.inv_sqrt_pi = 1.0 / sqrt(M_PI);
    k.fh2_pref = kM2A2 * (kHBAR2 / (24.0 * kKB * o.temperature));
    k.fh4_pref = kM2A4 * (kHBAR4 / (1152.0 * kKB2 * o.temperature * o.temperature));
    k.fh4_pref_es = kM2A4 * (kHBAR4 / (1152.0 * (kKB * kKB * o.temperature * o.temperature)));
    k.wolf = o.es_wolf && o.fh_order == 0;
    k.no_lj = o.no_lj;
    k.wolf_shift = erf(k.alpha * k.cutoff) / k.cutoff;
    k.wolf_ir2 = (1.0 / k.cutoff) * (1.0 / k.cutoff);
    return k;
}

void launch_pair_energy(const Cell& c, bool ortho, const AtomsDev& a, const PairOpts& o, const int2* tiles, int ntiles,
                        double* partial, int stride, cudaStream_t s) {
    if (ntiles <= 0) return;
    const PairConsts k = make_pair_consts(c, o);
#define LAUNCH(ORTHO, FH) \
    pair_energy_kernel<ORTHO, FH><<<ntiles, TILE, 0, s>>>(c, a, k, o.mixing, o.rd_only, tiles, partial, stride)
    if (ortho) {
        if (o.fh_order == 0)
            LAUNCH(true, 0);
        else if (o.fh_order == 2)
            LAUNCH(true, 2);
        else
            LAUNCH(true, 4);
    } else {
        if (o.fh_order == 0)
            LAUNCH(false, 0);
        else if (o.fh_order == 2)
            LAUNCH(false, 2);
        else
            LAUNCH(false, 4);
    }
#undef LAUNCH
}

// ------------------------------------------------------------------------------------------------
// rd_crystal (lj.c:175-216): the LJ term of EVERY pair that is not frozen-frozen -- same-molecule pairs included --
// summed over the periodic images l in (-(order-1)..order-1)^3 of the UN-imaged separation, inside the wider cutoff
// 2*cutoff*(order-0.5); same-molecule pairs skip their home image (lj.c:203).  All-pairs tiles of the list-ordered
// atoms (no cutoff culling applies: the wider cutoff exceeds the minimum-image range), the image vectors
// B^T l staged in shared memory in the reference's loop order, i atom in registers, j tile in shared memory.
// The Feynman-Hibbs correction uses the summed terms with the minimum-image distance, as lj_fh_corr() does
// (lj.c:247-248).  Fixed-order block sums, one partial per tile pair.
constexpr int CRYSTAL_MAX_IMG = 343;  // rd_crystal_order <= 4

template <bool ORTHO, int FH>
__global__ void __launch_bounds__(TILE) crystal_lj_kernel(Cell c, AtomsDev a, PairConsts k, int mixing, double rd_cutoff,
                                                          const double* __restrict__ img, int nimg,
                                                          const int2* __restrict__ tiles, double* __restrict__ partial,
                                                          int stride) {
    __shared__ double sx[TILE], sy[TILE], sz[TILE], se[TILE], ss[TILE], sm[TILE];
    __shared__ int smol[TILE], sfl[TILE];
    __shared__ double ix[CRYSTAL_MAX_IMG], iy[CRYSTAL_MAX_IMG], iz[CRYSTAL_MAX_IMG];
    __shared__ double red[32];

    const int2 t = tiles[blockIdx.x];
    const int tid = threadIdx.x;
    const int i = t.x * TILE + tid;
    {
        const int j = t.y * TILE + tid;
        sx[tid] = a.x[j];
        sy[tid] = a.y[j];
        sz[tid] = a.z[j];
        se[tid] = a.eps[j];
        ss[tid] = a.sig[j];
        sm[tid] = a.molmass[j];
        smol[tid] = a.mol[j];
        sfl[tid] = a.flags[j];
    }
    for (int m = tid; m < nimg; m += TILE) {
        ix[m] = img[3 * m + 0];
        iy[m] = img[3 * m + 1];
        iz[m] = img[3 * m + 2];
    }
    const double xi = a.x[i], yi = a.y[i], zi = a.z[i], ei = a.eps[i], si = a.sig[i], mi = a.molmass[i];
    const int moli = a.mol[i], fli = a.flags[i];
    __syncthreads();

    const int home = (nimg - 1) / 2;  // l = (0,0,0) sits in the middle of the symmetric loop nest
    const double cut2_lo = rd_cutoff * rd_cutoff * (1.0 - 1e-12), cut2_hi = rd_cutoff * rd_cutoff * (1.0 + 1e-12);
    const bool lj_i = (ei != 0.0) && (si != 0.0);
    const int j0 = (t.x == t.y) ? tid + 1 : 0;
    double rd = 0.0;
    int n_lj = 0;
    for (int jj = j0; jj < TILE; jj++) {
        const int fj = sfl[jj];
        if (!((fli & fj) & F_VALID) || ((fli & fj) & F_FROZEN)) continue;  // padding; frozen pairs (lj.c:193)
        const double ej = se[jj], sj = ss[jj];
        // a pair with a zero eps or sigma on either atom mixes to eps = 0 or sigma = 0 under every rule
        // (pairs.c:84-211) and adds exactly 0 (also through lj_fh_corr): skipped
        if (!(lj_i && ej != 0.0 && sj != 0.0)) continue;
        const double dx = xi - sx[jj], dy = yi - sy[jj], dz = zi - sz[jj];
        double mx = dx, my = dy, mz = dz;
        min_image<ORTHO>(c, mx, my, mz);
        const double rimg = sqrt(mx * mx + my * my + mz * mz);
        if (!(rimg - kSMALL_dR < rd_cutoff)) continue;  // lj.c:191
        double eps, sigma;
        bool att;
        mix_lj(mixing, ei, si, sqrt(fabs(ei)), ej, sj, sqrt(fabs(ej)), eps, sigma, att);  // once per pair, images inside
        const double sig2 = sigma * sigma;
        const bool same = (moli == smol[jj]);
        double s6 = 0.0, s12 = 0.0;
        for (int m = 0; m < nimg; m++) {
            if (same && m == home) continue;  // lj.c:203
            const double ax = ix[m] + dx, ay = iy[m] + dy, az = iz[m] + dz;
            const double r2 = ax * ax + ay * ay + az * az;
            // lj.c:213 `r > cutoff` without the square root except in a 1e-12-wide shell around the cutoff
            if (r2 > cut2_hi) continue;
            if (r2 >= cut2_lo && sqrt(r2) > rd_cutoff) continue;
            const double y = rsqrt_fast(r2);  // (|sigma| / r)^6 = (sigma^2 / r^2)^3: no sqrt, no division
            const double t2 = sig2 * (y * y);
            const double t6 = t2 * t2 * t2;
            s6 += t6;
            s12 = fma(t6, t6, s12);
        }
        const double term6 = s6, term12 = att ? 0.0 : s12;
        double u = 4.0 * eps * (term12 - term6);
        if (FH) {  // lj.c:29-50 with pair->rimg
            const double ir = 1.0 / rimg, ir2 = ir * ir, ir3 = ir2 * ir, ir4 = ir3 * ir;
            const double mured = kAMU2KG * mi * sm[jj] / (mi + sm[jj]);
            const double dE = -24.0 * eps * (2.0 * term12 - term6) * ir;
            const double d2E = 24.0 * eps * (26.0 * term12 - 7.0 * term6) * ir2;
            u += (k.fh2_pref / mured) * (d2E + 2.0 * dE * ir);
            if (FH >= 4) {
                const double d3E = -1344.0 * eps * (6.0 * term12 - term6) * ir3;
                const double d4E = 12096.0 * eps * (10.0 * term12 - term6) * ir4;
                u += (k.fh4_pref / (mured * mured)) * (15.0 * dE * ir3 + 4.0 * d3E * ir + d4E);
            }
        }
        rd += u;
        n_lj++;
    }
    double v = block_reduce<0>(rd, red);
    if (tid == 0) partial[0 * stride + blockIdx.x] = v;
    v = block_reduce<0>((double)n_lj, red);
    if (tid == 0) partial[1 * stride + blockIdx.x] = v;
}

// One CTA sums the two partial rows of crystal_lj_kernel in a fixed order: out[0] = energy, out[1] = pair count.
__global__ void __launch_bounds__(1024) crystal_reduce_kernel(const double* __restrict__ partial, int ntiles, int stride,
                                                              double* __restrict__ out) {
    __shared__ double red[32];
    for (int q = 0; q < 2; q++) {
        double acc = 0.0;
        for (int t = threadIdx.x; t < ntiles; t += blockDim.x) acc += partial[q * stride + t];
        const double v = block_reduce<0>(acc, red);
        if (threadIdx.x == 0) out[q] = v;
    }
}

int crystal_max_images() { return CRYSTAL_MAX_IMG; }

void launch_crystal_lj(const Cell& c, bool ortho, const AtomsDev& a, const PairOpts& o, double rd_cutoff, const double* img,
                       int nimg, const int2* tiles, int ntiles, double* partial, int stride, double* out, cudaStream_t s) {
    const PairConsts k = make_pair_consts(c, o);
    if (ntiles > 0) {
#define LAUNCH(ORTHO, FH) \
    crystal_lj_kernel<ORTHO, FH><<<ntiles, TILE, 0, s>>>(c, a, k, o.mixing, rd_cutoff, img, nimg, tiles, partial, stride)
        if (ortho) {
            if (o.fh_order == 0)
                LAUNCH(true, 0);
            else if (o.fh_order == 2)
                LAUNCH(true, 2);
            else
                LAUNCH(true, 4);
        } else {
            if (o.fh_order == 0)
                LAUNCH(false, 0);
            else if (o.fh_order == 2)
                LAUNCH(false, 2);
            else
                LAUNCH(false, 4);
        }
#undef LAUNCH
    }
    crystal_reduce_kernel<<<1, 1024, 0, s>>>(partial, ntiles > 0 ? ntiles : 0, stride, out);
}

// ------------------------------------------------------------------------------------------------
// Incremental evaluation (opt-in): after a move of a few atoms only the pairs that contain a moved atom
// change.  Thread j (list order; the arrays hold the NEW positions) walks the moved atoms m: pair terms with m at its
// new position minus pair terms with m at its old one -- pair_terms() itself, so exclusions, mixing, FH, the
// intramolecular erf term and the Wolf form are those of K1.  A pair of two moved atoms is taken once (at the smaller
// moved index) with both atoms old / both new.  Differences are formed per pair before any summation.
template <bool ORTHO, int FH>
__global__ void __launch_bounds__(TILE) pair_delta_kernel(Cell c, AtomsDev a, PairConsts k, int mixing, int rd_only,
                                                          const MovedRec* __restrict__ moved, int nmoved,
                                                          double* __restrict__ partial, int stride) {
    __shared__ MovedRec smv[INC_MAX_MOVED];
    __shared__ double red[32];
    const int tid = threadIdx.x;
    for (int m = tid; m < nmoved; m += TILE) smv[m] = moved[m];
    __syncthreads();
    const int j = blockIdx.x * TILE + tid;
    const double xj = a.x[j], yj = a.y[j], zj = a.z[j], qj = a.q[j], ej = a.eps[j], sj = a.sig[j], mj = a.molmass[j];
    const int molj = a.mol[j], fj = a.flags[j];
    const PairAtom pj = pair_atom<FH>(ej, sj, mj);
    int mine = -1;  // index of j in the moved list
    for (int m = 0; m < nmoved; m++)
        if (smv[m].idx == j) mine = m;
    const double xjo = mine >= 0 ? smv[mine].xo : xj, yjo = mine >= 0 ? smv[mine].yo : yj, zjo = mine >= 0 ? smv[mine].zo : zj;
    double d_rd = 0.0, d_es = 0.0, d_intra = 0.0;
    int d_nlj = 0, d_nes = 0;
    for (int m = 0; m < nmoved; m++) {
        if (mine >= 0 && mine <= m) continue;  // j itself, or a moved pair already taken at the smaller index
        const MovedRec& r = smv[m];
        const bool lj_i = (r.eps != 0.0) && (r.sig != 0.0) && !k.no_lj;
        PairAcc An = {0.0, 0.0, 0.0, kMinR2Init, 0, 0, 0}, Ao = {0.0, 0.0, 0.0, kMinR2Init, 0, 0, 0};
        const PairAtom pm = pair_atom<FH>(r.eps, r.sig, r.mm);
        pair_terms<ORTHO, FH>(c, k, mixing, rd_only, r.xn, r.yn, r.zn, r.q, pm, r.mol, r.fl, lj_i, xj, yj, zj, qj, pj, molj, fj, An);
        pair_terms<ORTHO, FH>(c, k, mixing, rd_only, r.xo, r.yo, r.zo, r.q, pm, r.mol, r.fl, lj_i, xjo, yjo, zjo, qj, pj, molj, fj,
                              Ao);
        d_rd += An.rd - Ao.rd;
        d_es += An.es - Ao.es;
        d_intra += An.intra - Ao.intra;
        d_nlj += An.n_lj - Ao.n_lj;
        d_nes += An.n_es - Ao.n_es;
    }
    double v = block_reduce<0>(d_rd, red);
    if (tid == 0) partial[0 * stride + blockIdx.x] = v;
    v = block_reduce<0>(d_es, red);
    if (tid == 0) partial[1 * stride + blockIdx.x] = v;
    v = block_reduce<0>(d_intra, red);
    if (tid == 0) partial[2 * stride + blockIdx.x] = v;
    v = block_reduce<0>((double)d_nlj, red);
    if (tid == 0) partial[3 * stride + blockIdx.x] = v;
    v = block_reduce<0>((double)d_nes, red);
    if (tid == 0) partial[4 * stride + blockIdx.x] = v;
}

__global__ void __launch_bounds__(1024) delta_reduce_kernel(const double* __restrict__ partial, int ncols, int stride,
                                                            double* __restrict__ out) {
    __shared__ double red[32];
    for (int q = 0; q < INC_NPART; q++) {
        double acc = 0.0;
        for (int t = threadIdx.x; t < ncols; t += blockDim.x) acc += partial[q * stride + t];
        const double v = block_reduce<0>(acc, red);
        if (threadIdx.x == 0) out[q] = v;
    }
}

void launch_pair_delta(const Cell& c, bool ortho, const AtomsDev& a, const PairOpts& o, const MovedRec* moved, int nmoved,
                       double* partial, int stride, double* out, cudaStream_t s) {
    const PairConsts k = make_pair_consts(c, o);
    const int nb = a.n_pad / TILE;
#define LAUNCH(ORTHO, FH) \
    pair_delta_kernel<ORTHO, FH><<<nb, TILE, 0, s>>>(c, a, k, o.mixing, o.rd_only, moved, nmoved, partial, stride)
    if (ortho) {
        if (o.fh_order == 0)
            LAUNCH(true, 0);
        else if (o.fh_order == 2)
            LAUNCH(true, 2);
        else
            LAUNCH(true, 4);
    } else {
        if (o.fh_order == 0)
            LAUNCH(false, 0);
        else if (o.fh_order == 2)
            LAUNCH(false, 2);
        else
            LAUNCH(false, 4);
    }
#undef LAUNCH
    delta_reduce_kernel<<<1, 1024, 0, s>>>(partial, nb, stride, out);
}

void launch_pair_reduce(const double* partial, int ntiles, int stride, double* out, cudaStream_t s) {
    pair_reduce_kernel<<<1, 1024, 0, s>>>(partial, ntiles, stride, out);
}

}  // namespace mpmc
