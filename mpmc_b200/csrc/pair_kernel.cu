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
//   cavity_absolute_check()  src/energy/energy.c:49-64
#include "kernels.h"

namespace mpmc {

struct PairConsts {
    double cutoff, alpha, a2, a3, a4;
    double fh2_pref, fh4_pref;  // M2A2*HBAR2/(24 KB T) and M2A4*HBAR4/(1152 KB2 T^2), before 1/mu_red powers
    double fh4_pref_es;         // coulombic.c:141 spells KB*KB instead of the KB2 constant (differs at 4.7e-10)
    double inv_sqrt_pi;
};

__device__ __forceinline__ void mix_lj(int mixing, double ei, double si, double ej, double sj, double& eps,
                                       double& sigma, bool& attractive_only) {
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
                eps = sqrt(ei * ej);
            } else {
                sigma = pow(0.5 * (si6 + sj6), 1. / 6.);
                eps = sqrt(ei * ej) * 2.0 * si3 * sj3 / (si6 + sj6);
            }
        } break;
        case MPMC_MIX_HALGREN:  // pairs.c:104-114
            if (si > 0.0 && sj > 0.0) sigma = (si * si * si + sj * sj * sj) / (si * si + sj * sj);
            if (ei > 0.0 && ej > 0.0) {
                double t = sqrt(ei) + sqrt(ej);
                eps = 4 * ei * ej / (t * t);
            }
            break;
        case MPMC_MIX_C6:  // pairs.c:194-199
            sigma = 0.5 * (si + sj);
            if (sigma != 0.0) {
                double sp = si + sj, sp3 = sp * sp * sp;
                eps = 64.0 * sqrt(ei * ej) * (si * si * si) * (sj * sj * sj) / (sp3 * sp3);
            }
            break;
        default:  // Lorentz-Berthelot, pairs.c:200-211
            if (si < 0.0 || sj < 0.0) {
                attractive_only = true;
                sigma = 0.5 * (fabs(si) + fabs(sj));
            } else if (si == 0 || sj == 0) {
                eps = sqrt(ei * ej);
            } else {
                sigma = 0.5 * (si + sj);
                eps = sqrt(ei * ej);
            }
    }
}

template <bool ORTHO, int FH>
__global__ void __launch_bounds__(TILE) pair_energy_kernel(Cell c, AtomsDev a, PairConsts k, int mixing, int rd_only,
                                                           const int2* __restrict__ tiles, double* __restrict__ partial,
                                                           int stride) {
    __shared__ double sx[TILE], sy[TILE], sz[TILE], sq[TILE], se[TILE], ss[TILE], sm[TILE];
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
        se[tid] = a.eps[j];
        ss[tid] = a.sig[j];
        sm[tid] = a.molmass[j];
        smol[tid] = a.mol[j];
        sfl[tid] = a.flags[j];
    }
    const double xi = a.x[i], yi = a.y[i], zi = a.z[i], qi = a.q[i], ei = a.eps[i], si = a.sig[i], mi = a.molmass[i];
    const int moli = a.mol[i], fli = a.flags[i];
    __syncthreads();

    double rd = 0.0, es = 0.0, intra = 0.0, minr = kMAXVALUE;
    int n_tested = 0, n_lj = 0, n_es = 0;
    const bool lj_i = (ei != 0.0) && (si != 0.0);
    const int j0 = (t.x == t.y) ? tid + 1 : 0;

    for (int jj = j0; jj < TILE; jj++) {
        const int fj = sfl[jj];
        // skip padding and frozen-frozen pairs (pairs.c:81, lj.c:193, coulombic.c:165)
        if (!((fli & fj) & F_VALID) || ((fli & fj) & F_FROZEN)) continue;
        double dx = xi - sx[jj], dy = yi - sy[jj], dz = zi - sz[jj];
        const double qq = qi * sq[jj];
        if (moli == smol[jj]) {
            // same molecule: rd and es excluded (pairs.c:61-64); charge-to-screen term with the
            // UN-imaged separation (coulombic.c:179-180)
            if (!rd_only && qq != 0.0) {
                double r = sqrt(dx * dx + dy * dy + dz * dz);
                intra += qq * erf(k.alpha * r) / r;
            }
            continue;
        }
        min_image<ORTHO>(c, dx, dy, dz);
        const double r2 = dx * dx + dy * dy + dz * dz;
        const double r = sqrt(r2);
        n_tested++;
        minr = fmin(minr, r);

        const double ej = se[jj], sj = ss[jj];
        const bool do_lj = (r - kSMALL_dR < k.cutoff) && lj_i && (ej != 0.0) && (sj != 0.0);
        const bool do_es = !rd_only && !(r > k.cutoff) && (qq != 0.0);
        if (!(do_lj || do_es)) continue;
        const double ir = 1.0 / r;
        double mured = 0.0;
        if (FH) mured = kAMU2KG * mi * sm[jj] / (mi + sm[jj]);

        if (do_lj) {
            double eps, sigma;
            bool att;
            mix_lj(mixing, ei, si, ej, sj, eps, sigma, att);
            const double sor = fabs(sigma) * ir;
            double s6 = sor * sor * sor;
            s6 *= s6;
            const double term6 = s6, term12 = att ? 0.0 : s6 * s6;
            double u = 4.0 * eps * (term12 - term6);
            if (FH) {  // lj.c:29-50
                const double ir2 = ir * ir, ir3 = ir2 * ir, ir4 = ir3 * ir;
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
        if (do_es) {
            const double erfc_term = erfc(k.alpha * r);
            double u = qq * erfc_term * ir;
            if (FH) {  // coulombic.c:132-141
                const double gaussian_term = exp(-k.a2 * r2);
                const double ir2 = ir * ir, ir3 = ir * ir2, ir4 = ir2 * ir2;
                const double g = gaussian_term * k.inv_sqrt_pi;
                const double du = -2.0 * k.alpha * g * ir - erfc_term * ir2;
                const double d2u = 4.0 * g * (k.a3 + ir2) + 2.0 * erfc_term * ir3;
                u += (k.fh2_pref / mured) * (d2u + 2.0 * du * ir);
                if (FH >= 4) {
                    const double d3u = g * (-8.0 * (k.a3 * k.a2) * r - 8.0 * k.a3 * ir - 12.0 * k.alpha * ir3) -
                                       6.0 * erfc_term * ir4;
                    const double d4u = g * (8.0 * k.a3 * k.a2 + 16.0 * k.a3 * k.a4 * r2 + 32.0 * k.a3 * ir2 + 48.0 * ir4) +
                                       24.0 * erfc_term * (ir4 * ir);
                    u += (k.fh4_pref_es / (mured * mured)) * (15.0 * du * ir3 + 4.0 * d3u * ir + d4u);
                }
            }
            es += u;
            n_es++;
        }
    }

    // fixed-order block reduction, one partial row per tile pair
    double v;
    v = block_reduce<0>(rd, red);
    if (tid == 0) partial[0 * stride + blockIdx.x] = v;
    v = block_reduce<0>(es, red);
    if (tid == 0) partial[1 * stride + blockIdx.x] = v;
    v = block_reduce<0>(intra, red);
    if (tid == 0) partial[2 * stride + blockIdx.x] = v;
    v = block_reduce<1>(minr, red);
    if (tid == 0) partial[3 * stride + blockIdx.x] = v;
    v = block_reduce<0>((double)n_tested, red);
    if (tid == 0) partial[4 * stride + blockIdx.x] = v;
    v = block_reduce<0>((double)n_lj, red);
    if (tid == 0) partial[5 * stride + blockIdx.x] = v;
    v = block_reduce<0>((double)n_es, red);
    if (tid == 0) partial[6 * stride + blockIdx.x] = v;
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

void launch_pair_energy(const Cell& c, bool ortho, const AtomsDev& a, const PairOpts& o, const int2* tiles, int ntiles,
                        double* partial, int stride, cudaStream_t s) {
    if (ntiles <= 0) return;
    PairConsts k;
    k.cutoff = c.cutoff;
    k.alpha = o.ewald_alpha;
    k.a2 = k.alpha * k.alpha;
    k.a3 = k.a2 * k.alpha;
    k.a4 = k.a3 * k.alpha;
    k.inv_sqrt_pi = 1.0 / sqrt(M_PI);
    k.fh2_pref = kM2A2 * (kHBAR2 / (24.0 * kKB * o.temperature));
    k.fh4_pref = kM2A4 * (kHBAR4 / (1152.0 * kKB2 * o.temperature * o.temperature));
    k.fh4_pref_es = kM2A4 * (kHBAR4 / (1152.0 * (kKB * kKB * o.temperature * o.temperature)));
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

void launch_pair_reduce(const double* partial, int ntiles, int stride, double* out, cudaStream_t s) {
    pair_reduce_kernel<<<1, 1024, 0, s>>>(partial, ntiles, stride, out);
}

}  // namespace mpmc
