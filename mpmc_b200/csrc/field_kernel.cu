// field_kernel.cu -- K3: static electric field at every atom from the permanent charges.
//
// Row-owned: thread t of CTA (bx, by) keeps atom i = row_begin + bx*128 + t in registers and walks
// the by-th slice of the compacted CHARGED-atom list staged through shared memory; partial fields go
// to epart[by][c][i] and are added in slice order by field_reduce (no atomics => deterministic).
//
// Reference loops replaced (relative to the reference tree):
//   thole_field_nopbc()  src/polarization/thole_field.c:39-68
//   thole_field_wolf()   src/polarization/thole_field.c:71-124 (+ polar_wolf_lookup.c:8-32)
//   real_term()          src/polarization/polar_ewald.c:38-74
// For the unordered pair (i<j) the reference adds q_j f(r) d to E_i and subtracts q_i f(r) d from
// E_j with d = r_i - r_j imaged; per row that is E_i = sum_j q_j f(r_ij) (r_i - r_j)_img.
#include "kernels.h"

namespace mpmc {

constexpr double kOneOverSqrtPiField = 0.56418958354;  // thole_field.c:10 (truncated constant, kept)
constexpr double kOneOverSqrtPiEwald = 0.5641895835477562869480794515607725858440506293289988;  // polar_ewald.c:6

// field of source j (charge qj) at atom i: one term of thole_field_nopbc / _wolf / real_term
template <bool ORTHO, int MODE>
__device__ __forceinline__ void field_term(const Cell& c, const FieldOpts& o, double xi, double yi, double zi, double qi,
                                           int moli, int fli, double xj, double yj, double zj, double qj, int molj, int fj,
                                           double& ex, double& ey, double& ez, int& cnt) {
    do {
            if (!((fli & fj) & F_VALID) || ((fli & fj) & F_FROZEN)) break;  // thole_field.c:49, polar_ewald.c:49
            const bool same = (moli == molj);
            if (MODE != FIELD_EWALD_REAL && same) break;  // thole_field.c:50
            double dx = xi - xj, dy = yi - yj, dz = zi - zj;
            if (MODE == FIELD_EWALD_REAL && same && dx == 0.0 && dy == 0.0 && dz == 0.0) break;  // the atom itself
            min_image<ORTHO>(c, dx, dy, dz);
            const double r2 = dx * dx + dy * dy + dz * dz;
            double f;  // E_i += q_j * f * d
            if (MODE == FIELD_NOPBC || MODE == FIELD_WOLF_A0) {
                // thole_field.c:55 (r - SMALL_dR < cutoff, r != 0) decided on r^2; the square root is only taken inside a
                // 1e-9 A shell around the cutoff, where the reference's own comparison is evaluated.  1/r from the
                // rsqrt seed + one third-order step (common.cuh) instead of sqrt + division: ~12 instead of ~45 instructions.
                if (r2 == 0.0) break;
                if (r2 >= o.in2) {
                    if (r2 > o.out2) break;
                    if (!(sqrt(r2) - kSMALL_dR < c.cutoff)) break;
                }
                const double ir = rsqrt_fast(r2);
                if (MODE == FIELD_NOPBC)
                    f = ir * ir * ir;
                else {
                    const double rR = 1.0 / c.cutoff;
                    f = (ir * ir - rR * rR) * ir;
                }
                const double qf0 = qj * f;
                ex += qf0 * dx;
                ey += qf0 * dy;
                ez += qf0 * dz;
                cnt++;
                break;
            }
            const double r = sqrt(r2);
            if (MODE == FIELD_EWALD_REAL) {
                if ((r > c.cutoff) || (r == 0.0)) break;  // polar_ewald.c:51
                const bool excluded = same || (qi == 0.0);   // pair es_excluded (pairs.c:61-77); q_j != 0 by construction
                const double g = 2.0 * o.a * kOneOverSqrtPiEwald * exp(-o.a * o.a * r2) * r;
                f = excluded ? (g - erf(o.a * r)) / (r * r2) : (g + erfc(o.a * r)) / (r2 * r);
            } else {
                if (!((r - kSMALL_dR < c.cutoff) && (r != 0.0))) break;  // thole_field.c:55
                if (MODE == FIELD_NOPBC) {
                    f = 1.0 / (r * r * r);
                } else {
                    const double rr = 1.0 / r;
                    if (MODE == FIELD_WOLF_A0) {
                        const double rR = 1.0 / c.cutoff;
                        f = (rr * rr - rR * rR) * rr;
                    } else {
                        double bigmess;
                        if (MODE == FIELD_WOLF_LOOKUP) {  // polar_wolf_lookup.c:28-32 (truncating index)
                            const int kx = (int)(r * 1000);
                            bigmess = (kx >= o.table_max) ? 0.0 : o.table[kx];
                        } else {
                            bigmess = erfc(o.a * r) * rr * rr + 2.0 * o.a * kOneOverSqrtPiField * exp(-o.a * o.a * r * r) * rr;
                        }
                        f = (bigmess - o.cutoffterm) * rr;
                    }
                }
            }
            const double qf = qj * f;
            ex += qf * dx;
            ey += qf * dy;
            ez += qf * dz;
            cnt++;
    } while (0);
}

template <bool ORTHO, int MODE>
__global__ void __launch_bounds__(TILE) static_field_kernel(Cell c, AtomsDev a, SubsetDev ch, FieldOpts o, int row_begin,
                                                            int row_end, int tiles_per_split,
                                                            double* __restrict__ epart, double* __restrict__ count_part) {
    __shared__ double sx[TILE], sy[TILE], sz[TILE], sq[TILE];
    __shared__ int smol[TILE], sfl[TILE];
    __shared__ double red[32];
    const int tid = threadIdx.x;
    const int i = row_begin + blockIdx.x * TILE + tid;
    const bool live = i < row_end;
    const int ii = live ? i : row_begin;
    const double xi = a.x[ii], yi = a.y[ii], zi = a.z[ii], qi = a.q[ii];
    const int moli = a.mol[ii], fli = live ? a.flags[ii] : 0;

    double ex = 0.0, ey = 0.0, ez = 0.0;
    int cnt = 0;
    const int ntiles = ch.n_pad / TILE;
    const int t0 = blockIdx.y * tiles_per_split, t1 = min(t0 + tiles_per_split, ntiles);
    for (int jt = t0; jt < t1; jt++) {
        __syncthreads();
        {
            const int j = jt * TILE + tid;
            sx[tid] = ch.x[j];
            sy[tid] = ch.y[j];
            sz[tid] = ch.z[j];
            sq[tid] = ch.q[j];
            smol[tid] = ch.mol[j];
            sfl[tid] = ch.flags[j];
        }
        __syncthreads();
        for (int jj = 0; jj < TILE; jj++) {
            field_term<ORTHO, MODE>(c, o, xi, yi, zi, qi, moli, fli, sx[jj], sy[jj], sz[jj], sq[jj], smol[jj], sfl[jj], ex, ey, ez,
                                    cnt);
        }
    }
    if (live) {
        double* e = epart + (size_t)blockIdx.y * 3 * a.n_pad;
        e[0 * a.n_pad + i] = ex;
        e[1 * a.n_pad + i] = ey;
        e[2 * a.n_pad + i] = ez;
    }
    if (count_part) {
        double v = block_reduce<0>((double)cnt, red);
        if (tid == 0) count_part[blockIdx.y * gridDim.x + blockIdx.x] = v;
    }
}

// ------------------------------------------------------------------------------------------------
// Cluster form of K3 (the default): spatially sorted image of all atoms, one WARP per 32-atom i-cluster
// (row-owned: lane l accumulates the field at atom ci*32+l over ALL sources, no atomics, no partial rows).
// The j-clusters are scanned in batches of 32 (lane l tests cluster kb+l with cluster_gap2); clusters
// proven to lie beyond the cutoff are skipped: thole_field.c:55 / polar_ewald.c:51 add nothing for them.
// ------------------------------------------------------------------------------------------------
struct __align__(16) FieldRec {
    double x, y, z, q;
    int mol, fl;
};

constexpr int FC_WARPS = 8;  // warps sharing one i-cluster's passing j-clusters (4 left ~13 warps per SM at 16k atoms)
template <bool ORTHO, int MODE>
__global__ void __launch_bounds__(32 * FC_WARPS) field_cluster_kernel(Cell c, SubsetDev a, FieldOpts o, BoxLen box, int clus_begin,
                                                             int clus_end, double cull2, int accumulate,
                                                             double* __restrict__ ef, int ef_stride,
                                                             double* __restrict__ count_part) {
    // One CTA per i-cluster: its warps scan interleaved batches of j-clusters (a warp's serial scan is the
    // critical path of this latency-bound kernel) and their partial fields are added in warp order.
    __shared__ FieldRec rec[FC_WARPS][32];
    __shared__ double sred[FC_WARPS][4][32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int ci = clus_begin + blockIdx.x;
    if (ci >= clus_end) return;  // whole CTA
    const int nclus = a.n_pad >> 5;
    const int i = ci * 32 + lane;
    const double xi = a.x[i], yi = a.y[i], zi = a.z[i], qi = a.q[i];
    const int moli = a.mol[i], fli = a.flags[i];
    double ex = 0.0, ey = 0.0, ez = 0.0;
    int cnt = 0;
    __shared__ unsigned smask[CULL_MAX_BATCHES];
    const int nb = min((nclus + 31) >> 5, CULL_MAX_BATCHES);
    cull_masks(box, a.cbox, nclus, ci, cull2, 0, nb, smask);
    int pos = 0;
    for (int b = 0; b < nb; b++) {
        unsigned mask = smask[b];
        while (mask) {
            const int bit = __ffs(mask) - 1;
            mask &= mask - 1;
            if (((pos++) % FC_WARPS) != w) continue;  // warp-uniform: every FC_WARPS-th passing cluster is this warp's
            const int cj = b * 32 + bit;
            __syncwarp();
            {
                const int j = cj * 32 + lane;
                FieldRec r;
                r.x = a.x[j];
                r.y = a.y[j];
                r.z = a.z[j];
                r.q = a.q[j];
                r.mol = a.mol[j];
                r.fl = a.flags[j];
                rec[w][lane] = r;
            }
            __syncwarp();
            const bool self_cluster = (cj == ci);
            for (int s = 0; s < 32; s++) {
                const FieldRec& r = rec[w][s];
                if (r.q == 0.0) continue;                  // a source without charge adds exact zeros
                if (self_cluster && s == lane) continue;   // the atom itself
                field_term<ORTHO, MODE>(c, o, xi, yi, zi, qi, moli, fli, r.x, r.y, r.z, r.q, r.mol, r.fl, ex, ey, ez, cnt);
            }
        }
    }
    sred[w][0][lane] = ex;
    sred[w][1][lane] = ey;
    sred[w][2][lane] = ez;
    sred[w][3][lane] = (double)cnt;
    __syncthreads();
    if (w != 0) return;
    double cntd = sred[0][3][lane];
    ex = sred[0][0][lane], ey = sred[0][1][lane], ez = sred[0][2][lane];
#pragma unroll
    for (int q = 1; q < FC_WARPS; q++) {  // warp order: fixed association
        ex += sred[q][0][lane];
        ey += sred[q][1][lane];
        ez += sred[q][2][lane];
        cntd += sred[q][3][lane];
    }
    if (fli & F_VALID) {
        const int g = a.idx[i];
        if (accumulate) {
            ef[0 * ef_stride + g] += ex;
            ef[1 * ef_stride + g] += ey;
            ef[2 * ef_stride + g] += ez;
        } else {
            ef[0 * ef_stride + g] = ex;
            ef[1 * ef_stride + g] = ey;
            ef[2 * ef_stride + g] = ez;
        }
    }
    if (count_part) {
        const double v = warp_sum(cntd);
        if (lane == 0) count_part[ci - clus_begin] = v;
    }
}

__global__ void field_reduce_kernel(const double* __restrict__ epart, int jsplits, int n_pad, int row_begin, int row_end,
                                    int accumulate, double* __restrict__ ef) {
    const int i = row_begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= row_end) return;
    for (int cdim = 0; cdim < 3; cdim++) {
        double acc = accumulate ? ef[cdim * n_pad + i] : 0.0;
        for (int js = 0; js < jsplits; js++) acc += epart[((size_t)js * 3 + cdim) * n_pad + i];
        ef[cdim * n_pad + i] = acc;
    }
}

void launch_static_field(const Cell& c, bool ortho, const AtomsDev& a, const SubsetDev& charged, const FieldOpts& o,
                         int row_begin, int row_end, int jsplits, double* epart, double* count_part, cudaStream_t s) {
    const int rows = row_end - row_begin;
    if (rows <= 0) return;
    const int ntiles = charged.n_pad / TILE;
    const int tps = (ntiles + jsplits - 1) / jsplits;
    dim3 grid((rows + TILE - 1) / TILE, jsplits);
    FieldOpts fo = o;
    fo.in2 = (c.cutoff - 1e-9) * (c.cutoff - 1e-9);
    fo.out2 = (c.cutoff + 1e-9) * (c.cutoff + 1e-9);
#define LAUNCH(ORTHO, MODE)                                                                                          \
    static_field_kernel<ORTHO, MODE><<<grid, TILE, 0, s>>>(c, a, charged, fo, row_begin, row_end, tps, epart, count_part)
#define DISPATCH(ORTHO)                                     \
    switch (o.mode) {                                       \
        case FIELD_NOPBC: LAUNCH(ORTHO, FIELD_NOPBC); break; \
        case FIELD_WOLF_A0: LAUNCH(ORTHO, FIELD_WOLF_A0); break; \
        case FIELD_WOLF: LAUNCH(ORTHO, FIELD_WOLF); break;  \
        case FIELD_WOLF_LOOKUP: LAUNCH(ORTHO, FIELD_WOLF_LOOKUP); break; \
        default: LAUNCH(ORTHO, FIELD_EWALD_REAL);           \
    }
    if (ortho) {
        DISPATCH(true)
    } else {
        DISPATCH(false)
    }
#undef DISPATCH
#undef LAUNCH
}

void launch_static_field_clusters(const Cell& c, bool ortho, const SubsetDev& a, const FieldOpts& o, int clus_begin,
                                  int clus_end, double cull, bool accumulate, double* ef, int ef_stride, double* count_part,
                                  cudaStream_t s) {
    const int nown = clus_end - clus_begin;
    if (nown <= 0) return;
    const double cull2 = (ortho && cull > 0.0) ? cull * cull : 0.0;
    const BoxLen box = box_len(c);
    const int grid = nown;
    FieldOpts fo = o;
    fo.in2 = (c.cutoff - 1e-9) * (c.cutoff - 1e-9);
    fo.out2 = (c.cutoff + 1e-9) * (c.cutoff + 1e-9);
#define LAUNCH(ORTHO, MODE)                                                                                                  \
    field_cluster_kernel<ORTHO, MODE><<<grid, 32 * FC_WARPS, 0, s>>>(c, a, fo, box, clus_begin, clus_end, cull2, accumulate ? 1 : 0, \
                                                                     ef, ef_stride, count_part)
#define DISPATCH(ORTHO)                                     \
    switch (o.mode) {                                       \
        case FIELD_NOPBC: LAUNCH(ORTHO, FIELD_NOPBC); break; \
        case FIELD_WOLF_A0: LAUNCH(ORTHO, FIELD_WOLF_A0); break; \
        case FIELD_WOLF: LAUNCH(ORTHO, FIELD_WOLF); break;  \
        case FIELD_WOLF_LOOKUP: LAUNCH(ORTHO, FIELD_WOLF_LOOKUP); break; \
        default: LAUNCH(ORTHO, FIELD_EWALD_REAL);           \
    }
    if (ortho) {
        DISPATCH(true)
    } else {
        DISPATCH(false)
    }
#undef DISPATCH
#undef LAUNCH
}

void launch_field_reduce(const double* epart, int jsplits, int n_pad, int row_begin, int row_end, bool accumulate,
                         double* ef, cudaStream_t s) {
    const int rows = row_end - row_begin;
    if (rows <= 0) return;
    field_reduce_kernel<<<(rows + 255) / 256, 256, 0, s>>>(epart, jsplits, n_pad, row_begin, row_end, accumulate ? 1 : 0, ef);
}

}  // namespace mpmc
