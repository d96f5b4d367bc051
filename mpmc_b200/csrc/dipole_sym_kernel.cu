// dipole_sym_kernel.cu -- K4 (Jacobi family): matrix-free dipole-field-tensor contraction that uses
// the symmetry of the tensor.
//
// The reference stores A_ij = A_ji once per pair (thole_amatrix, src/polarization/thole_matrix.c:73-143,
// "set the lower half of the matrix" :132-134) and contract_dipoles() streams every row
// (src/polarization/thole_iterative.c:27-59).  Here T_ij is never stored; it is recomputed from the
// two positions, ONCE per unordered pair, and applied in both directions:
//     E_i -= T_ij mu_j        E_j -= T_ij mu_i          (T_ij = T_ji, even in d_ij)
// which halves the geometry / damping work of the row-by-row form in dipole_kernel.cu.
//
// Work decomposition (np = padded polarizable atoms, T = np/128 tiles):
//   * tile pairs (I, J = (I+k) mod T), k = 0..T/2, cover every unordered tile pair exactly once
//     (for even T the offset k = T/2 is taken for I < T/2 only); every I has the same amount of
//     work, so the I tiles are the unit of multi-GPU sharding;
//   * one CTA = one I tile x a range of offsets k; thread t keeps atom i = I*128+t (position, dipole,
//     field accumulator) in registers for the whole range;
//   * the J tile (x, y, z, mu) is staged in shared memory; warp w walks it in 4 clusters of 32:
//     lane l meets j = (l+s) mod 32 at step s, so the 32 lanes always touch 32 different j --
//     the J-side accumulator travels with j from lane to lane by one warp shuffle per step and
//     is back home after 32 steps.  No atomics anywhere: the four warps' J-side sums are added in
//     warp order in shared memory and written to a per-offset partial row, epartJ[k-1][3][np];
//     dipole_sym_reduce_kernel adds the partial rows in a fixed order (bit-reproducible run to run).
//   * per (i-cluster, j-cluster) the kernel proves from the clusters' bounding boxes whether every
//     pair is so far apart that the exponential damping factors are exactly 1.0 in fp64
//     (e^{-lr}(...) < 2^-54 for l r >= 48): then exp() and the damping polynomial are skipped -- the
//     result is bit-identical to evaluating them.
//   * minimum image: rint() runs on the conversion pipe (FRND.F64), not on the fp64 FMA pipe.
#include "kernels.h"

namespace mpmc {

namespace {

struct SymConsts {
    double l, l2h, l3s;       // l, l^2/2, l^3/6
    double wd1, wd2, ircut3;  // wolf-full (thole_matrix.c:123-130)
    double far2;              // r^2 beyond which damp1 == damp2 == 1.0 exactly; <= 0: never
    double Lx, Ly, Lz, iLx, iLy, iLz;  // orthorhombic box (cluster classification only)
};

template <bool ORTHO>
__device__ __forceinline__ void min_image_xu(const Cell& c, double& dx, double& dy, double& dz) {
    if (ORTHO) {
        dx = fma(-c.B[0][0], rint(c.R[0][0] * dx), dx);
        dy = fma(-c.B[1][1], rint(c.R[1][1] * dy), dy);
        dz = fma(-c.B[2][2], rint(c.R[2][2] * dz), dz);
    } else {
        const double i0 = rint(c.R[0][0] * dx + c.R[1][0] * dy + c.R[2][0] * dz);
        const double i1 = rint(c.R[0][1] * dx + c.R[1][1] * dy + c.R[2][1] * dz);
        const double i2 = rint(c.R[0][2] * dx + c.R[1][2] * dy + c.R[2][2] * dz);
        dx -= c.B[0][0] * i0 + c.B[1][0] * i1 + c.B[2][0] * i2;
        dy -= c.B[0][1] * i0 + c.B[1][1] * i1 + c.B[2][1] * i2;
        dz -= c.B[0][2] * i0 + c.B[1][2] * i1 + c.B[2][2] * i2;
    }
}

// c1 = damp1/r^3 (delta_pq coefficient), c2 = 3 damp2/r^5 (d_p d_q coefficient); thole_matrix.c:78-131
template <int DAMP, bool WF, bool FAR>
__device__ __forceinline__ void tensor_coeffs(const SymConsts& k, double r2, double aux_i, double aux_j, bool same_mol,
                                              double& c1, double& c2) {
    const double ir = rsqrt(r2);
    const double ir2 = ir * ir;
    const double ir3 = ir * ir2;
    double damp1 = 1.0, damp2 = 1.0;
    if (DAMP == MPMC_DAMP_EXPONENTIAL) {
        if (!FAR) {
            const double r = r2 * ir;
            const double lr = k.l * r;
            const double e = exp(-lr);
            damp1 = 1.0 - e * (k.l2h * r2 + lr + 1.0);
            damp2 = damp1 - e * (k.l3s * r2 * r);
        }
    } else if (DAMP == MPMC_DAMP_LINEAR) {
        const double r = r2 * ir;
        const double s = k.l * (aux_i * aux_j);
        const double v = r / s;
        if (r < s) {
            const double v3 = v * v * v;
            damp1 = (4.0 - 3.0 * v) * v3;
            damp2 = v3 * v;
        }
    } else {  // off: es_excluded pairs are switched off entirely (thole_matrix.c:92-97)
        const bool excl = same_mol || aux_i == 0.0 || aux_j == 0.0;
        damp1 = damp2 = excl ? 0.0 : 1.0;
    }
    if (DAMP == MPMC_DAMP_EXPONENTIAL && FAR) {
        c1 = ir3;
        c2 = (3.0 * ir3) * ir2;
    } else {
        c1 = damp1 * ir3;
        c2 = 3.0 * damp2 * ir3 * ir2;
    }
    if (WF) {
        if (DAMP == MPMC_DAMP_OFF) {
            c1 -= damp1 * k.ircut3;
            c2 -= 3.0 * damp2 * ir2 * k.ircut3;
        } else {
            c1 -= k.wd1;
            c2 -= k.wd2 * ir2;
        }
    }
}

// lower bound of the minimum-image distance^2 between two clusters (orthorhombic box):
// per axis |minimg(ci - cj)| - (hi + hj), clamped at 0
__device__ __forceinline__ double cluster_gap2(const SymConsts& k, const double* __restrict__ a, const double* __restrict__ b,
                                               int nclus, int ca, int cb) {
    double g2 = 0.0;
    {
        double d = a[0 * nclus + ca] - b[0 * nclus + cb];
        d = fabs(fma(-k.Lx, rint(d * k.iLx), d)) - (a[3 * nclus + ca] + b[3 * nclus + cb]);
        d = fmax(d, 0.0);
        g2 = fma(d, d, g2);
    }
    {
        double d = a[1 * nclus + ca] - b[1 * nclus + cb];
        d = fabs(fma(-k.Ly, rint(d * k.iLy), d)) - (a[4 * nclus + ca] + b[4 * nclus + cb]);
        d = fmax(d, 0.0);
        g2 = fma(d, d, g2);
    }
    {
        double d = a[2 * nclus + ca] - b[2 * nclus + cb];
        d = fabs(fma(-k.Lz, rint(d * k.iLz), d)) - (a[5 * nclus + ca] + b[5 * nclus + cb]);
        d = fmax(d, 0.0);
        g2 = fma(d, d, g2);
    }
    return g2;
}

struct SymArgs {
    int n, np, T;
    int tile_begin;  // first I tile of this launch (gridDim.x tiles follow)
    int K;           // T/2: offsets 0..K
    int kper;        // offsets per gridDim.y slice
    const double *x, *y, *z, *aux;
    const int* mol;
    const double* mu;    // [3][np]
    const double* cbox;  // [6][np/32] centre xyz, half extent xyz
    double* epartI;      // [gridDim.y][3][np]
    double* epartJ;      // [K][3][np]
};

// one (i-cluster in registers) x (j-cluster of 32 in shared memory) block, both directions
template <bool ORTHO, int DAMP, bool WF, bool FAR, bool CHECK>
__device__ __forceinline__ void cluster_pair(const Cell& c, const SymConsts& k, int lane, int i, int jbase_global, int n,
                                             const double* __restrict__ sx, const double* __restrict__ sy,
                                             const double* __restrict__ sz, const double* __restrict__ smx,
                                             const double* __restrict__ smy, const double* __restrict__ smz,
                                             const double* __restrict__ saux, const int* __restrict__ smol, double xi,
                                             double yi, double zi, double mxi, double myi, double mzi, double auxi, int moli,
                                             double& exi, double& eyi, double& ezi, double& ajx, double& ajy, double& ajz) {
    ajx = ajy = ajz = 0.0;
#pragma unroll 2
    for (int s = 0; s < 32; s++) {
        const int jl = (lane + s) & 31;
        double dx = xi - sx[jl], dy = yi - sy[jl], dz = zi - sz[jl];
        const double mxj = smx[jl], myj = smy[jl], mzj = smz[jl];
        min_image_xu<ORTHO>(c, dx, dy, dz);
        const double r2 = dx * dx + dy * dy + dz * dz;
        double c1, c2;
        double auxj = 0.0;
        bool same = false;
        if (DAMP != MPMC_DAMP_EXPONENTIAL) auxj = saux[jl];
        if (DAMP == MPMC_DAMP_OFF) same = (moli == smol[jl]);
        tensor_coeffs<DAMP, WF, FAR>(k, r2, auxi, auxj, same, c1, c2);
        if (CHECK) {
            const bool valid = (i < n) && (jbase_global + jl < n);
            c1 = valid ? c1 : 0.0;
            c2 = valid ? c2 : 0.0;
        }
        const double dmj = (dx * mxj + dy * myj + dz * mzj) * c2;
        const double dmi = (dx * mxi + dy * myi + dz * mzi) * c2;
        exi = fma(dmj, dx, exi);
        eyi = fma(dmj, dy, eyi);
        ezi = fma(dmj, dz, ezi);
        exi = fma(-c1, mxj, exi);
        eyi = fma(-c1, myj, eyi);
        ezi = fma(-c1, mzj, ezi);
        ajx = fma(dmi, dx, ajx);
        ajy = fma(dmi, dy, ajy);
        ajz = fma(dmi, dz, ajz);
        ajx = fma(-c1, mxi, ajx);
        ajy = fma(-c1, myi, ajy);
        ajz = fma(-c1, mzi, ajz);
        // the accumulator follows its j: lane l needs the one lane l+1 holds now
        const int src = (lane + 1) & 31;
        ajx = __shfl_sync(0xffffffffu, ajx, src);
        ajy = __shfl_sync(0xffffffffu, ajy, src);
        ajz = __shfl_sync(0xffffffffu, ajz, src);
    }
}

template <bool ORTHO, int DAMP, bool WF>
__global__ void __launch_bounds__(TILE) dipole_sym_kernel(Cell c, SymConsts k, SymArgs a) {
    __shared__ double sx[TILE], sy[TILE], sz[TILE], smx[TILE], smy[TILE], smz[TILE], saux[TILE];
    __shared__ int smol[TILE];
    __shared__ double sej[4][3][TILE];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int I = a.tile_begin + blockIdx.x;
    const int np = a.np, T = a.T, nclus = np >> 5;
    const int i = I * TILE + tid;
    const double xi = a.x[i], yi = a.y[i], zi = a.z[i];
    const double mxi = a.mu[i], myi = a.mu[np + i], mzi = a.mu[2 * np + i];
    const double auxi = (DAMP == MPMC_DAMP_EXPONENTIAL) ? 0.0 : a.aux[i];
    const int moli = (DAMP == MPMC_DAMP_OFF) ? a.mol[i] : 0;
    const int ci = I * 4 + w;
    double exi = 0.0, eyi = 0.0, ezi = 0.0;

    const int k0 = blockIdx.y * a.kper, k1 = min(k0 + a.kper, a.K + 1);
    for (int kk = k0; kk < k1; kk++) {
        if (kk > 0 && 2 * kk == T && I >= kk) continue;  // even T: the half-way offset belongs to I < T/2
        int J = I + kk;
        if (J >= T) J -= T;
        __syncthreads();
        {
            const int j = J * TILE + tid;
            sx[tid] = a.x[j];
            sy[tid] = a.y[j];
            sz[tid] = a.z[j];
            smx[tid] = a.mu[j];
            smy[tid] = a.mu[np + j];
            smz[tid] = a.mu[2 * np + j];
            if (DAMP != MPMC_DAMP_EXPONENTIAL) saux[tid] = a.aux[j];
            if (DAMP == MPMC_DAMP_OFF) smol[tid] = a.mol[j];
        }
        __syncthreads();
        if (kk == 0) {
            // diagonal tile: every ordered pair inside the tile, I side only
            const int jn = min(TILE, a.n - J * TILE);
            if (i < a.n) {
                for (int jj = 0; jj < jn; jj++) {
                    if (jj == tid) continue;
                    double dx = xi - sx[jj], dy = yi - sy[jj], dz = zi - sz[jj];
                    min_image_xu<ORTHO>(c, dx, dy, dz);
                    const double r2 = dx * dx + dy * dy + dz * dz;
                    double c1, c2;
                    tensor_coeffs<DAMP, WF, false>(k, r2, auxi, saux[jj], moli == smol[jj], c1, c2);
                    const double dmj = (dx * smx[jj] + dy * smy[jj] + dz * smz[jj]) * c2;
                    exi = fma(dmj, dx, exi);
                    eyi = fma(dmj, dy, eyi);
                    ezi = fma(dmj, dz, ezi);
                    exi = fma(-c1, smx[jj], exi);
                    eyi = fma(-c1, smy[jj], eyi);
                    ezi = fma(-c1, smz[jj], ezi);
                }
            }
            continue;
        }
        const bool check = ((I + 1) * TILE > a.n) || ((J + 1) * TILE > a.n);  // tiles that hold padding
        // classification of the 4 (i-cluster, j-cluster) blocks of this warp: lanes 0..3 test one each
        unsigned farmask = 0;
        if (ORTHO && DAMP == MPMC_DAMP_EXPONENTIAL && k.far2 > 0.0) {
            bool f = false;
            if (lane < 4) f = cluster_gap2(k, a.cbox, a.cbox, nclus, ci, J * 4 + lane) >= k.far2;
            farmask = __ballot_sync(0xffffffffu, f) & 0xfu;
        }
#pragma unroll 1
        for (int b = 0; b < 4; b++) {
            const int jc = (w + b) & 3;
            const int o = jc * 32;
            double ajx, ajy, ajz;
            const bool far = (farmask >> jc) & 1u;
#define CP(FAR, CHECK)                                                                                                \
    cluster_pair<ORTHO, DAMP, WF, FAR, CHECK>(c, k, lane, i, J * TILE + o, a.n, sx + o, sy + o, sz + o, smx + o, smy + o, \
                                              smz + o, saux + o, smol + o, xi, yi, zi, mxi, myi, mzi, auxi, moli, exi,  \
                                              eyi, ezi, ajx, ajy, ajz)
            if (check) {
                CP(false, true);
            } else if (far) {
                CP(true, false);
            } else {
                CP(false, false);
            }
#undef CP
            sej[w][0][o + lane] = ajx;
            sej[w][1][o + lane] = ajy;
            sej[w][2][o + lane] = ajz;
        }
        __syncthreads();
        {
            double* ej = a.epartJ + (size_t)(kk - 1) * 3 * np + (size_t)J * TILE + tid;
#pragma unroll
            for (int cc = 0; cc < 3; cc++) ej[(size_t)cc * np] = ((sej[0][cc][tid] + sej[1][cc][tid]) + sej[2][cc][tid]) + sej[3][cc][tid];
        }
    }
    double* e = a.epartI + (size_t)blockIdx.y * 3 * np;
    e[i] = exi;
    e[np + i] = eyi;
    e[2 * np + i] = ezi;
}

// E_ind[c][i] = sum_y epartI[y][c][i] (if tile(i) was an I tile of this launch) + sum_k epartJ[k-1][c][i] over the
// offsets k whose I = tile(i) - k was processed.  Fixed order: y ascending, then k ascending in 8 interleaved lanes
// combined in lane order.  One CTA = 32 consecutive atoms x 3 components x 8 k-lanes.
__global__ void __launch_bounds__(256) dipole_sym_reduce_kernel(int np, int T, int K, int tile_begin, int tile_end, int ysplits,
                                                                const double* __restrict__ epartI,
                                                                const double* __restrict__ epartJ,
                                                                double* __restrict__ esum) {
    __shared__ double part[8][3][32];
    const int li = threadIdx.x & 31, lk = threadIdx.x >> 5;
    const int i = blockIdx.x * 32 + li;
    const int J = i / TILE;
    double s0 = 0.0, s1 = 0.0, s2 = 0.0;
    if (lk == 0 && J >= tile_begin && J < tile_end) {
        for (int y = 0; y < ysplits; y++) {
            const double* e = epartI + (size_t)y * 3 * np;
            s0 += e[i];
            s1 += e[np + i];
            s2 += e[2 * np + i];
        }
    }
    for (int kk = 1 + lk; kk <= K; kk += 8) {
        int I = J - kk;
        if (I < 0) I += T;
        if (I < tile_begin || I >= tile_end) continue;
        if (2 * kk == T && I >= kk) continue;
        const double* e = epartJ + (size_t)(kk - 1) * 3 * np;
        s0 += e[i];
        s1 += e[np + i];
        s2 += e[2 * np + i];
    }
    part[lk][0][li] = s0;
    part[lk][1][li] = s1;
    part[lk][2][li] = s2;
    __syncthreads();
    if (threadIdx.x < 96) {
        const int cc = threadIdx.x >> 5;
        double s = part[0][cc][li];
#pragma unroll
        for (int q = 1; q < 8; q++) s += part[q][cc][li];
        esum[(size_t)cc * np + i] = s;
    }
}

// centre and half extent of every 32-atom cluster (valid atoms only)
__global__ void __launch_bounds__(256) cluster_box_kernel(int n, int np, const double* __restrict__ x,
                                                          const double* __restrict__ y, const double* __restrict__ z,
                                                          double* __restrict__ cbox) {
    const int lane = threadIdx.x & 31;
    const int cl = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int nclus = np >> 5;
    if (cl >= nclus) return;
    const int i = cl * 32 + lane;
    const bool live = i < n;
    const double px[3] = {live ? x[i] : 0.0, live ? y[i] : 0.0, live ? z[i] : 0.0};
    const unsigned any = __ballot_sync(0xffffffffu, live);
#pragma unroll
    for (int d = 0; d < 3; d++) {
        double lo = live ? px[d] : 1.0e300, hi = live ? px[d] : -1.0e300;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo = fmin(lo, __shfl_xor_sync(0xffffffffu, lo, o));
            hi = fmax(hi, __shfl_xor_sync(0xffffffffu, hi, o));
        }
        if (lane == 0) {
            cbox[d * nclus + cl] = any ? 0.5 * (lo + hi) : 0.0;
            // round the half extent up by one ulp-scale so the bound stays conservative
            cbox[(3 + d) * nclus + cl] = any ? 0.5 * (hi - lo) * (1.0 + 1e-14) + 1e-300 : 1.0e30;
        }
    }
}

SymConsts make_sym_consts(const Cell& c, bool ortho, const DipoleOpts& o) {
    SymConsts k;
    const double l = o.damp;
    k.l = l;
    k.l2h = 0.5 * l * l;
    k.l3s = l * l * l / 6.0;
    const double rcut = c.cutoff, rcut3 = rcut * rcut * rcut;
    k.ircut3 = 1.0 / rcut3;
    k.wd1 = k.wd2 = 0.0;
    if (o.wolf_full && o.damp_type == MPMC_DAMP_EXPONENTIAL) {
        const double e = exp(-l * rcut);
        const double w1 = 1.0 - e * (0.5 * l * l * rcut * rcut + l * rcut + 1.0);
        const double w2 = w1 - e * (l * l * l * rcut3 / 6.0);
        k.wd1 = w1 / rcut3;
        k.wd2 = 3.0 * w2 / rcut3;
    }
    // l r >= 48  =>  e^{-lr} (l^2 r^2/2 + l r + 1) < 1.8e-18 and e^{-lr} (lr)^3/6 < 2.7e-17, both below 2^-54
    // = 5.55e-17: 1.0 - x rounds to 1.0 exactly, so damp1 == damp2 == 1.0 bit for bit.
    k.far2 = 0.0;
    if (ortho && o.damp_type == MPMC_DAMP_EXPONENTIAL && l > 0.0) {
        const double rf = 48.0 / l * (1.0 + 1e-12);
        k.far2 = rf * rf;
    }
    k.Lx = c.B[0][0];
    k.Ly = c.B[1][1];
    k.Lz = c.B[2][2];
    k.iLx = c.R[0][0];
    k.iLy = c.R[1][1];
    k.iLz = c.R[2][2];
    return k;
}

}  // namespace

void launch_cluster_boxes(const SubsetDev& pol, double* cbox, cudaStream_t s) {
    const int nclus = pol.n_pad / 32;
    if (nclus <= 0) return;
    cluster_box_kernel<<<(nclus * 32 + 255) / 256, 256, 0, s>>>(pol.n, pol.n_pad, pol.x, pol.y, pol.z, cbox);
}

int dipole_sym_partial_rows(int n_pad) {
    const int K = (n_pad / TILE) / 2;
    return K > 1 ? K : 1;
}

void launch_dipole_contract_sym(const Cell& c, bool ortho, const SubsetDev& pol, const DipoleOpts& o, const double* mu,
                                const double* cbox, int tile_begin, int tile_end, int ysplits, double* epartI,
                                double* epartJ, double* esum, cudaStream_t s) {
    const int np = pol.n_pad, T = np / TILE, K = T / 2;
    const int ntile = tile_end - tile_begin;
    if (T <= 0) return;
    const SymConsts k = make_sym_consts(c, ortho, o);
    if (ysplits < 1) ysplits = 1;
    if (ysplits > K + 1) ysplits = K + 1;
    SymArgs a;
    a.n = pol.n;
    a.np = np;
    a.T = T;
    a.tile_begin = tile_begin;
    a.K = K;
    a.kper = (K + 1 + ysplits - 1) / ysplits;
    ysplits = (K + 1 + a.kper - 1) / a.kper;
    a.x = pol.x;
    a.y = pol.y;
    a.z = pol.z;
    a.aux = pol.aux;
    a.mol = pol.mol;
    a.mu = mu;
    a.cbox = cbox;
    a.epartI = epartI;
    a.epartJ = epartJ;
    if (ntile > 0) {
        dim3 grid(ntile, ysplits);
#define LAUNCH(ORTHO, DAMP, WF) dipole_sym_kernel<ORTHO, DAMP, WF><<<grid, TILE, 0, s>>>(c, k, a)
#define DISPATCH(ORTHO)                                              \
    if (o.damp_type == MPMC_DAMP_EXPONENTIAL) {                      \
        if (o.wolf_full) LAUNCH(ORTHO, MPMC_DAMP_EXPONENTIAL, true); \
        else LAUNCH(ORTHO, MPMC_DAMP_EXPONENTIAL, false);            \
    } else if (o.damp_type == MPMC_DAMP_LINEAR) {                    \
        LAUNCH(ORTHO, MPMC_DAMP_LINEAR, false);                      \
    } else {                                                         \
        if (o.wolf_full) LAUNCH(ORTHO, MPMC_DAMP_OFF, true);         \
        else LAUNCH(ORTHO, MPMC_DAMP_OFF, false);                    \
    }
        if (ortho) {
            DISPATCH(true)
        } else {
            DISPATCH(false)
        }
#undef DISPATCH
#undef LAUNCH
    }
    dipole_sym_reduce_kernel<<<np / 32, 256, 0, s>>>(np, T, K, tile_begin, tile_end, ysplits, epartI, epartJ, esum);
}

}  // namespace mpmc
