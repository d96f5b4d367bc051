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
//   * minimum image: rint() runs on the conversion pipe (FRND.F64), not on the fp64 FMA pipe.  Where the boxes of an
//     (i-cluster, j-cluster) pair prove that EVERY atom pair picks the same image n (per axis |centre distance - n L| + half
//     extents < L/2), the i positions are moved by -n L once per cluster pair and the per-pair image step (3 DMUL + 3 FRND
//     + 3 DFMA of ~41 fp64 instructions) is left out: about half of the cluster pairs of a MOF cell.  n = 0 gives the same
//     bits as the per-pair form; n != 0 rounds x_i - n L once where the per-pair form rounds (x_i - x_j) - n L (1 ulp of d).
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <tuple>
#include <vector>

#include "kernels.h"

namespace mpmc {

namespace {

struct SymConsts {
    double l, l2h, l3s;       // l, l^2/2, l^3/6
    double wd1, wd2, ircut3;  // wolf-full (thole_matrix.c:123-130)
    double far2;              // r^2 beyond which damp1 == damp2 == 1.0 exactly; <= 0: never
    BoxLen box;               // orthorhombic box (cluster classification only)
    double ec[12];                     // 1/k!, k = 0..11 (exp_neg_fast)
    int fastexp;                       // l * (largest possible r) < 700: exp_neg_fast is valid
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

struct SymArgs {
    int n, np, S;    // polarizable atoms, padded, super tiles (np / 256)
    int Y;           // offset slices OF THIS LAUNCH: work items (I super tile, slice y0 + y), numbered y-major: c = y * S + I
    int y0;          // first slice of this launch, jrow0 = first offset whose J-side row it writes (row = kk - jrow0):
    int jrow0;       // a contraction is split into launches over slice ranges when its J-side rows would not fit the cap
    int rank, world; // this launch owns the items c with c % world == rank (round robin: every rank gets the same mix of
                     // near -- damped, expensive -- and far tile pairs; contiguous shares left the ranks 9 % apart)
    int K;           // S/2: offsets 0..K
    int kper;        // offsets per gridDim.y slice
    int hsplit;      // gridDim.z.  1: a CTA walks both halves of the J super tile; 2: one half each (finer grain: the last
                     // wave of a launch costs less); 4 / 8: one half, and 2 / 1 of its four j-clusters (launches of
                     // less than two waves of CTAs: the single item is then the critical path)
    int unishift;    // 1: cluster pairs with one image for all their atom pairs skip the per-pair minimum image
    const double *x, *y, *z, *aux;
    const int* mol;
    const double* mu;    // [3][np]
    const double* cbox;  // [7][np/32] centre xyz, half extent xyz, 1.0 if no atom of the cluster sits in another image than its first
    double* epartI;      // [gridDim.y][3][np]
    double* epartJ;      // [K][3][np]
};

#ifndef SYM_UNROLL
#define SYM_UNROLL 4  // 8 overflows the instruction cache (ncu: stall_no_instruction), 2 and 4 time alike
#endif
#ifndef SYM_MINB
#define SYM_MINB 4
#endif
#ifndef SYM_HSPLIT_MAX
#define SYM_HSPLIT_MAX 8
#endif
constexpr int kSymUnroll = SYM_UNROLL;
// unroll factor of the 32-step loop per class of cluster pair (far / damped, with / without one image for all pairs);
// SYM_UNR_DAMPUNI 0: damped pairs always take the per-pair minimum image.  The kernel is sensitive to the size of its hot
// loops (instruction cache): with all four classes at 4 the one-image path LOST 6.7 % (39.66 against 37.16 ms per energy()
// at mof_h2_63k), with the two damped classes at 2 it gains 1.3 % (36.33 against 36.82; profiles/r02_k4_one_image.md)
#ifndef SYM_UNR_FAR
#define SYM_UNR_FAR SYM_UNROLL
#endif
#ifndef SYM_UNR_FARUNI
#define SYM_UNR_FARUNI SYM_UNROLL
#endif
#ifndef SYM_UNR_DAMP
#define SYM_UNR_DAMP 2
#endif
#ifndef SYM_UNR_DAMPUNI
#define SYM_UNR_DAMPUNI 2
#endif
constexpr int STILE = 256;  // atoms per super tile: 2 per thread
constexpr int JREC = 6;     // doubles per staged j record: x y z mx my mz

// c1 = damp1/r^3, c2 = 3 damp2/r^5 from r^2 (exponential damping, fast math helpers)
template <bool WF, bool FAR, bool FASTEXP>
__device__ __forceinline__ void coeffs_exp(const SymConsts& k, double r2, double& c1, double& c2) {
    const double ir = rsqrt_fast(r2);
    const double ir2 = ir * ir;
    const double ir3 = ir * ir2;
    if (FAR) {
        c1 = ir3;
        c2 = (3.0 * ir3) * ir2;
    } else {
        const double r = r2 * ir;
        const double lr = k.l * r;
        const double e = FASTEXP ? exp_neg_fast(k, lr) : exp(-lr);
        const double damp1 = 1.0 - e * (k.l2h * r2 + lr + 1.0);
        const double damp2 = damp1 - e * (k.l3s * r2 * r);
        c1 = damp1 * ir3;
        c2 = 3.0 * damp2 * ir3 * ir2;
    }
    if (WF) {
        c1 -= k.wd1;
        c2 -= k.wd2 * ir2;
    }
}

// Do all atom pairs of clusters (ca, cb) pick the same periodic image?  Both clusters must hold their atoms in ONE image
// (row 6 of the boxes: then the box bounds the stored coordinates), and per axis |D - n L| + half extents must stay below
// L/2 by a margin (1e-6 A: ~1e-8 of the box, nine orders above the rounding of rint(d / L)), D = centre distance,
// n = rint(D / L): then rint((x_i - x_j) / L) = n for every pair.  Returns bit 30 | (n_x + 512) | (n_y + 512) << 10 |
// (n_z + 512) << 20, or 0.
constexpr int kNoImage = 512 | (512 << 10) | (512 << 20);
__device__ __forceinline__ int uniform_image(const BoxLen& k, const double* __restrict__ cbox, int nclus, int ca, int cb) {
    if (cbox[6 * nclus + ca] != 1.0 || cbox[6 * nclus + cb] != 1.0) return 0;
    const double L[3] = {k.Lx, k.Ly, k.Lz}, iL[3] = {k.iLx, k.iLy, k.iLz};
    int code = 1 << 30;
#pragma unroll
    for (int d = 0; d < 3; d++) {
        const double D = cbox[d * nclus + ca] - cbox[d * nclus + cb];
        const double n = rint(D * iL[d]);
        const double rem = fabs(fma(-L[d], n, D)) + (cbox[(3 + d) * nclus + ca] + cbox[(3 + d) * nclus + cb]);
        if (!(rem < 0.5 * L[d] - 1e-6) || !(fabs(n) <= 500.0)) return 0;
        code |= ((int)n + 512) << (10 * d);
    }
    return code;
}

// One j-cluster of 32 (records staged twice in a row, so lane+s never wraps) against the two i atoms a
// lane keeps in registers; T once per unordered pair, applied in both directions.
template <bool ORTHO, int DAMP, bool WF, bool FAR, bool CHECK, bool FASTEXP, bool UNI, int UNR>
__device__ __forceinline__ void cluster_pair2(const Cell& c, const SymConsts& k, int lane, int ia, int jbase_global, int n,
                                              const double* __restrict__ rec, const double* __restrict__ saux,
                                              const int* __restrict__ smol, const double (&pi)[2][6],
                                              const double (&auxi)[2], const int (&moli)[2], double (&ei)[2][3],
                                              double& ajx, double& ajy, double& ajz, const bool uni_rt = false) {
    ajx = ajy = ajz = 0.0;
    const double* base = rec + lane * JREC;
#pragma unroll UNR
    for (int s = 0; s < 32; s++) {
        const double2 pxy = *reinterpret_cast<const double2*>(base + s * JREC);
        const double2 pzm = *reinterpret_cast<const double2*>(base + s * JREC + 2);
        const double2 pmm = *reinterpret_cast<const double2*>(base + s * JREC + 4);
        const double xj = pxy.x, yj = pxy.y, zj = pzm.x, mxj = pzm.y, myj = pmm.x, mzj = pmm.y;
        double auxj = 0.0;
        int molj = 0;
        if (DAMP != MPMC_DAMP_EXPONENTIAL) auxj = saux[lane + s];
        if (DAMP == MPMC_DAMP_OFF) molj = smol[lane + s];
#pragma unroll
        for (int u = 0; u < 2; u++) {
            double dx = pi[u][0] - xj, dy = pi[u][1] - yj, dz = pi[u][2] - zj;
            if (!UNI && !uni_rt) min_image_xu<ORTHO>(c, dx, dy, dz);  // UNI: pi is already in the image every pair of this cluster pair picks
            const double r2 = dx * dx + dy * dy + dz * dz;
            double c1, c2;
            if (DAMP == MPMC_DAMP_EXPONENTIAL)
                coeffs_exp<WF, FAR, FASTEXP>(k, r2, c1, c2);
            else
                tensor_coeffs<DAMP, WF, false>(k, r2, auxi[u], auxj, moli[u] == molj, c1, c2);
            if (CHECK) {
                const bool valid = (ia + 32 * u < n) && (jbase_global + ((lane + s) & 31) < n);
                c1 = valid ? c1 : 0.0;
                c2 = valid ? c2 : 0.0;
            }
            const double dmj = (dx * mxj + dy * myj + dz * mzj) * c2;
            const double dmi = (dx * pi[u][3] + dy * pi[u][4] + dz * pi[u][5]) * c2;
            ei[u][0] = fma(dmj, dx, ei[u][0]);
            ei[u][1] = fma(dmj, dy, ei[u][1]);
            ei[u][2] = fma(dmj, dz, ei[u][2]);
            ei[u][0] = fma(-c1, mxj, ei[u][0]);
            ei[u][1] = fma(-c1, myj, ei[u][1]);
            ei[u][2] = fma(-c1, mzj, ei[u][2]);
            ajx = fma(dmi, dx, ajx);
            ajy = fma(dmi, dy, ajy);
            ajz = fma(dmi, dz, ajz);
            ajx = fma(-c1, pi[u][3], ajx);
            ajy = fma(-c1, pi[u][4], ajy);
            ajz = fma(-c1, pi[u][5], ajz);
        }
        // the accumulator follows its j: lane l needs the one lane l+1 holds now
        const int src = (lane + 1) & 31;
        ajx = __shfl_sync(0xffffffffu, ajx, src);
        ajy = __shfl_sync(0xffffffffu, ajy, src);
        ajz = __shfl_sync(0xffffffffu, ajz, src);
    }
}

template <bool ORTHO, int DAMP, bool WF, bool FASTEXP>
__global__ void __launch_bounds__(TILE, SYM_MINB) dipole_sym_kernel(Cell c, SymConsts k, SymArgs a) {
    // j records of one half super tile: 4 clusters x (32 records stored twice) x 6 doubles
    __shared__ __align__(16) double srec[4][64][JREC];
    __shared__ double saux[4][64];
    __shared__ int smol[4][64];
    __shared__ double sej[4][3][TILE];
    __shared__ double spi[2][3][TILE];  // the positions as stored; pi[u][0..2] hold them in the image of the current cluster pair
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int np = a.np, S = a.S, nclus = np >> 5;
    const int cidx = a.rank + (int)blockIdx.x * a.world;  // this rank's blockIdx.x-th item
    if (cidx >= S * a.Y) return;                          // (whole CTA, before any barrier)
    const int I = cidx % S, ylocal = cidx / S, yslice = a.y0 + ylocal;
    // thread t keeps atoms ia and ia + 32: warp w owns the i-clusters 2w and 2w+1 of the super tile
    const int ia = I * STILE + w * 64 + lane;
    double pi[2][6], auxi[2] = {0.0, 0.0}, ei[2][3] = {{0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}};
    int moli[2] = {0, 0};
#pragma unroll
    for (int u = 0; u < 2; u++) {
        const int i = ia + 32 * u;
        pi[u][0] = spi[u][0][tid] = a.x[i];  // (read back by this thread only: no barrier needed)
        pi[u][1] = spi[u][1][tid] = a.y[i];
        pi[u][2] = spi[u][2][tid] = a.z[i];
        pi[u][3] = a.mu[i];
        pi[u][4] = a.mu[np + i];
        pi[u][5] = a.mu[2 * np + i];
        if (DAMP != MPMC_DAMP_EXPONENTIAL) auxi[u] = a.aux[i];
        if (DAMP == MPMC_DAMP_OFF) moli[u] = a.mol[i];
    }
    const int ci = (I * STILE + w * 64) >> 5;  // first of this warp's two i-clusters

    const int k0 = yslice * a.kper, k1 = min(k0 + a.kper, a.K + 1);
    for (int kk = k0; kk < k1; kk++) {
        if (kk > 0 && 2 * kk == S && I >= kk) continue;  // even S: the half-way offset belongs to I < S/2
        int J = I + kk;
        if (J >= S) J -= S;
        // gridDim.z = hsplit slices of the J super tile: halves, then groups of the half's four j-clusters
        const int zq = a.hsplit >= 2 ? a.hsplit >> 1 : 1, zb = (int)blockIdx.z % zq;
        const int h0 = a.hsplit >= 2 ? (int)blockIdx.z / zq : 0, h1 = a.hsplit >= 2 ? h0 + 1 : 2;
        const int b0 = zb * (4 / zq), b1 = b0 + 4 / zq;  // this CTA's j-clusters of the half
#pragma unroll 1
        for (int half = h0; half < h1; half++) {
            const int jb = J * STILE + half * TILE;  // first atom of this half
            __syncthreads();
            {
                const int j = jb + tid;
                const int cl = tid >> 5, l = tid & 31;
                const double xj = a.x[j], yj = a.y[j], zj = a.z[j];
                const double mx = a.mu[j], my = a.mu[np + j], mz = a.mu[2 * np + j];
#pragma unroll
                for (int d = 0; d < 2; d++) {
                    double* r = &srec[cl][l + 32 * d][0];
                    r[0] = xj;
                    r[1] = yj;
                    r[2] = zj;
                    r[3] = mx;
                    r[4] = my;
                    r[5] = mz;
                }
                if (DAMP != MPMC_DAMP_EXPONENTIAL) saux[cl][l] = saux[cl][l + 32] = a.aux[j];
                if (DAMP == MPMC_DAMP_OFF) smol[cl][l] = smol[cl][l + 32] = a.mol[j];
            }
            __syncthreads();
            if (kk == 0) {
                // diagonal super tile: every ordered pair inside it, I side only
                const int jn = min(32 * b1, a.n - jb);
#pragma unroll
                for (int u = 0; u < 2; u++)
#pragma unroll
                    for (int d = 0; d < 3; d++) pi[u][d] = spi[u][d][tid];
#pragma unroll
                for (int u = 0; u < 2; u++) {
                    const int i = ia + 32 * u;
                    if (i >= a.n) continue;
                    for (int jj = 32 * b0; jj < jn; jj++) {
                        if (jb + jj == i) continue;
                        const double* r = &srec[jj >> 5][jj & 31][0];
                        double dx = pi[u][0] - r[0], dy = pi[u][1] - r[1], dz = pi[u][2] - r[2];
                        min_image_xu<ORTHO>(c, dx, dy, dz);
                        const double r2 = dx * dx + dy * dy + dz * dz;
                        double c1, c2;
                        tensor_coeffs<DAMP, WF, false>(k, r2, auxi[u], saux[jj >> 5][jj & 31], moli[u] == smol[jj >> 5][jj & 31],
                                                       c1, c2);
                        const double dmj = (dx * r[3] + dy * r[4] + dz * r[5]) * c2;
                        ei[u][0] = fma(dmj, dx, ei[u][0]);
                        ei[u][1] = fma(dmj, dy, ei[u][1]);
                        ei[u][2] = fma(dmj, dz, ei[u][2]);
                        ei[u][0] = fma(-c1, r[3], ei[u][0]);
                        ei[u][1] = fma(-c1, r[4], ei[u][1]);
                        ei[u][2] = fma(-c1, r[5], ei[u][2]);
                    }
                }
                continue;
            }
            const bool check = ((I + 1) * STILE > a.n) || (jb + TILE > a.n);  // tiles that hold padding
            // (2 i-clusters) x (4 j-clusters): lanes 0..7 test one combination each
            unsigned farmask = 0;
            if (ORTHO && DAMP == MPMC_DAMP_EXPONENTIAL && k.far2 > 0.0) {
                bool f = false;
                if (lane < 8) f = cluster_gap2(k.box, a.cbox, nclus, ci + (lane >> 2), a.cbox, nclus, (jb >> 5) + (lane & 3)) >= k.far2;
                farmask = __ballot_sync(0xffffffffu, f);
                farmask = farmask & (farmask >> 4) & 0xfu;  // both i-clusters far from j-cluster b
            }
            // one image for all 32 x 32 pairs of (i-cluster, j-cluster)?  bit 30 + the image numbers (10 bits each, +512)
            int ucode = 0;
            if (ORTHO && a.unishift && !check && lane < 8)
                ucode = uniform_image(k.box, a.cbox, nclus, ci + (lane >> 2), (jb >> 5) + (lane & 3));
#pragma unroll 1
            for (int b = b0; b < b1; b++) {
                const int jc = b;  // all four warps on the same j-cluster: their i-clusters are neighbours, so the
                                   // far / damped classes coincide and nobody waits long at the barrier
                double ajx, ajy, ajz;
                const bool far = (farmask >> jc) & 1u;
                bool uni = false;
                if (ORTHO) {
                    const int cu0 = __shfl_sync(0xffffffffu, ucode, jc), cu1 = __shfl_sync(0xffffffffu, ucode, 4 + jc);
                    uni = (cu0 & cu1 & (1 << 30)) != 0;  // warp-uniform
                    if (SYM_UNR_DAMPUNI == 0) uni = uni && far;
                    const int cu[2] = {uni ? cu0 : kNoImage, uni ? cu1 : kNoImage};
#pragma unroll
                    for (int u = 0; u < 2; u++) {  // x - 0 L = x exactly: a cluster pair without a common image keeps the stored positions
                        pi[u][0] = fma(-k.box.Lx, (double)((cu[u] & 1023) - 512), spi[u][0][tid]);
                        pi[u][1] = fma(-k.box.Ly, (double)(((cu[u] >> 10) & 1023) - 512), spi[u][1][tid]);
                        pi[u][2] = fma(-k.box.Lz, (double)(((cu[u] >> 20) & 1023) - 512), spi[u][2][tid]);
                    }
                }
#define CP(FAR, CHECK, UNI, UNR)                                                                                                     \
    cluster_pair2<ORTHO, DAMP, WF, FAR, CHECK, FASTEXP, UNI, UNR>(c, k, lane, ia, jb + jc * 32, a.n, &srec[jc][0][0], &saux[jc][0], \
                                                                  &smol[jc][0], pi, auxi, moli, ei, ajx, ajy, ajz)
#ifdef SYM_UNI_RUNTIME
#define CPR(FAR)                                                                                                                  \
    cluster_pair2<ORTHO, DAMP, WF, FAR, false, FASTEXP, false, kSymUnroll>(c, k, lane, ia, jb + jc * 32, a.n, &srec[jc][0][0], &saux[jc][0], \
                                                               &smol[jc][0], pi, auxi, moli, ei, ajx, ajy, ajz, uni)
                if (check) {
                    CP(false, true, false, kSymUnroll);
                } else if (far) {
                    CPR(true);
                } else {
                    CPR(false);
                }
#undef CPR
#else
                if (check) {
                    CP(false, true, false, 2);
                } else if (ORTHO && uni && far) {
                    CP(true, false, true, SYM_UNR_FARUNI);
                } else if (ORTHO && SYM_UNR_DAMPUNI > 0 && uni) {
                    CP(false, false, true, (SYM_UNR_DAMPUNI > 0 ? SYM_UNR_DAMPUNI : 1));
                } else if (far) {
                    CP(true, false, false, SYM_UNR_FAR);
                } else {
                    CP(false, false, false, SYM_UNR_DAMP);
                }
#endif
#undef CP
                sej[w][0][jc * 32 + lane] = ajx;
                sej[w][1][jc * 32 + lane] = ajy;
                sej[w][2][jc * 32 + lane] = ajz;
            }
            __syncthreads();
            if ((tid >> 5) >= b0 && (tid >> 5) < b1) {  // the j atoms of this CTA's clusters (the other slices write the rest of the row)
                double* ej = a.epartJ + (size_t)(kk - a.jrow0) * 3 * np + (size_t)jb + tid;
#pragma unroll
                for (int cc = 0; cc < 3; cc++)
                    ej[(size_t)cc * np] = ((sej[0][cc][tid] + sej[1][cc][tid]) + sej[2][cc][tid]) + sej[3][cc][tid];
            }
        }
    }
    double* e = a.epartI + (size_t)(ylocal * a.hsplit + blockIdx.z) * 3 * np;
#pragma unroll
    for (int u = 0; u < 2; u++) {
        const int i = ia + 32 * u;
        e[i] = ei[u][0];
        e[np + i] = ei[u][1];
        e[2 * np + i] = ei[u][2];
    }
}

// E_ind[c][i] = sum of the partial rows this launch wrote for atom i: the I-side rows epartI[y] of the items (tile(i), y)
// and the J-side rows epartJ[k-1] of the items (tile(i) - k, slice(k)) it owned.  Which rows those are is a pure
// function of (super tiles, slices, rank, world): the host lists them once per layout (sym_reduce_list: CSR over the
// super tiles, entry = row index, bit 30 set for a J-side row, in the fixed order y ascending then k ascending), so the
// kernel touches nothing it does not own -- its time falls with 1/world like the contraction's (the loop over every
// (y, k) with an ownership test per step cost 70 us at any world size: 13 % of a contraction at 8 GPUs).
// One CTA = 32 consecutive atoms x 3 components x 8 lanes over the list; the lanes' sums are combined in lane order.
__global__ void __launch_bounds__(256) dipole_sym_reduce_kernel(int np, const int* __restrict__ offs, const int* __restrict__ rows,
                                                                const double* __restrict__ epartI,
                                                                const double* __restrict__ epartJ, int accumulate,
                                                                double* __restrict__ esum) {
    __shared__ double part[8][3][32];
    const int li = threadIdx.x & 31, lk = threadIdx.x >> 5;
    const int i = blockIdx.x * 32 + li;
    const int J = i / STILE;
    double s0 = 0.0, s1 = 0.0, s2 = 0.0;
    const int e0 = offs[J], e1 = offs[J + 1];
#pragma unroll 4
    for (int e = e0 + lk; e < e1; e += 8) {
        const int r = rows[e];
        const double* src = ((r & (1 << 30)) ? epartJ : epartI) + (size_t)(r & ((1 << 30) - 1)) * 3 * np;
        s0 += src[i];
        s1 += src[np + i];
        s2 += src[2 * np + i];
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
        esum[(size_t)cc * np + i] = accumulate ? esum[(size_t)cc * np + i] + s : s;  // later slice ranges add to the first
    }
}

// Raises this rank's flag in every peer's memory.  Launched right after dipole_sym_reduce_kernel on the same stream:
// the kernel boundary makes the partial field complete in this GPU's memory, the system-scope fence + release
// store order it before the flag for readers on other GPUs.  (A fence in every CTA of the reduce kernel instead of
// this 3 us launch cost 24 us per contraction at 8 GPUs.)
__global__ void peer_signal_kernel(PeerSignal sig) {
    const int r = threadIdx.x;
    if (r >= sig.world) return;
    __threadfence_system();
    unsigned long long* f = sig.flags[r] + sig.me;
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(f), "l"(sig.seq) : "memory");
}

// centre and half extent of every 32-atom cluster (valid atoms only).  Orthorhombic boxes: positions are
// taken as minimum-image displacements from the cluster's first atom, so a cluster whose atoms sit in
// different periodic images (MPMC keeps unwrapped coordinates) still gets a tight box; the gap test in
// cluster_gap2 is invariant under lattice translations of either cluster.
__global__ void __launch_bounds__(256) cluster_box_kernel(int n, int np, double Lx, double Ly, double Lz, double iLx,
                                                          double iLy, double iLz, int ortho, const double* __restrict__ x,
                                                          const double* __restrict__ y, const double* __restrict__ z,
                                                          double* __restrict__ cbox) {
    const int lane = threadIdx.x & 31;
    const int cl = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int nclus = np >> 5;
    if (cl >= nclus) return;
    const int i = cl * 32 + lane;
    const bool live = i < n;
    double px[3] = {live ? x[i] : 0.0, live ? y[i] : 0.0, live ? z[i] : 0.0};
    const unsigned any = __ballot_sync(0xffffffffu, live);
    const double L[3] = {Lx, Ly, Lz}, iL[3] = {iLx, iLy, iLz};
    bool one_image = true;  // every atom in the image of the first: the box bounds the stored coordinates
#pragma unroll
    for (int d = 0; d < 3; d++) {
        const double p0 = __shfl_sync(0xffffffffu, px[d], 0);  // lane 0 is live whenever any lane is
        double v = px[d] - p0;
        if (ortho) {
            const double m = rint(v * iL[d]);
            one_image = one_image && (m == 0.0 || !live);
            v = fma(-L[d], m, v);
        }
        double lo = live ? v : 1.0e300, hi = live ? v : -1.0e300;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo = fmin(lo, __shfl_xor_sync(0xffffffffu, lo, o));
            hi = fmax(hi, __shfl_xor_sync(0xffffffffu, hi, o));
        }
        if (lane == 0) {
            cbox[d * nclus + cl] = any ? p0 + 0.5 * (lo + hi) : 0.0;
            // half extent padded so that rounding in centre / extent can never make the bound optimistic
            cbox[(3 + d) * nclus + cl] = any ? 0.5 * (hi - lo) * (1.0 + 1e-12) + 1e-9 : 1.0e30;
        }
    }
    const bool all_one = __all_sync(0xffffffffu, one_image);
    if (lane == 0) cbox[6 * nclus + cl] = (any && ortho && all_one) ? 1.0 : 0.0;
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
        // MPMC_B200_SYM_FAR_LR (A/B): a smaller product trades the bit-exact skip for more far cluster pairs; at l r = 40 the
        // damping factors differ from 1 by 4.5e-14 at most (e^{-40} 40^3 / 6)
        static const double far_lr = [] {
            const char* v = getenv("MPMC_B200_SYM_FAR_LR");
            const double x = v ? atof(v) : 48.0;
            return x >= 30.0 && x <= 48.0 ? x : 48.0;
        }();
        const double rf = far_lr / l * (1.0 + 1e-12);
        k.far2 = rf * rf;
    }
    {
        double f = 1.0;
        for (int q = 0; q < 12; q++) {
            k.ec[q] = 1.0 / f;
            f *= (double)(q + 1);
        }
        // largest minimum-image separation is below half the sum of the lattice vector lengths
        double rmax = 0.0;
        for (int q = 0; q < 3; q++) rmax += 0.5 * sqrt(c.B[q][0] * c.B[q][0] + c.B[q][1] * c.B[q][1] + c.B[q][2] * c.B[q][2]);
        k.fastexp = (l >= 0.0 && l * rmax < 700.0) ? 1 : 0;
    }
    if (const char* v = getenv("MPMC_B200_SYM_FAR")) {  // A/B timing: "off" evaluates the damping of every pair (same bits)
        if (!strcmp(v, "off")) k.far2 = 0.0;
#ifdef MPMC_B200_TIMING_EXPERIMENTS  // never in the product build: treating every cluster pair as far gives WRONG fields
        if (!strcmp(v, "all") && k.far2 > 0.0) k.far2 = 1e-300;
#endif
    }
    k.box = box_len(c);
    return k;
}

}  // namespace

// The partial rows a launch owns, per super tile (see dipole_sym_reduce_kernel), cached on the device per
// (device, layout): [S + 1] CSR offsets followed by the entries.  The layout changes only with the padded atom count
// or the shard, so the steady state is a map lookup.
struct ReduceListKey {
    int dev, S, K, y0, Y, kper, hsplit, rank, world;
    bool operator<(const ReduceListKey& o) const {
        return std::tie(dev, S, K, y0, Y, kper, hsplit, rank, world) <
               std::tie(o.dev, o.S, o.K, o.y0, o.Y, o.kper, o.hsplit, o.rank, o.world);
    }
};
// slices [y0, y0 + Y) of a contraction with offsets 0..K in slices of kper
static const int* sym_reduce_list(int S, int K, int y0, int Y, int kper, int hsplit, int rank, int world, cudaStream_t st) {
    static std::map<ReduceListKey, int*> cache;
    static std::mutex mu;
    std::lock_guard<std::mutex> lock(mu);
    ReduceListKey key;
    cudaGetDevice(&key.dev);
    key.S = S, key.K = K, key.y0 = y0, key.Y = Y, key.kper = kper, key.hsplit = hsplit, key.rank = rank, key.world = world;
    auto it = cache.find(key);
    if (it != cache.end()) return it->second;
    if (cache.size() > 64) {  // uVT runs walk through many sizes: drop the old layouts (no launch is in flight on them
        for (auto& kv : cache) {  // once the stream has been synchronised)
            cudaSetDevice(kv.first.dev);
            cudaDeviceSynchronize();
            cudaFree(kv.second);
        }
        cache.clear();
        cudaSetDevice(key.dev);
    }
    std::vector<int> offs(S + 1, 0), rows;
    const int jrow0 = std::max(y0 * kper, 1), kend = std::min((y0 + Y) * kper - 1, K);
    for (int J = 0; J < S; J++) {
        offs[J] = (int)rows.size();
        for (int y = 0; y < Y * hsplit; y++)  // I side: item (J, y0 + y / hsplit) wrote row y
            if (((y / hsplit) * S + J) % world == rank) rows.push_back(y);
        for (int kk = jrow0; kk <= kend; kk++) {  // J side: item (J - kk, slice of kk) wrote row kk - jrow0
            int I = J - kk;
            if (I < 0) I += S;
            if (((kk / kper - y0) * S + I) % world != rank) continue;
            if (2 * kk == S && I >= kk) continue;  // even S: the half-way offset belongs to I < S/2
            rows.push_back((kk - jrow0) | (1 << 30));
        }
    }
    offs[S] = (int)rows.size();
    std::vector<int> all(offs);
    all.insert(all.end(), rows.begin(), rows.end());
    int* dev = nullptr;
    if (cudaMalloc(&dev, sizeof(int) * all.size()) != cudaSuccess) return nullptr;
    // pageable source: the copy is staged before the call returns, and stream order puts it ahead of the reduce kernel
    if (cudaMemcpyAsync(dev, all.data(), sizeof(int) * all.size(), cudaMemcpyHostToDevice, st) != cudaSuccess) {
        cudaFree(dev);
        return nullptr;
    }
    cudaStreamSynchronize(st);
    cache[key] = dev;
    return dev;
}

void launch_cluster_boxes(const Cell& c, bool ortho, const SubsetDev& pol, double* cbox, cudaStream_t s) {
    const int nclus = pol.n_pad / 32;
    if (nclus <= 0) return;
    cluster_box_kernel<<<(nclus * 32 + 255) / 256, 256, 0, s>>>(pol.n, pol.n_pad, c.B[0][0], c.B[1][1], c.B[2][2], c.R[0][0],
                                                                c.R[1][1], c.R[2][2], ortho ? 1 : 0, pol.x, pol.y, pol.z, cbox);
}

void launch_peer_signal(const PeerSignal& sig, cudaStream_t s) {
    if (sig.world > 1) peer_signal_kernel<<<1, 32, 0, s>>>(sig);
}

int dipole_sym_tile() { return STILE; }

// J-side partial rows are one [3][n_pad] row per offset of a launch: N^2/512 doubles for a whole contraction (102 MB at
// 63k atoms, 2 GB at 256k, 17 GB at 1M).  Above MPMC_B200_SYM_PART_MB (default 1024) a contraction is cut into launches
// over slice ranges that reuse one buffer of that size, so the resident image stays O(N) + the cap.
static size_t sym_part_cap_bytes() {
    static const size_t cap = [] {
        const char* v = getenv("MPMC_B200_SYM_PART_MB");
        const double mb = v ? atof(v) : 1024.0;
        return (size_t)((mb > 0.0 ? mb : 1024.0) * 1048576.0);
    }();
    return cap;
}
static int sym_slices_per_launch(int n_pad, int kper, int Y) {
    const size_t row_bytes = (size_t)3 * n_pad * sizeof(double);
    const size_t rows_cap = std::max<size_t>(sym_part_cap_bytes() / row_bytes, 1);
    int spr = (int)std::min<size_t>(rows_cap / (size_t)kper, (size_t)Y);
    return spr < 1 ? 1 : spr;
}
int dipole_sym_partial_rows(int n_pad, int ysplits) {
    const int S = n_pad / STILE, K = S / 2;
    if (ysplits < 1) ysplits = 1;
    if (ysplits > K + 1) ysplits = K + 1;
    const int kper = (K + 1 + ysplits - 1) / ysplits;
    const int Y = (K + 1 + kper - 1) / kper;
    const int rows = sym_slices_per_launch(n_pad, kper, Y) * kper;
    return rows > 1 ? rows : 1;
}

// gridDim.z of a contraction's launches (see SymArgs::hsplit); the I-side partial rows are [hsplit * ysplits][3][n_pad]
int dipole_sym_hsplit(int n_pad, int ysplits, int world) {
    static const int forced = [] {
        const char* v = getenv("MPMC_B200_SYM_HSPLIT");
        const int h = v ? atoi(v) : 0;
        return (h == 1 || h == 2 || h == 4 || h == 8) ? h : 0;
    }();
    if (forced) return forced;
    const int S = n_pad / STILE, K = S / 2;
    if (S <= 0) return 2;
    ysplits = std::max(1, std::min(ysplits, K + 1));
    const int kper = (K + 1 + ysplits - 1) / ysplits, Y = (K + 1 + kper - 1) / kper;
    int sms = 148, dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const long mine = ((long)S * Y + std::max(world, 1) - 1) / std::max(world, 1);
    // measured on one GPU (call 73, ms per energy() with gridDim.z 2 / 4 / 8): 4.8k polarizable atoms 1.66 / 1.54 / 1.49, 5.2k
    // 1.44 / 1.30 / 1.25, 7.2k 2.70 / 2.67 / 2.69, 9.6k 4.23 / 4.11 / 4.13, 15k 2.36 / 2.35 / 2.45: finer while the launch is
    // below ~2.75 waves of resident CTAs
    const long limit = 11L * sms * SYM_MINB / 4;
    int h = 2;
    while (h < SYM_HSPLIT_MAX && mine * h < limit) h *= 2;
    return h;
}

void launch_dipole_contract_sym(const Cell& c, bool ortho, const SubsetDev& pol, const DipoleOpts& o, const double* mu,
                                const double* cbox, int rank, int world, int ysplits, double* epartI,
                                double* epartJ, double* esum, cudaStream_t s, const PeerSignal* sig) {
    const int np = pol.n_pad, S = np / STILE, K = S / 2;
    if (S <= 0) return;
    const SymConsts k = make_sym_consts(c, ortho, o);
    if (ysplits < 1) ysplits = 1;
    if (ysplits > K + 1) ysplits = K + 1;
    SymArgs a;
    a.n = pol.n;
    a.np = np;
    a.S = S;
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
    int sms = 148;
    {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
    // Splitting the J halves of a super tile over gridDim.z halves the work item.  It pays where a rank's items make
    // fewer than ~8 waves of resident CTAs -- 4 and 8 GPUs at 63k atoms: 0.978 -> 0.951 and 0.528 -> 0.501 ms per
    // contraction (profiles/r02_multi_gpu.md; in round 1 it lost 8 % because the partial-row reduce then cost twice as
    // much, which the list-driven reduce no longer does) -- and nowhere else.
    // Round 2, end: with the list-driven reduce the half items also pay on ONE GPU -- 9.6k polarizable atoms 0.283 -> 0.253 ms
    // per contraction (1.3 -> 2.6 waves), 15k -6 %, 31k -3 %, 46k (mof_h2_63k, 9.5 -> 19 waves) 3.674 -> 3.597 ms
    // (profiles/r02_k4_hsplit.md) -- so they are the default everywhere; MPMC_B200_SYM_HSPLIT=1 restores whole items.
    // Last change of round 2: launches of less than two waves of half items (4k-12k polarizable atoms on one GPU) are
    // bound by the single item; there the half is cut further into groups of its j-clusters (profiles/r02_sym_min.md).
    a.hsplit = dipole_sym_hsplit(np, ysplits, world);
    (void)sms;
    a.unishift = 1;  // MPMC_B200_SYM_UNISHIFT=0: minimum image per pair everywhere (A/B)
    if (const char* v = getenv("MPMC_B200_SYM_UNISHIFT")) a.unishift = atoi(v) != 0 ? 1 : 0;
    // Sharding unit: the work item (I tile, offset slice y) -- every item has the same number of pairs; the items are
    // dealt out round robin over their y-major order, so the ranks differ by at most one item and hold the same mix of
    // damped and far tile pairs.  One CTA per owned item, in that order (heavy near-diagonal slices first).
    a.rank = rank;
    a.world = world;
    const int spr = sym_slices_per_launch(np, a.kper, ysplits);
    for (int y0 = 0, round = 0; y0 < ysplits; y0 += spr, round++) {
        const int Yr = std::min(spr, ysplits - y0);
        const int items = S * Yr;
        a.Y = Yr;
        a.y0 = y0;
        a.jrow0 = std::max(y0 * a.kper, 1);
        const int mine = items > rank ? (items - rank + world - 1) / world : 0;
        if (mine > 0) {
            dim3 grid(mine, 1, a.hsplit);
#define LAUNCH(ORTHO, DAMP, WF, FE) dipole_sym_kernel<ORTHO, DAMP, WF, FE><<<grid, TILE, 0, s>>>(c, k, a)
#define DISPATCH(ORTHO)                                                         \
    if (o.damp_type == MPMC_DAMP_EXPONENTIAL) {                                 \
        if (k.fastexp) {                                                        \
            if (o.wolf_full) LAUNCH(ORTHO, MPMC_DAMP_EXPONENTIAL, true, true);  \
            else LAUNCH(ORTHO, MPMC_DAMP_EXPONENTIAL, false, true);             \
        } else {                                                                \
            if (o.wolf_full) LAUNCH(ORTHO, MPMC_DAMP_EXPONENTIAL, true, false); \
            else LAUNCH(ORTHO, MPMC_DAMP_EXPONENTIAL, false, false);            \
        }                                                                       \
    } else if (o.damp_type == MPMC_DAMP_LINEAR) {                               \
        LAUNCH(ORTHO, MPMC_DAMP_LINEAR, false, false);                          \
    } else {                                                                    \
        if (o.wolf_full) LAUNCH(ORTHO, MPMC_DAMP_OFF, true, false);             \
        else LAUNCH(ORTHO, MPMC_DAMP_OFF, false, false);                        \
    }
            if (ortho) {
                DISPATCH(true)
            } else {
                DISPATCH(false)
            }
#undef DISPATCH
#undef LAUNCH
        }
        const int* list = sym_reduce_list(S, K, y0, Yr, a.kper, a.hsplit, rank, world, s);
        if (list) dipole_sym_reduce_kernel<<<np / 32, 256, 0, s>>>(np, list, list + S + 1, epartI, epartJ, round > 0 ? 1 : 0, esum);
    }
    if (sig && sig->world > 1) peer_signal_kernel<<<1, 32, 0, s>>>(*sig);
}

}  // namespace mpmc
