// dipole_kernel.cu -- K4 matrix-free dipole-field-tensor contraction, K5 dipole update, and the
// blocked Gauss-Seidel sweep, over the compacted list of polarizable atoms.
//
// The reference builds a dense 3N x 3N matrix A (thole_amatrix, src/polarization/thole_matrix.c:38-146)
// and streams it once per iteration (contract_dipoles, src/polarization/thole_iterative.c:27-59).
// Here T_ij is recomputed in registers from the two positions every time it is needed:
//     T_ij = delta_pq * damp1 / r^3  -  3 d_p d_q * damp2 / r^5          (thole_matrix.c:118-131)
//     E_ind,i = - sum_{j != i} T_ij mu_j                                  (thole_iterative.c:40-45)
// with exponential damping (thole_matrix.c:108-116)
//     damp1 = 1 - e^{-lr} (l^2 r^2 / 2 + l r + 1),   damp2 = damp1 - e^{-lr} l^3 r^3 / 6
// and the linear / off / wolf-full variants of the same switch.  All pairs, minimum image, no
// cutoff, no exclusions -- exactly the reference's loop bounds.  Atoms with alpha == 0 carry mu == 0
// in the reference and are simply not in the list.
#include <cooperative_groups.h>

#include "kernels.h"

namespace mpmc {

struct DipoleConsts {
    double l, l2h, l3s;       // l, l^2/2, l^3/6
    double wd1, wd2, ircut3;  // wolf-full: wdamp1/rcut^3, 3*wdamp2/rcut^3 (0 when off)
    double ec[12];            // 1/k! for exp_neg_fast
    int fastexp;              // l * (largest possible r) < 700: exp_neg_fast is valid
};

// contribution of source (sx,sy,sz,mu) to the field at (xi,yi,zi); aux_i/aux_j: alpha^(1/6) (linear)
// or charge (damping off); same_mol only matters for damping off.
// c1 = damp1 / r^3 (delta_pq coefficient), c2 = 3 damp2 / r^5 (d_p d_q coefficient) of T_ij from r^2
template <int DAMP, bool WF>
__device__ __forceinline__ void tensor_coeff(const DipoleConsts& k, double r2, double aux_i, double aux_j, bool same_mol,
                                             double& c1, double& c2) {
    const double ir = rsqrt_fast(r2);
    const double r = r2 * ir;
    const double ir2 = ir * ir;
    const double ir3 = ir * ir2;
    double damp1, damp2;
    if (DAMP == MPMC_DAMP_EXPONENTIAL) {
        const double lr = k.l * r;
        const double e = k.fastexp ? exp_neg_fast(k, lr) : exp(-lr);
        damp1 = 1.0 - e * (k.l2h * r2 + lr + 1.0);
        damp2 = damp1 - e * (k.l3s * r2 * r);
    } else if (DAMP == MPMC_DAMP_LINEAR) {
        const double s = k.l * (aux_i * aux_j);
        const double v = r / s;
        if (r < s) {
            const double v3 = v * v * v;
            damp1 = (4.0 - 3.0 * v) * v3;
            damp2 = v3 * v;
        } else {
            damp1 = damp2 = 1.0;
        }
    } else {  // off: es_excluded pairs are switched off entirely (thole_matrix.c:92-97)
        const bool excl = same_mol || aux_i == 0.0 || aux_j == 0.0;
        damp1 = damp2 = excl ? 0.0 : 1.0;
    }
    c1 = damp1 * ir3;
    c2 = 3.0 * damp2 * ir3 * ir2;
    if (WF) {                              // thole_matrix.c:123-130
        if (DAMP == MPMC_DAMP_OFF) {
            c1 -= damp1 * k.ircut3;
            c2 -= 3.0 * damp2 * ir2 * k.ircut3;
        } else {
            c1 -= k.wd1;
            c2 -= k.wd2 * ir2;
        }
    }
}

template <bool ORTHO, int DAMP, bool WF>
__device__ __forceinline__ void tensor_apply(const Cell& c, const DipoleConsts& k, double dx, double dy, double dz,
                                             double mx, double my, double mz, double aux_i, double aux_j, bool same_mol,
                                             double& ex, double& ey, double& ez) {
    min_image<ORTHO>(c, dx, dy, dz);
    const double r2 = dx * dx + dy * dy + dz * dz;
    double c1, c2;
    tensor_coeff<DAMP, WF>(k, r2, aux_i, aux_j, same_mol, c1, c2);
    const double dm = (dx * mx + dy * my + dz * mz) * c2;
    ex += dm * dx - c1 * mx;
    ey += dm * dy - c1 * my;
    ez += dm * dz - c1 * mz;
}

template <bool ORTHO, int DAMP, bool WF>
__global__ void __launch_bounds__(TILE) dipole_contract_kernel(Cell c, SubsetDev p, DipoleConsts k,
                                                               const double* __restrict__ aux,
                                                               const double* __restrict__ mu, int row_begin, int row_end,
                                                               int tiles_per_split, int skip_tile,
                                                               double* __restrict__ epart) {
    __shared__ double sx[TILE], sy[TILE], sz[TILE], smx[TILE], smy[TILE], smz[TILE], saux[TILE];
    __shared__ int smol[TILE];
    const int tid = threadIdx.x;
    const int i = row_begin + blockIdx.x * TILE + tid;  // rows are tile aligned; i < n_pad always
    const double xi = p.x[i], yi = p.y[i], zi = p.z[i];
    const double auxi = (DAMP == MPMC_DAMP_EXPONENTIAL) ? 0.0 : aux[i];
    const int moli = (DAMP == MPMC_DAMP_OFF) ? p.mol[i] : 0;
    const int np = p.n_pad;
    double ex = 0.0, ey = 0.0, ez = 0.0;
    const int ntiles = np / TILE;
    const int t0 = blockIdx.y * tiles_per_split, t1 = min(t0 + tiles_per_split, ntiles);
    for (int jt = t0; jt < t1; jt++) {
        if (jt == skip_tile) continue;
        __syncthreads();
        {
            const int j = jt * TILE + tid;
            sx[tid] = p.x[j];
            sy[tid] = p.y[j];
            sz[tid] = p.z[j];
            smx[tid] = mu[j];
            smy[tid] = mu[np + j];
            smz[tid] = mu[2 * np + j];
            if (DAMP != MPMC_DAMP_EXPONENTIAL) saux[tid] = aux[j];
            if (DAMP == MPMC_DAMP_OFF) smol[tid] = p.mol[j];
        }
        __syncthreads();
        const bool checked = (jt * TILE == (i - tid)) || (jt == ntiles - 1);  // own tile, or the padded last tile
        if (checked) {
            const int jn = min(TILE, p.n - jt * TILE);
            for (int jj = 0; jj < jn; jj++) {
                if (jt * TILE + jj == i) continue;
                tensor_apply<ORTHO, DAMP, WF>(c, k, xi - sx[jj], yi - sy[jj], zi - sz[jj], smx[jj], smy[jj], smz[jj], auxi,
                                              saux[jj], moli == smol[jj], ex, ey, ez);
            }
        } else {
#pragma unroll 4
            for (int jj = 0; jj < TILE; jj++)
                tensor_apply<ORTHO, DAMP, WF>(c, k, xi - sx[jj], yi - sy[jj], zi - sz[jj], smx[jj], smy[jj], smz[jj], auxi,
                                              saux[jj], moli == smol[jj], ex, ey, ez);
        }
    }
    double* e = epart + (size_t)blockIdx.y * 3 * np;
    e[i] = ex;
    e[np + i] = ey;
    e[2 * np + i] = ez;
}

static DipoleConsts make_consts(const Cell& c, const DipoleOpts& o) {
    DipoleConsts k;
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
    fill_exp_coeffs(k.ec);
    double rmax = 0.0;  // largest minimum-image separation is below half the sum of the lattice vector lengths
    for (int q = 0; q < 3; q++) rmax += 0.5 * sqrt(c.B[q][0] * c.B[q][0] + c.B[q][1] * c.B[q][1] + c.B[q][2] * c.B[q][2]);
    k.fastexp = (l >= 0.0 && l * rmax < 700.0) ? 1 : 0;
    return k;
}

static void contract_dispatch(const Cell& c, bool ortho, const SubsetDev& pol, const DipoleOpts& o, const double* mu,
                              int row_begin, int row_end, int jsplits, int skip_tile, double* epart, cudaStream_t s) {
    const int rows = row_end - row_begin;
    if (rows <= 0) return;
    const DipoleConsts k = make_consts(c, o);
    const int ntiles = pol.n_pad / TILE;
    const int tps = (ntiles + jsplits - 1) / jsplits;
    dim3 grid(rows / TILE, jsplits);
    const double* aux = pol.aux;
#define LAUNCH(ORTHO, DAMP, WF) \
    dipole_contract_kernel<ORTHO, DAMP, WF><<<grid, TILE, 0, s>>>(c, pol, k, aux, mu, row_begin, row_end, tps, skip_tile, epart)
#define DISPATCH(ORTHO)                                                        \
    if (o.damp_type == MPMC_DAMP_EXPONENTIAL) {                                \
        if (o.wolf_full) LAUNCH(ORTHO, MPMC_DAMP_EXPONENTIAL, true);           \
        else LAUNCH(ORTHO, MPMC_DAMP_EXPONENTIAL, false);                      \
    } else if (o.damp_type == MPMC_DAMP_LINEAR) {                              \
        LAUNCH(ORTHO, MPMC_DAMP_LINEAR, false); /* wdamp stays 0 for linear */ \
    } else {                                                                   \
        if (o.wolf_full) LAUNCH(ORTHO, MPMC_DAMP_OFF, true);                   \
        else LAUNCH(ORTHO, MPMC_DAMP_OFF, false);                              \
    }
    if (ortho) {
        DISPATCH(true)
    } else {
        DISPATCH(false)
    }
#undef DISPATCH
#undef LAUNCH
}

// Field of the dipoles at atoms that are NOT in the polarizable list (alpha == 0).  The reference's
// palmo_contraction() (thole_iterative.c:119-141) runs over every atom, so non-polarizable atoms
// get ef_induced_change = - sum_j T_ij mu_j as well (their rows of A exist, thole_matrix.c:73-143);
// the values never enter an energy (mu_i = 0) but are written back to atom_t.
// blockIdx.y = slice of the source tiles (tiles_per_split each): with ~140 row tiles at 63k atoms an unsplit launch left
// every SM with one 4-warp CTA walking 46k sources (7 ms); the slices' partial rows are added in slice order.
template <bool ORTHO, int DAMP, bool WF>
__global__ void __launch_bounds__(TILE) dipole_field_at_kernel(Cell c, SubsetDev rows, SubsetDev p, DipoleConsts k,
                                                               const double* __restrict__ mu, int row_begin,
                                                               int tiles_per_split, double* __restrict__ out) {
    __shared__ double sx[TILE], sy[TILE], sz[TILE], smx[TILE], smy[TILE], smz[TILE], saux[TILE];
    __shared__ int smol[TILE];
    const int tid = threadIdx.x;
    const int i = row_begin + blockIdx.x * TILE + tid;
    const double xi = rows.x[i], yi = rows.y[i], zi = rows.z[i];
    const double auxi = rows.aux[i];
    const int moli = rows.mol[i];
    const int np = p.n_pad;
    double ex = 0.0, ey = 0.0, ez = 0.0;
    const int t0 = blockIdx.y * tiles_per_split, t1 = min(t0 + tiles_per_split, np / TILE);
    for (int jt = t0; jt < t1; jt++) {
        __syncthreads();
        const int j = jt * TILE + tid;
        sx[tid] = p.x[j];
        sy[tid] = p.y[j];
        sz[tid] = p.z[j];
        smx[tid] = mu[j];
        smy[tid] = mu[np + j];
        smz[tid] = mu[2 * np + j];
        saux[tid] = p.aux[j];
        smol[tid] = p.mol[j];
        __syncthreads();
        const int jn = min(TILE, p.n - jt * TILE);
#pragma unroll 2
        for (int jj = 0; jj < jn; jj++)
            tensor_apply<ORTHO, DAMP, WF>(c, k, xi - sx[jj], yi - sy[jj], zi - sz[jj], smx[jj], smy[jj], smz[jj], auxi,
                                          saux[jj], moli == smol[jj], ex, ey, ez);
    }
    const bool live = i < rows.n;
    double* o = out + (size_t)blockIdx.y * 3 * rows.n_pad;
    o[i] = live ? ex : 0.0;
    o[rows.n_pad + i] = live ? ey : 0.0;
    o[2 * rows.n_pad + i] = live ? ez : 0.0;
}

// out[c][i] = sum over slices (slice order) of part[slice][c][i], rows [row_begin, row_end)
__global__ void field_at_reduce_kernel(const double* __restrict__ part, int splits, int np, int row_begin, int row_end,
                                       double* __restrict__ out) {
    const int i = row_begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= row_end) return;
    for (int cdim = 0; cdim < 3; cdim++) {
        double e = 0.0;
        for (int js = 0; js < splits; js++) e += part[((size_t)js * 3 + cdim) * np + i];
        out[(size_t)cdim * np + i] = e;
    }
}

int dipole_field_at_splits(int row_tiles, int src_tiles) {
    int sms = 148, dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int js = (4 * sms + row_tiles - 1) / (row_tiles > 0 ? row_tiles : 1);
    if (js > src_tiles) js = src_tiles;
    return js < 1 ? 1 : js;
}

void launch_dipole_field_at(const Cell& c, bool ortho, const SubsetDev& rows, const SubsetDev& pol, const DipoleOpts& o,
                            const double* mu, int row_begin, int row_end, double* out, cudaStream_t s, double* part) {
    const int n = row_end - row_begin;
    if (n <= 0 || pol.n <= 0) return;
    const DipoleConsts k = make_consts(c, o);
    const int src_tiles = pol.n_pad / TILE;
    const int js = part ? dipole_field_at_splits(n / TILE, src_tiles) : 1;
    const int tps = (src_tiles + js - 1) / js;
    const dim3 grid(n / TILE, (src_tiles + tps - 1) / tps);
    double* dst = part ? part : out;
#define LAUNCH(ORTHO, DAMP, WF) \
    dipole_field_at_kernel<ORTHO, DAMP, WF><<<grid, TILE, 0, s>>>(c, rows, pol, k, mu, row_begin, tps, dst)
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
    if (part) field_at_reduce_kernel<<<(n + 255) / 256, 256, 0, s>>>(part, (int)grid.y, rows.n_pad, row_begin, row_end, out);
}

void launch_dipole_contract(const Cell& c, bool ortho, const SubsetDev& pol, const DipoleOpts& o, const double* mu,
                            int row_begin, int row_end, int jsplits, double* epart, cudaStream_t s) {
    contract_dispatch(c, ortho, pol, o, mu, row_begin, row_end, jsplits, -1, epart, s);
}

// ------------------------------------------------------------------------------------------------
// K5
// ------------------------------------------------------------------------------------------------
__global__ void dipole_init_kernel(SubsetDev p, const double* __restrict__ ef_static, int atoms_n_pad, double gscale,
                                   double* __restrict__ es, double* __restrict__ mu, int row_begin, int row_end,
                                   const double* __restrict__ mu_start) {
    const int i = row_begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= row_end) return;
    const int np = p.n_pad;
    const bool live = i < p.n;
    const int g = live ? p.idx[i] : 0;
    const double al = live ? p.alpha[i] : 0.0;
    for (int cdim = 0; cdim < 3; cdim++) {
        const double e = live ? ef_static[cdim * atoms_n_pad + g] : 0.0;
        es[cdim * np + i] = e;
        // thole_iterative.c:19-21; opt-in warm start: the converged dipoles of the previous evaluation (atom space)
        mu[cdim * np + i] = (mu_start && live) ? mu_start[cdim * atoms_n_pad + g] : (al * e) * gscale;
    }
}

// converged dipoles -> atom space, kept as the starting point of the next solve (opt-in warm start)
__global__ void dipole_keep_kernel(SubsetDev p, const double* __restrict__ mu, int atoms_n_pad, double* __restrict__ mu_keep) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.n) return;
    const int g = p.idx[i];
    for (int cdim = 0; cdim < 3; cdim++) mu_keep[cdim * atoms_n_pad + g] = mu[cdim * p.n_pad + i];
}

__global__ void dipole_update_kernel(int n, int np, const double* __restrict__ alpha, UpdateOpts o,
                                     const double* __restrict__ epart, int jsplits, const double* __restrict__ es,
                                     const double* __restrict__ mu, double* __restrict__ mu_old,
                                     double* __restrict__ mu_new, double* __restrict__ eind, double* __restrict__ rrms,
                                     unsigned long long* __restrict__ max_err2_bits, int row_begin, int row_end,
                                     double* mu_commit) {
    __shared__ double red[32];
    const int i = row_begin + blockIdx.x * blockDim.x + threadIdx.x;
    double worst = 0.0;
    if (i < row_end && i < n) {
        const double al = alpha[i];
        double num = 0.0, den = 0.0;
        const double wold = (mu_commit && o.esor) ? exp(-o.gamma * o.iteration) : 0.0;
        for (int cdim = 0; cdim < 3; cdim++) {
            double e = 0.0;
            for (int js = 0; js < jsplits; js++) e += epart[((size_t)js * 3 + cdim) * np + i];
            eind[cdim * np + i] = e;
            const double old = mu[cdim * np + i];
            if (o.save_old) mu_old[cdim * np + i] = old;
            const double nw = al * (es[cdim * np + i] + e);  // thole_iterative.c:49
            mu_new[cdim * np + i] = nw;
            const double d = nw - old;
            worst = fmax(worst, d * d);  // are_we_done_yet, thole_iterative.c:108-111
            num += d * d;
            den += nw * nw;
            if (mu_commit) {  // commit fused (no palmo pass needs the old dipoles): thole_iterative.c:242-252
                double v = nw;
                if (o.sor)
                    v = o.gamma * nw + (1.0 - o.gamma) * old;
                else if (o.esor)
                    v = (1.0 - wold) * nw + wold * old;
                mu_commit[cdim * np + i] = v;
            }
        }
        if (o.do_rrms) {  // calc_dipole_rrms, thole_iterative.c:69-77
            double v = sqrt(num / den);
            if (!isfinite(v)) v = 0.0;
            rrms[i] = v;
        }
    }
    worst = block_reduce<2>(worst, red);
    // non-negative doubles order like their bit patterns => integer atomicMax is exact and order independent
    if (threadIdx.x == 0 && max_err2_bits) atomicMax(max_err2_bits, (unsigned long long)__double_as_longlong(worst));
}

__global__ void dipole_commit_kernel(int n, int np, UpdateOpts o, const double* __restrict__ mu_old,
                                     const double* __restrict__ mu_new, double* __restrict__ mu, int row_begin,
                                     int row_end) {
    const int i = row_begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= row_end || i >= n) return;
    for (int cdim = 0; cdim < 3; cdim++) {
        const double nw = mu_new[cdim * np + i];
        double v;
        if (o.sor)
            v = o.gamma * nw + (1.0 - o.gamma) * mu_old[cdim * np + i];
        else if (o.esor) {
            const double w = exp(-o.gamma * o.iteration);
            v = (1.0 - w) * nw + w * mu_old[cdim * np + i];
        } else
            v = nw;
        mu[cdim * np + i] = v;
    }
}

// ---- peer-memory forms (see PeerWait in kernels.h) ----------------------------------------------
__device__ __forceinline__ void peer_wait(const PeerWait& pw) {
    if (threadIdx.x == 0) {
        unsigned long long t0, t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        for (int r = 0; r < pw.world; r++) {
            for (;;) {
                unsigned long long v;
                asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(pw.my_flags + r) : "memory");
                if (v >= pw.seq) break;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                if (t1 - t0 > 5000000000ull) {  // 5 s: a peer died; fail the evaluation instead of hanging the GPU
                    *pw.timeout_flag = 1.0;
                    break;
                }
            }
        }
    }
    __syncthreads();
}

__global__ void __launch_bounds__(256) dipole_update_peer_kernel(int n, int np, const double* __restrict__ alpha, UpdateOpts o,
                                                                 PeerWait pw, const double* __restrict__ es,
                                                                 double* __restrict__ mu, double* __restrict__ mu_old,
                                                                 double* __restrict__ mu_new, double* __restrict__ eind,
                                                                 double* __restrict__ rrms,
                                                                 unsigned long long* __restrict__ max_err2_bits, int commit) {
    __shared__ double red[32];
    peer_wait(pw);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    double worst = 0.0;
    if (i < n) {
        const double al = alpha[i];
        double num = 0.0, den = 0.0;
        const double wold = (commit && o.esor) ? exp(-o.gamma * o.iteration) : 0.0;
        for (int cdim = 0; cdim < 3; cdim++) {
            double e = 0.0;
            for (int r = 0; r < pw.world; r++) e += __ldcv(pw.src[r] + (size_t)cdim * np + i);  // rank order, L1 bypassed
            eind[cdim * np + i] = e;
            const double old = mu[cdim * np + i];
            if (o.save_old) mu_old[cdim * np + i] = old;
            const double nw = al * (es[cdim * np + i] + e);  // thole_iterative.c:49
            mu_new[cdim * np + i] = nw;
            const double d = nw - old;
            worst = fmax(worst, d * d);
            num += d * d;
            den += nw * nw;
            if (commit) {  // thole_iterative.c:242-252
                double v = nw;
                if (o.sor)
                    v = o.gamma * nw + (1.0 - o.gamma) * old;
                else if (o.esor)
                    v = (1.0 - wold) * nw + wold * old;
                mu[cdim * np + i] = v;
            }
        }
        if (o.do_rrms) {
            double v = sqrt(num / den);
            if (!isfinite(v)) v = 0.0;
            rrms[i] = v;
        }
    }
    worst = block_reduce<2>(worst, red);
    if (threadIdx.x == 0 && max_err2_bits) atomicMax(max_err2_bits, (unsigned long long)__double_as_longlong(worst));
}

__global__ void __launch_bounds__(256) palmo_finish_peer_kernel(int n, int np, PeerWait pw, const double* __restrict__ eind,
                                                                double* __restrict__ eic) {
    peer_wait(pw);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    for (int cdim = 0; cdim < 3; cdim++) {
        double e = 0.0;
        for (int r = 0; r < pw.world; r++) e += __ldcv(pw.src[r] + (size_t)cdim * np + i);
        eic[cdim * np + i] = -eind[cdim * np + i] + e;  // thole_iterative.c:129-130
    }
}

// out[i] = sum_r src[r][i] in rank order: the all-gather of row-owned data (the static field: every row is written by
// exactly one rank and zero elsewhere, so the sum IS the gather and every rank gets the same bits)
__global__ void __launch_bounds__(256) peer_sum_kernel(PeerWait pw, size_t n, double* __restrict__ out) {
    peer_wait(pw);
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double e = 0.0;
    for (int r = 0; r < pw.world; r++) e += __ldcv(pw.src[r] + i);
    out[i] = e;
}

// The device scalars of an evaluation combined over the ranks, in rank order (identical bits on every rank):
// sums, except slot `min_slot` (smallest intermolecular distance) which takes the minimum.
__global__ void peer_scalars_kernel(PeerWait pw, int n, int min_slot, double* __restrict__ out) {
    peer_wait(pw);
    const int i = threadIdx.x;
    if (i >= n) return;
    double e = __ldcv(pw.src[0] + i);
    for (int r = 1; r < pw.world; r++) {
        const double v = __ldcv(pw.src[r] + i);
        e = (i == min_slot) ? fmin(e, v) : e + v;
    }
    out[i] = e;
}

__global__ void palmo_finish_kernel(int n, int np, const double* __restrict__ epart, int jsplits,
                                    const double* __restrict__ eind, double* __restrict__ eic, int row_begin, int row_end) {
    const int i = row_begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= row_end || i >= n) return;
    for (int cdim = 0; cdim < 3; cdim++) {
        double e = -eind[cdim * np + i];  // thole_iterative.c:129-130
        for (int js = 0; js < jsplits; js++) e += epart[((size_t)js * 3 + cdim) * np + i];
        eic[cdim * np + i] = e;
    }
}

__global__ void __launch_bounds__(1024) polar_energy_kernel(int n, int np, const double* __restrict__ mu,
                                                            const double* __restrict__ es, const double* __restrict__ eic,
                                                            const double* __restrict__ rrms, int row_begin, int row_end,
                                                            double* __restrict__ out) {
    __shared__ double red[32];
    double u = 0.0, up = 0.0, rr = 0.0;
    for (int i = row_begin + threadIdx.x; i < row_end && i < n; i += blockDim.x) {
        for (int cdim = 0; cdim < 3; cdim++) {
            const double m = mu[cdim * np + i];
            u += m * es[cdim * np + i];
            if (eic) up += m * eic[cdim * np + i];
        }
        if (rrms) {
            const double v = rrms[i];
            if (isfinite(v)) rr += v;
        }
    }
    double v = block_reduce<0>(u, red);
    if (threadIdx.x == 0) out[0] = v;
    v = block_reduce<0>(up, red);
    if (threadIdx.x == 0) out[1] = v;
    v = block_reduce<0>(rr, red);
    if (threadIdx.x == 0) out[2] = v;
}

void launch_dipole_init(const SubsetDev& pol, const double* ef_static, int atoms_n_pad, double gamma_scale, double* es,
                        double* mu, int row_begin, int row_end, cudaStream_t s, const double* mu_start) {
    const int rows = row_end - row_begin;
    if (rows <= 0) return;
    dipole_init_kernel<<<(rows + 255) / 256, 256, 0, s>>>(pol, ef_static, atoms_n_pad, gamma_scale, es, mu, row_begin, row_end,
                                                          mu_start);
}

void launch_dipole_keep(const SubsetDev& pol, const double* mu, int atoms_n_pad, double* mu_keep, cudaStream_t s) {
    if (pol.n <= 0) return;
    dipole_keep_kernel<<<(pol.n + 255) / 256, 256, 0, s>>>(pol, mu, atoms_n_pad, mu_keep);
}

void launch_dipole_update(const SubsetDev& pol, const UpdateOpts& o, const double* epart, int jsplits, const double* es,
                          const double* mu, double* mu_old, double* mu_new, double* eind, double* rrms,
                          unsigned long long* max_err2_bits, int row_begin, int row_end, cudaStream_t s, double* mu_commit) {
    const int rows = row_end - row_begin;
    if (rows <= 0) return;
    dipole_update_kernel<<<(rows + 255) / 256, 256, 0, s>>>(pol.n, pol.n_pad, pol.alpha, o, epart, jsplits, es, mu, mu_old,
                                                          mu_new, eind, rrms, max_err2_bits, row_begin, row_end, mu_commit);
}

void launch_dipole_commit(const SubsetDev& pol, const UpdateOpts& o, const double* mu_old, const double* mu_new, double* mu,
                          int row_begin, int row_end, cudaStream_t s) {
    const int rows = row_end - row_begin;
    if (rows <= 0) return;
    dipole_commit_kernel<<<(rows + 255) / 256, 256, 0, s>>>(pol.n, pol.n_pad, o, mu_old, mu_new, mu, row_begin, row_end);
}

void launch_dipole_update_peer(const SubsetDev& pol, const UpdateOpts& o, const PeerWait& pw, const double* es, double* mu,
                               double* mu_old, double* mu_new, double* eind, double* rrms,
                               unsigned long long* max_err2_bits, int commit, cudaStream_t s) {
    if (pol.n_pad <= 0) return;
    dipole_update_peer_kernel<<<(pol.n_pad + 255) / 256, 256, 0, s>>>(pol.n, pol.n_pad, pol.alpha, o, pw, es, mu, mu_old, mu_new,
                                                                    eind, rrms, max_err2_bits, commit);
}

void launch_palmo_finish_peer(const SubsetDev& pol, const PeerWait& pw, const double* eind, double* eic, cudaStream_t s) {
    if (pol.n_pad <= 0) return;
    palmo_finish_peer_kernel<<<(pol.n_pad + 255) / 256, 256, 0, s>>>(pol.n, pol.n_pad, pw, eind, eic);
}

void launch_peer_sum(const PeerWait& pw, size_t n, double* out, cudaStream_t s) {
    if (n == 0) return;
    peer_sum_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(pw, n, out);
}

void launch_peer_scalars(const PeerWait& pw, int n, int min_slot, double* out, cudaStream_t s) {
    peer_scalars_kernel<<<1, 64, 0, s>>>(pw, n, min_slot, out);
}

void launch_palmo_finish(const SubsetDev& pol, const double* epart, int jsplits, const double* eind, double* eic,
                         int row_begin, int row_end, cudaStream_t s) {
    const int rows = row_end - row_begin;
    if (rows <= 0) return;
    palmo_finish_kernel<<<(rows + 255) / 256, 256, 0, s>>>(pol.n, pol.n_pad, epart, jsplits, eind, eic, row_begin, row_end);
}

void launch_polar_energy(const SubsetDev& pol, const double* mu, const double* es, const double* eic, const double* rrms,
                         int row_begin, int row_end, double* out, cudaStream_t s) {
    polar_energy_kernel<<<1, 1024, 0, s>>>(pol.n, pol.n_pad, mu, es, eic, rrms, row_begin, row_end, out);
}

// ------------------------------------------------------------------------------------------------
// Gauss-Seidel (polar_gs / polar_gs_ranked): blocked forward substitution.
//
// contract_dipoles() with polar_gs overwrites mu[index] inside the i loop
// (thole_iterative.c:52-53), so atom i sees the NEW dipoles of every atom processed before it and
// the OLD dipoles of the rest.  `pol` is already permuted into processing order.  For block b of
// 128 consecutive atoms:
//   panel  (parallel):   E_i = - sum_{j not in b} T_ij mu_j   -- earlier blocks hold new, later old
//   inner  (one CTA):    add the in-block part with old dipoles, then for k = 0..127 in order
//                        finalise mu_k and push T_ik (mu_k_new - mu_k_old) to the rows i > k.
// ------------------------------------------------------------------------------------------------
template <bool ORTHO, int DAMP, bool WF>
__global__ void __launch_bounds__(TILE) gs_block_kernel(Cell c, SubsetDev p, DipoleConsts k, const double* __restrict__ aux,
                                                        const double* __restrict__ alpha, const double* __restrict__ epart,
                                                        int jsplits, const double* __restrict__ es, double* __restrict__ mu,
                                                        double* __restrict__ mu_new, double* __restrict__ eind, int block) {
    __shared__ double sx[TILE], sy[TILE], sz[TILE], smx[TILE], smy[TILE], smz[TILE], saux[TILE];
    __shared__ int smol[TILE];
    __shared__ double sdx, sdy, sdz;
    const int tid = threadIdx.x, np = p.n_pad;
    const int i = block * TILE + tid;
    const bool live = i < p.n;
    const double xi = p.x[i], yi = p.y[i], zi = p.z[i];
    const double auxi = (DAMP == MPMC_DAMP_EXPONENTIAL) ? 0.0 : aux[i];
    const int moli = p.mol[i];
    sx[tid] = xi;
    sy[tid] = yi;
    sz[tid] = zi;
    smx[tid] = mu[i];
    smy[tid] = mu[np + i];
    smz[tid] = mu[2 * np + i];
    saux[tid] = auxi;
    smol[tid] = moli;
    double ex = 0.0, ey = 0.0, ez = 0.0;
    for (int js = 0; js < jsplits; js++) {
        ex += epart[((size_t)js * 3 + 0) * np + i];
        ey += epart[((size_t)js * 3 + 1) * np + i];
        ez += epart[((size_t)js * 3 + 2) * np + i];
    }
    __syncthreads();
    const int jn = min(TILE, p.n - block * TILE);
    if (live)
        for (int jj = 0; jj < jn; jj++) {
            if (jj == tid) continue;
            tensor_apply<ORTHO, DAMP, WF>(c, k, xi - sx[jj], yi - sy[jj], zi - sz[jj], smx[jj], smy[jj], smz[jj], auxi,
                                          saux[jj], moli == smol[jj], ex, ey, ez);
        }
    const double al = live ? alpha[i] : 0.0;
    const double esx = es[i], esy = es[np + i], esz = es[2 * np + i];
    for (int kk = 0; kk < jn; kk++) {
        if (tid == kk) {
            const double nx = al * (esx + ex), ny = al * (esy + ey), nz = al * (esz + ez);
            sdx = nx - smx[kk];
            sdy = ny - smy[kk];
            sdz = nz - smz[kk];
            mu_new[i] = nx;
            mu_new[np + i] = ny;
            mu_new[2 * np + i] = nz;
            mu[i] = nx;
            mu[np + i] = ny;
            mu[2 * np + i] = nz;
            eind[i] = ex;
            eind[np + i] = ey;
            eind[2 * np + i] = ez;
        }
        __syncthreads();
        if (live && tid > kk)
            tensor_apply<ORTHO, DAMP, WF>(c, k, xi - sx[kk], yi - sy[kk], zi - sz[kk], sdx, sdy, sdz, auxi, saux[kk],
                                          moli == smol[kk], ex, ey, ez);
        __syncthreads();
    }
}

void launch_gs_sweep(const Cell& c, bool ortho, const SubsetDev& pol, const DipoleOpts& o, const int* /*order*/,
                     const double* es, double* mu, double* mu_new, double* eind, double* scratch, cudaStream_t s) {
    const DipoleConsts k = make_consts(c, o);
    const int ntiles = pol.n_pad / TILE;
    const int jsplits = ntiles < 148 ? ntiles : 148;
    const double* aux = pol.aux;
    const double* alpha = pol.alpha;
    for (int b = 0; b < ntiles; b++) {
        contract_dispatch(c, ortho, pol, o, mu, b * TILE, (b + 1) * TILE, jsplits, b, scratch, s);
#define LAUNCH(ORTHO, DAMP, WF) \
    gs_block_kernel<ORTHO, DAMP, WF><<<1, TILE, 0, s>>>(c, pol, k, aux, alpha, scratch, jsplits, es, mu, mu_new, eind, b)
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
}

// ------------------------------------------------------------------------------------------------
// Gauss-Seidel in delta form.  With mu_new = mu_old + delta the field atom i sees in a sweep is
//     E_i = - sum_{j != i} T_ij mu_old_j  -  sum_{j < i} T_ij delta_j
// The first sum is one Jacobi contraction with the old dipoles: the symmetric K4 kernel does it for all rows
// at once (T_ij once per unordered pair).  The second is strictly lower triangular and is resolved block by
// block, right-looking: gs_delta_block_kernel finishes the 128 atoms of block b in order (E already holds the
// pushes of all earlier blocks) and leaves delta_b; gs_push_kernel subtracts T_ib delta_b from every later row
// (row-owned, no atomics).  Per sweep: N^2/2 (symmetric) + N^2/2 (pushes) pair evaluations in ~2 N/128 short
// launches, instead of N^2 in the panel form above whose 148-CTA panels leave the GPU mostly idle.
// Same algebra as contract_dipoles() with polar_gs (thole_iterative.c:38-53); only the association of the
// sums differs (1e-16 relative per term).
// ------------------------------------------------------------------------------------------------
template <bool ORTHO, int DAMP, bool WF>
__global__ void __launch_bounds__(TILE) gs_delta_block_kernel(Cell c, SubsetDev p, DipoleConsts k, const double* __restrict__ aux,
                                                              const double* __restrict__ alpha, const double* __restrict__ esum,
                                                              const double* __restrict__ es, double* __restrict__ mu,
                                                              double* __restrict__ mu_new, double* __restrict__ eind,
                                                              double* __restrict__ delta, int block) {
    __shared__ double sx[TILE], sy[TILE], sz[TILE], saux[TILE];
    __shared__ int smol[TILE];
    __shared__ double sdx, sdy, sdz;
    const int tid = threadIdx.x, np = p.n_pad;
    const int i = block * TILE + tid;
    const bool live = i < p.n;
    const double xi = p.x[i], yi = p.y[i], zi = p.z[i];
    const double auxi = (DAMP == MPMC_DAMP_EXPONENTIAL) ? 0.0 : aux[i];
    const int moli = p.mol[i];
    sx[tid] = xi;
    sy[tid] = yi;
    sz[tid] = zi;
    saux[tid] = auxi;
    smol[tid] = moli;
    double ex = esum[i], ey = esum[np + i], ez = esum[2 * np + i];
    const double ox = mu[i], oy = mu[np + i], oz = mu[2 * np + i];
    const double al = live ? alpha[i] : 0.0;
    const double esx = es[i], esy = es[np + i], esz = es[2 * np + i];
    const int jn = min(TILE, p.n - block * TILE);
    if (!live) {
        delta[i] = 0.0;
        delta[np + i] = 0.0;
        delta[2 * np + i] = 0.0;
    }
    __syncthreads();
    for (int kk = 0; kk < jn; kk++) {
        if (tid == kk) {
            const double nx = al * (esx + ex), ny = al * (esy + ey), nz = al * (esz + ez);  // thole_iterative.c:49-53
            sdx = nx - ox;
            sdy = ny - oy;
            sdz = nz - oz;
            mu_new[i] = nx;
            mu_new[np + i] = ny;
            mu_new[2 * np + i] = nz;
            mu[i] = nx;
            mu[np + i] = ny;
            mu[2 * np + i] = nz;
            eind[i] = ex;
            eind[np + i] = ey;
            eind[2 * np + i] = ez;
            delta[i] = sdx;
            delta[np + i] = sdy;
            delta[2 * np + i] = sdz;
        }
        __syncthreads();
        if (live && tid > kk)
            tensor_apply<ORTHO, DAMP, WF>(c, k, xi - sx[kk], yi - sy[kk], zi - sz[kk], sdx, sdy, sdz, auxi, saux[kk],
                                          moli == smol[kk], ex, ey, ez);
        __syncthreads();
    }
}

// (c1, c2) of every in-block pair (i > k) of every 128-atom block, pair (i, k) at i (i-1)/2 + k: they depend on the
// positions only, so one launch serves all sweeps of an evaluation; the sequential block kernel then reads them from
// shared memory instead of evaluating rsqrt / exp inside its 128-step dependency chain.
constexpr int GS_PAIRS = TILE * (TILE - 1) / 2;
template <bool ORTHO, int DAMP, bool WF>
__global__ void __launch_bounds__(TILE) gs_block_coeff_kernel(Cell c, SubsetDev p, DipoleConsts k, const double* __restrict__ aux,
                                                              double2* __restrict__ coef) {
    __shared__ double sx[TILE], sy[TILE], sz[TILE], saux[TILE];
    __shared__ int smol[TILE];
    const int tid = threadIdx.x, b = blockIdx.x;
    const int j = b * TILE + tid;
    sx[tid] = p.x[j];
    sy[tid] = p.y[j];
    sz[tid] = p.z[j];
    saux[tid] = (DAMP == MPMC_DAMP_EXPONENTIAL) ? 0.0 : aux[j];
    smol[tid] = p.mol[j];
    __syncthreads();
    const int jn = min(TILE, p.n - b * TILE);
    double2* out = coef + (size_t)b * GS_PAIRS;
    for (int idx = tid; idx < GS_PAIRS; idx += TILE) {
        int i = (int)((1.0f + sqrtf(1.0f + 8.0f * (float)idx)) * 0.5f);
        while (i * (i - 1) / 2 > idx) i--;
        while ((i + 1) * i / 2 <= idx) i++;
        const int kk = idx - i * (i - 1) / 2;
        double2 v = make_double2(0.0, 0.0);
        if (i < jn) {
            double dx = sx[i] - sx[kk], dy = sy[i] - sy[kk], dz = sz[i] - sz[kk];
            min_image<ORTHO>(c, dx, dy, dz);
            tensor_coeff<DAMP, WF>(k, dx * dx + dy * dy + dz * dz, saux[i], saux[kk], smol[i] == smol[kk], v.x, v.y);
        }
        out[idx] = v;
    }
}

template <bool ORTHO>
__global__ void __launch_bounds__(TILE) gs_delta_block_table_kernel(Cell c, SubsetDev p, const double* __restrict__ alpha,
                                                                    const double2* __restrict__ coef,
                                                                    const double* __restrict__ esum, const double* __restrict__ es,
                                                                    double* __restrict__ mu, double* __restrict__ mu_new,
                                                                    double* __restrict__ eind, double* __restrict__ delta, int block) {
    extern __shared__ double2 sc[];  // [GS_PAIRS]
    __shared__ double sx[TILE], sy[TILE], sz[TILE];
    __shared__ double sdx, sdy, sdz;
    const int tid = threadIdx.x, np = p.n_pad;
    const int i = block * TILE + tid;
    const bool live = i < p.n;
    {
        const double2* src = coef + (size_t)block * GS_PAIRS;
        for (int idx = tid; idx < GS_PAIRS; idx += TILE) sc[idx] = src[idx];
    }
    const double xi = p.x[i], yi = p.y[i], zi = p.z[i];
    sx[tid] = xi;
    sy[tid] = yi;
    sz[tid] = zi;
    double ex = esum[i], ey = esum[np + i], ez = esum[2 * np + i];
    const double ox = mu[i], oy = mu[np + i], oz = mu[2 * np + i];
    const double al = live ? alpha[i] : 0.0;
    const double esx = es[i], esy = es[np + i], esz = es[2 * np + i];
    const int jn = min(TILE, p.n - block * TILE);
    if (!live) {
        delta[i] = 0.0;
        delta[np + i] = 0.0;
        delta[2 * np + i] = 0.0;
    }
    const double2* row = sc + tid * (tid - 1) / 2;
    __syncthreads();
    for (int kk = 0; kk < jn; kk++) {
        if (tid == kk) {
            const double nx = al * (esx + ex), ny = al * (esy + ey), nz = al * (esz + ez);  // thole_iterative.c:49-53
            sdx = nx - ox;
            sdy = ny - oy;
            sdz = nz - oz;
            mu_new[i] = nx;
            mu_new[np + i] = ny;
            mu_new[2 * np + i] = nz;
            mu[i] = nx;
            mu[np + i] = ny;
            mu[2 * np + i] = nz;
            eind[i] = ex;
            eind[np + i] = ey;
            eind[2 * np + i] = ez;
            delta[i] = sdx;
            delta[np + i] = sdy;
            delta[2 * np + i] = sdz;
        }
        __syncthreads();
        if (live && tid > kk) {
            const double2 cf = row[kk];
            double dx = xi - sx[kk], dy = yi - sy[kk], dz = zi - sz[kk];
            min_image<ORTHO>(c, dx, dy, dz);
            const double mx = sdx, my = sdy, mz = sdz;
            const double dm = (dx * mx + dy * my + dz * mz) * cf.y;
            ex += dm * dx - cf.x * mx;
            ey += dm * dy - cf.x * my;
            ez += dm * dz - cf.x * mz;
        }
        __syncthreads();
    }
}

// Warp-synchronous form of the sequential block kernel.  The 128 atoms of a block are four sub-blocks of 32, one per
// warp.  -T_ik of every pair inside a sub-block is tabulated (full symmetric 3x3, 93 KB of shared memory) before the
// chain starts, so one step of the chain is: candidate delta of lane k (3 dependent fp64 ops) -> three 64-bit shuffles
// -> nine FMAs in chains of three; no block barrier, no rsqrt/exp.  Between sub-blocks (three block barriers) the
// later warps apply the 32 finished deltas to their rows with the ordinary pair evaluation (independent pairs, pipelined).
constexpr int GS_SUB = 32, GS_SUBPAIRS = GS_SUB * (GS_SUB - 1) / 2;  // 496
template <bool ORTHO, int DAMP, bool WF>
__global__ void __launch_bounds__(TILE) gs_delta_block_warp_kernel(Cell c, SubsetDev p, DipoleConsts k, const double* __restrict__ aux,
                                                                   const double* __restrict__ alpha, const double* __restrict__ esum,
                                                                   const double* __restrict__ es, double* __restrict__ mu,
                                                                   double* __restrict__ mu_new, double* __restrict__ eind,
                                                                   double* __restrict__ delta, int block) {
    extern __shared__ double tens[];  // [4][GS_SUBPAIRS][6]: -T = c2 d d^T - c1 I  (xx xy xz yy yz zz)
    __shared__ double sx[TILE], sy[TILE], sz[TILE], saux[TILE], sdl[3][TILE];
    __shared__ int smol[TILE];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5, np = p.n_pad;
    const int i = block * TILE + tid;
    const int jn = min(TILE, p.n - block * TILE);
    const bool live = tid < jn;
    const double xi = p.x[i], yi = p.y[i], zi = p.z[i];
    const double auxi = (DAMP == MPMC_DAMP_EXPONENTIAL) ? 0.0 : aux[i];
    const int moli = p.mol[i];
    sx[tid] = xi;
    sy[tid] = yi;
    sz[tid] = zi;
    saux[tid] = auxi;
    smol[tid] = moli;
    double ex = esum[i], ey = esum[np + i], ez = esum[2 * np + i];
    const double ox = mu[i], oy = mu[np + i], oz = mu[2 * np + i];
    const double al = live ? alpha[i] : 0.0;
    const double esx = es[i], esy = es[np + i], esz = es[2 * np + i];
    __syncthreads();
    for (int idx = tid; idx < 4 * GS_SUBPAIRS; idx += TILE) {
        const int s = idx / GS_SUBPAIRS, q = idx - s * GS_SUBPAIRS;
        int a = (int)((1.0f + sqrtf(1.0f + 8.0f * (float)q)) * 0.5f);
        while (a * (a - 1) / 2 > q) a--;
        while ((a + 1) * a / 2 <= q) a++;
        const int b = q - a * (a - 1) / 2;  // pair (a, b), b < a, inside sub-block s
        const int ia = s * GS_SUB + a, ib = s * GS_SUB + b;
        double m[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
        if (ia < jn) {  // ib < ia: both atoms exist
            double dx = sx[ia] - sx[ib], dy = sy[ia] - sy[ib], dz = sz[ia] - sz[ib];
            min_image<ORTHO>(c, dx, dy, dz);
            double c1, c2;
            tensor_coeff<DAMP, WF>(k, dx * dx + dy * dy + dz * dz, saux[ia], saux[ib], smol[ia] == smol[ib], c1, c2);
            m[0] = c2 * dx * dx - c1;
            m[1] = c2 * dx * dy;
            m[2] = c2 * dx * dz;
            m[3] = c2 * dy * dy - c1;
            m[4] = c2 * dy * dz;
            m[5] = c2 * dz * dz - c1;
        }
        double* t = tens + (size_t)idx * 6;
#pragma unroll
        for (int e = 0; e < 6; e++) t[e] = m[e];
    }
    __syncthreads();
    for (int s = 0; s < 4; s++) {
        if (w == s) {
            const double* trow = tens + ((size_t)s * GS_SUBPAIRS + (size_t)lane * (lane - 1) / 2) * 6;
            for (int kk = 0; kk < GS_SUB; kk++) {
                // every lane forms its candidate; lane kk's is final (its field is complete): thole_iterative.c:49-53
                const double cx = al * (esx + ex) - ox, cy = al * (esy + ey) - oy, cz = al * (esz + ez) - oz;
                const double bx = __shfl_sync(0xffffffffu, cx, kk), by = __shfl_sync(0xffffffffu, cy, kk),
                             bz = __shfl_sync(0xffffffffu, cz, kk);
                if (lane > kk) {
                    const double* t = trow + kk * 6;
                    ex += t[0] * bx + t[1] * by + t[2] * bz;
                    ey += t[1] * bx + t[3] * by + t[4] * bz;
                    ez += t[2] * bx + t[4] * by + t[5] * bz;
                }
            }
            const double nx = al * (esx + ex), ny = al * (esy + ey), nz = al * (esz + ez);
            sdl[0][tid] = nx - ox;
            sdl[1][tid] = ny - oy;
            sdl[2][tid] = nz - oz;
            if (live) {
                mu_new[i] = nx;
                mu_new[np + i] = ny;
                mu_new[2 * np + i] = nz;
                mu[i] = nx;
                mu[np + i] = ny;
                mu[2 * np + i] = nz;
                eind[i] = ex;
                eind[np + i] = ey;
                eind[2 * np + i] = ez;
            }
            delta[i] = live ? nx - ox : 0.0;
            delta[np + i] = live ? ny - oy : 0.0;
            delta[2 * np + i] = live ? nz - oz : 0.0;
        }
        if (s == 3) break;
        __syncthreads();
        if (w > s && live) {
            const int j1 = min((s + 1) * GS_SUB, jn);
#pragma unroll 2
            for (int j = s * GS_SUB; j < j1; j++)
                tensor_apply<ORTHO, DAMP, WF>(c, k, xi - sx[j], yi - sy[j], zi - sz[j], sdl[0][j], sdl[1][j], sdl[2][j], auxi,
                                              saux[j], moli == smol[j], ex, ey, ez);
        }
    }
}

// esum_i -= sum_{k in block} T_ik delta_k for the rows i >= row_begin.  CTA = 64 rows x 4 column groups of 32:
// the four partial sums of a row are added in group order (fixed association).
constexpr int PUSH_ROWS = 64;
template <bool ORTHO, int DAMP, bool WF>
__global__ void __launch_bounds__(4 * PUSH_ROWS) gs_push_kernel(Cell c, SubsetDev p, DipoleConsts k, const double* __restrict__ aux,
                                                               const double* __restrict__ delta, int block, int row_begin,
                                                               double* __restrict__ esum) {
    __shared__ double sx[TILE], sy[TILE], sz[TILE], sdx[TILE], sdy[TILE], sdz[TILE], saux[TILE];
    __shared__ int smol[TILE];
    __shared__ double red[3][3][PUSH_ROWS];
    const int tid = threadIdx.x, np = p.n_pad;
    if (tid < TILE) {
        const int j = block * TILE + tid;
        sx[tid] = p.x[j];
        sy[tid] = p.y[j];
        sz[tid] = p.z[j];
        sdx[tid] = delta[j];
        sdy[tid] = delta[np + j];
        sdz[tid] = delta[2 * np + j];
        saux[tid] = (DAMP == MPMC_DAMP_EXPONENTIAL) ? 0.0 : aux[j];
        smol[tid] = p.mol[j];
    }
    __syncthreads();
    const int r = tid % PUSH_ROWS, g = tid / PUSH_ROWS;
    const int i = row_begin + blockIdx.x * PUSH_ROWS + r;
    const bool live = i < p.n;
    double ex = 0.0, ey = 0.0, ez = 0.0;
    if (live) {
        const double xi = p.x[i], yi = p.y[i], zi = p.z[i];
        const double auxi = (DAMP == MPMC_DAMP_EXPONENTIAL) ? 0.0 : aux[i];
        const int moli = p.mol[i];
        const int jn = min(TILE, p.n - block * TILE);
        const int j0 = g * 32, j1 = min(j0 + 32, jn);
#pragma unroll 2
        for (int jj = j0; jj < j1; jj++)
            tensor_apply<ORTHO, DAMP, WF>(c, k, xi - sx[jj], yi - sy[jj], zi - sz[jj], sdx[jj], sdy[jj], sdz[jj], auxi, saux[jj],
                                          moli == smol[jj], ex, ey, ez);
    }
    if (g > 0) {
        red[g - 1][0][r] = ex;
        red[g - 1][1][r] = ey;
        red[g - 1][2][r] = ez;
    }
    __syncthreads();
    if (g == 0 && live) {
        ex = ((ex + red[0][0][r]) + red[1][0][r]) + red[2][0][r];
        ey = ((ey + red[0][1][r]) + red[1][1][r]) + red[2][1][r];
        ez = ((ez + red[0][2][r]) + red[1][2][r]) + red[2][2][r];
        esum[i] += ex;  // tensor_apply accumulates -T m
        esum[np + i] += ey;
        esum[2 * np + i] += ez;
    }
}

// ------------------------------------------------------------------------------------------------
// Fused Jacobi-family solve for small systems: thole_iterative() (thole_iterative.c:168-259) in ONE cooperative launch.
// Below a few thousand polarizable atoms an iteration is launch- and latency-bound (h2_fluid_1000: 600 polarizable atoms,
// 53 us per iteration for a contraction + an update launch whose CTAs each walk 128 columns in sequence).  Here the
// grid keeps running across iterations: per iteration every CTA takes (32 rows x 128 columns) items -- 512 threads =
// 32 rows x 16 lanes, 8 columns per lane, the lanes' sums added in lane order in shared memory -> one partial row per
// 128-column range -- grid barrier -- every thread updates its rows from the partial rows (the arithmetic of
// dipole_update_kernel / dipole_commit_kernel: new_mu, SOR / ESOR mix, rrms, the are_we_done_yet residual) -- grid
// barrier.  Convergence (polar_precision), the 128-iteration bail-out (:199-210) and the Palmo pass (:119-141) are
// decided inside the kernel; the host reads the iteration count with the other scalars.  Fixed-order sums, no atomics
// on any field (the residual's atomicMax on non-negative doubles is order independent).
// ------------------------------------------------------------------------------------------------
constexpr int FS_THREADS = 512, FS_ROWS = 32, FS_LANES = 16, FS_COLS = 128;

struct FusedArgs {
    int n, np, njr;  // polarizable atoms, padded, 128-column ranges
    const double *x, *y, *z, *aux, *alpha, *es;
    const int* mol;
    double *mu, *mu_old, *mu_new, *eind, *eic, *rrms;
    double* part;             // [njr][3][np]
    double* mu_palmo;         // [3][np]: the dipoles the Palmo pass saw (its rows for the non-polarizable atoms come later)
    unsigned long long* err;  // [2] residual words, alternating by iteration parity
    double* out;              // [2]: iterations done, iter_success (1 = bailed out)
    int max_iter, precision_mode, palmo;
    double allowed;
    UpdateOpts uo;
};

template <bool ORTHO, int DAMP, bool WF>
__global__ void __launch_bounds__(FS_THREADS, 1) dipole_fused_solver_kernel(Cell c, DipoleConsts k, FusedArgs a) {
    namespace cg = cooperative_groups;
    cg::grid_group grid = cg::this_grid();
    __shared__ double sx[FS_COLS], sy[FS_COLS], sz[FS_COLS], smx[FS_COLS], smy[FS_COLS], smz[FS_COLS], saux[FS_COLS];
    __shared__ int smol[FS_COLS];
    __shared__ double red[FS_LANES][3][FS_ROWS];
    const int tid = threadIdx.x, r = tid & (FS_ROWS - 1), l = tid >> 5;
    const int np = a.np, n = a.n;
    const int nrg = (n + FS_ROWS - 1) / FS_ROWS, items = nrg * a.njr;
    const int gtid = blockIdx.x * FS_THREADS + tid, gthreads = gridDim.x * FS_THREADS;

    // part[jr][.][i] = - sum_{j in range jr, j != i} T_ij mu_j for every row i
    auto contract = [&](const double* __restrict__ mu) {
        for (int item = blockIdx.x; item < items; item += gridDim.x) {
            const int rg = item % nrg, jr = item / nrg;
            const int i = rg * FS_ROWS + r;
            __syncthreads();
            if (tid < FS_COLS) {
                const int j = jr * FS_COLS + tid;  // < np: np is a multiple of 128
                sx[tid] = a.x[j];
                sy[tid] = a.y[j];
                sz[tid] = a.z[j];
                smx[tid] = __ldcg(mu + j);
                smy[tid] = __ldcg(mu + np + j);
                smz[tid] = __ldcg(mu + 2 * np + j);
                saux[tid] = (DAMP == MPMC_DAMP_EXPONENTIAL) ? 0.0 : a.aux[j];
                smol[tid] = a.mol[j];
            }
            __syncthreads();
            double ex = 0.0, ey = 0.0, ez = 0.0;
            if (i < n) {
                const double xi = a.x[i], yi = a.y[i], zi = a.z[i];
                const double auxi = (DAMP == MPMC_DAMP_EXPONENTIAL) ? 0.0 : a.aux[i];
                const int moli = a.mol[i];
#pragma unroll 4
                for (int q = 0; q < FS_COLS / FS_LANES; q++) {
                    const int jj = l + FS_LANES * q, j = jr * FS_COLS + jj;
                    if (j < n && j != i)
                        tensor_apply<ORTHO, DAMP, WF>(c, k, xi - sx[jj], yi - sy[jj], zi - sz[jj], smx[jj], smy[jj], smz[jj], auxi,
                                                      saux[jj], moli == smol[jj], ex, ey, ez);
                }
            }
            red[l][0][r] = ex;
            red[l][1][r] = ey;
            red[l][2][r] = ez;
            __syncthreads();
            if (tid < 3 * FS_ROWS) {
                const int cdim = tid >> 5, rr = tid & 31, ii = rg * FS_ROWS + rr;
                double s = red[0][cdim][rr];
#pragma unroll
                for (int q = 1; q < FS_LANES; q++) s += red[q][cdim][rr];
                if (ii < np) a.part[((size_t)jr * 3 + cdim) * np + ii] = s;
            }
        }
    };

    int it = 0, fail = 0;
    bool keep = true;
    while (keep) {
        it++;
        if (it >= kMAX_ITERATION_COUNT && a.precision_mode) {  // thole_iterative.c:199-210: mu = alpha E_s, eic = 0
            for (int i = gtid; i < n; i += gthreads)
                for (int cdim = 0; cdim < 3; cdim++) {
                    a.mu[cdim * np + i] = a.alpha[i] * a.es[cdim * np + i];
                    a.eic[cdim * np + i] = 0.0;
                }
            fail = 1;
            break;
        }
        contract(a.mu);
        grid.sync();
        if (gtid == 0) a.err[(it + 1) & 1] = 0ull;  // next iteration's word: last read before this iteration's first barrier
        // ---- update (dipole_update_kernel): E_ind, new_mu, residual, rrms; commit unless a Palmo pass needs the old dipoles
        {
            UpdateOpts o = a.uo;
            o.iteration = it;
            const bool commit = !a.palmo;
            const double wold = (commit && o.esor) ? exp(-o.gamma * o.iteration) : 0.0;
            double worst = 0.0;
            for (int i = gtid; i < n; i += gthreads) {
                const double al = a.alpha[i];
                double num = 0.0, den = 0.0;
                for (int cdim = 0; cdim < 3; cdim++) {
                    double e = 0.0;
                    for (int js = 0; js < a.njr; js++) e += __ldcg(a.part + ((size_t)js * 3 + cdim) * np + i);
                    a.eind[cdim * np + i] = e;
                    const double old = a.mu[cdim * np + i];
                    if (o.save_old) a.mu_old[cdim * np + i] = old;
                    const double nw = al * (a.es[cdim * np + i] + e);  // thole_iterative.c:49
                    a.mu_new[cdim * np + i] = nw;
                    const double d = nw - old;
                    worst = fmax(worst, d * d);
                    num += d * d;
                    den += nw * nw;
                    if (commit) {  // thole_iterative.c:242-252
                        double v = nw;
                        if (o.sor)
                            v = o.gamma * nw + (1.0 - o.gamma) * old;
                        else if (o.esor)
                            v = (1.0 - wold) * nw + wold * old;
                        a.mu[cdim * np + i] = v;
                    }
                }
                if (o.do_rrms) {
                    double v = sqrt(num / den);
                    if (!isfinite(v)) v = 0.0;
                    a.rrms[i] = v;
                }
            }
            if (a.precision_mode) {
                worst = warp_max(worst);
                if ((tid & 31) == 0 && worst > 0.0) atomicMax(a.err + (it & 1), (unsigned long long)__double_as_longlong(worst));
            }
        }
        grid.sync();
        if (a.precision_mode) {
            const unsigned long long bits = *reinterpret_cast<volatile unsigned long long*>(a.err + (it & 1));
            keep = __longlong_as_double((long long)bits) > a.allowed;
        } else
            keep = it != a.max_iter;
        if (a.palmo) {
            if (!keep) {  // palmo_contraction (:119-141) with the dipoles as they are before this iteration's commit
                if (a.mu_palmo)
                    for (int i = gtid; i < np; i += gthreads)
                        for (int cdim = 0; cdim < 3; cdim++) a.mu_palmo[cdim * np + i] = a.mu[cdim * np + i];
                contract(a.mu);
                grid.sync();
                for (int i = gtid; i < n; i += gthreads)
                    for (int cdim = 0; cdim < 3; cdim++) {
                        double e = -a.eind[cdim * np + i];  // thole_iterative.c:129-130
                        for (int js = 0; js < a.njr; js++) e += __ldcg(a.part + ((size_t)js * 3 + cdim) * np + i);
                        a.eic[cdim * np + i] = e;
                    }
            }
            // commit (dipole_commit_kernel): mu = new_mu | SOR | ESOR mix
            for (int i = gtid; i < n; i += gthreads)
                for (int cdim = 0; cdim < 3; cdim++) {
                    const double nw = a.mu_new[cdim * np + i];
                    double v = nw;
                    if (a.uo.sor)
                        v = a.uo.gamma * nw + (1.0 - a.uo.gamma) * a.mu_old[cdim * np + i];
                    else if (a.uo.esor) {
                        const double w = exp(-a.uo.gamma * it);
                        v = (1.0 - w) * nw + w * a.mu_old[cdim * np + i];
                    }
                    a.mu[cdim * np + i] = v;
                }
            grid.sync();
        }
    }
    if (gtid == 0) {
        a.out[0] = (double)it;
        a.out[1] = (double)fail;
    }
}

// false -- nothing launched -- when the fused solve does not apply here (no cooperative launch, too many items per CTA)
bool launch_dipole_fused_solver(const Cell& c, bool ortho, const SubsetDev& pol, const DipoleOpts& o, const UpdateOpts& uo,
                                int max_iter, double allowed, bool precision_mode, bool palmo, const double* es, double* mu,
                                double* mu_old, double* mu_new, double* eind, double* eic, double* rrms, double* part,
                                double* mu_palmo, unsigned long long* err2, double* out2, cudaStream_t s) {
    if (pol.n <= 0) return false;
    int dev = 0, sms = 0, coop = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
    if (!coop) return false;
    const DipoleConsts k = make_consts(c, o);
    FusedArgs a;
    a.n = pol.n;
    a.np = pol.n_pad;
    a.njr = (pol.n + FS_COLS - 1) / FS_COLS;
    a.x = pol.x, a.y = pol.y, a.z = pol.z, a.aux = pol.aux, a.alpha = pol.alpha, a.es = es, a.mol = pol.mol;
    a.mu = mu, a.mu_old = mu_old, a.mu_new = mu_new, a.eind = eind, a.eic = eic, a.rrms = rrms;
    a.part = part;
    a.mu_palmo = mu_palmo;
    a.err = err2;
    a.out = out2;
    a.max_iter = max_iter;
    a.precision_mode = precision_mode ? 1 : 0;
    a.palmo = palmo ? 1 : 0;
    a.allowed = allowed;
    a.uo = uo;
    const int items = ((pol.n + FS_ROWS - 1) / FS_ROWS) * a.njr;
    Cell cc = c;
    DipoleConsts kk = k;
    void* args[] = {&cc, &kk, &a};
    cudaError_t err = cudaErrorInvalidValue;
#define LAUNCH(ORTHO, DAMP, WF)                                                                      \
    {                                                                                                \
        const void* fn = (const void*)dipole_fused_solver_kernel<ORTHO, DAMP, WF>;                   \
        int per_sm = 0;                                                                              \
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, FS_THREADS, 0);                   \
        if (per_sm < 1) return false;                                                                \
        int gridx = per_sm * sms;                                                                    \
        if (gridx > items) gridx = items;                                                            \
        err = cudaLaunchCooperativeKernel(fn, dim3(gridx), dim3(FS_THREADS), args, 0, s);            \
    }
#define DISPATCH(ORTHO)                                              \
    if (o.damp_type == MPMC_DAMP_EXPONENTIAL) {                      \
        if (o.wolf_full) LAUNCH(ORTHO, MPMC_DAMP_EXPONENTIAL, true)  \
        else LAUNCH(ORTHO, MPMC_DAMP_EXPONENTIAL, false)             \
    } else if (o.damp_type == MPMC_DAMP_LINEAR) {                    \
        LAUNCH(ORTHO, MPMC_DAMP_LINEAR, false)                       \
    } else {                                                         \
        if (o.wolf_full) LAUNCH(ORTHO, MPMC_DAMP_OFF, true)          \
        else LAUNCH(ORTHO, MPMC_DAMP_OFF, false)                     \
    }
    if (ortho) {
        DISPATCH(true)
    } else {
        DISPATCH(false)
    }
#undef DISPATCH
#undef LAUNCH
    if (err != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return true;
}

// ------------------------------------------------------------------------------------------------
// Persistent form of the delta sweep: ONE launch per sweep instead of two per 128-atom block (726 at 63k atoms).
// CTA 0 (the resolver) walks the blocks in order and never leaves its SM; every other CTA (the pushers, one per SM)
// owns a fixed set of row blocks and applies each finished delta_b to them as soon as the resolver publishes it:
//
//   resolver, block b:  E_b = (Jacobi part) + pushacc_b (pushers: deltas of blocks <= b-2) + its own push of b-1
//                       -> the 4 x 32-step warp-synchronous chain of gs_delta_block_warp_kernel, with the in-sub-block
//                          tensors read from a table made once per ordering (gs_block_tensor_kernel)
//                       -> publish delta_b (flag ready[b]) -> push delta_b into the rows of block b+1 itself: that one
//                          push is the only one on the critical path, so it never crosses a flag
//   pusher p, row blocks r = 2+p, 2+p+P, ...:  for b = 0, 1, ...: wait ready[b]; acc_r -= T_rb delta_b for its r >= b+2
//                       (most urgent first); when b = r-2 the sum is complete: publish it (pushacc, flag pushed[r])
//
// Look-ahead: while the resolver works on block b+1 the pushers spread delta_b over rows > b+2; only (b -> b+2) has a
// deadline, one block period away.  Every sum has a fixed order (b ascending, column quarters in order): bit-reproducible.
// Flags carry the sweep's epoch, so they are never cleared; every wait has a 2 s time-out that ends the kernel with
// an error flag instead of hanging the GPU.  Launched cooperatively (co-residency of all CTAs is what the waits need).
// ------------------------------------------------------------------------------------------------
constexpr int GSP_THREADS = 512;
constexpr int GSP_MAXOWN = 8;                 // row blocks a pusher may own (shared-memory accumulators)
constexpr int GS_TENS = 4 * GS_SUBPAIRS * 6;  // doubles of one block's in-sub-block tensor table
constexpr int GSP_RES_DOUBLES = 2 * GS_TENS + 4 * TILE + 3 * TILE + 9 * TILE + 12 * TILE + TILE / 2;  // table double buffered
constexpr int GSP_PUSH_DOUBLES = 7 * TILE + TILE / 2 + GSP_MAXOWN * 12 * TILE;
constexpr int GSP_SMEM_DOUBLES = GSP_RES_DOUBLES > GSP_PUSH_DOUBLES ? GSP_RES_DOUBLES : GSP_PUSH_DOUBLES;

struct GsSweepArgs {
    int n, np, nb;  // polarizable atoms, padded, live 128-atom blocks
    const double *x, *y, *z, *aux, *alpha;
    const int* mol;
    const double *esum, *es, *tens;
    const double2* ctab;  // [nb][2][128 cols][128 rows] (c1, c2) of the pairs (block b, block b) and (block b+1, block b)
    double *mu, *mu_new, *eind, *delta, *pushacc;
    unsigned *ready, *pushed;
    unsigned epoch;
    double* timeout_flag;
    unsigned long long* trace;  // optional [nb][8] globaltimer stamps of the resolver (MPMC_B200_GS_TRACE): wait, load, chain, post, push
};

#define GSP_STAMP(slot)                                                                \
    if (a.trace && tid == 0) {                                                         \
        unsigned long long t__;                                                        \
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t__));                        \
        a.trace[(size_t)b * 8 + (slot)] = t__;                                         \
    }

// M = -alpha_a T_ab = alpha_a (c2 d d^T - c1 I)  (xx xy xz yy yz zz) of every pair (a > b) inside the four 32-atom
// sub-blocks of every block: what a finished delta_b adds to the CANDIDATE delta of a later atom a of its sub-block
template <bool ORTHO, int DAMP, bool WF>
__global__ void __launch_bounds__(TILE) gs_block_tensor_kernel(Cell c, SubsetDev p, DipoleConsts k, const double* __restrict__ aux,
                                                               double* __restrict__ tens) {
    __shared__ double sx[TILE], sy[TILE], sz[TILE], saux[TILE], sal[TILE];
    __shared__ int smol[TILE];
    const int tid = threadIdx.x, b = blockIdx.x;
    const int j = b * TILE + tid;
    sx[tid] = p.x[j];
    sy[tid] = p.y[j];
    sz[tid] = p.z[j];
    saux[tid] = (DAMP == MPMC_DAMP_EXPONENTIAL) ? 0.0 : aux[j];
    smol[tid] = p.mol[j];
    sal[tid] = j < p.n ? p.alpha[j] : 0.0;
    __syncthreads();
    const int jn = min(TILE, p.n - b * TILE);
    double* out = tens + (size_t)b * GS_TENS;
    for (int idx = tid; idx < 4 * GS_SUBPAIRS; idx += TILE) {
        const int s = idx / GS_SUBPAIRS, q = idx - s * GS_SUBPAIRS;
        int a = (int)((1.0f + sqrtf(1.0f + 8.0f * (float)q)) * 0.5f);
        while (a * (a - 1) / 2 > q) a--;
        while ((a + 1) * a / 2 <= q) a++;
        const int bb = q - a * (a - 1) / 2;  // pair (a, bb), bb < a, inside sub-block s
        const int ia = s * GS_SUB + a, ib = s * GS_SUB + bb;
        double m[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
        if (ia < jn) {
            double dx = sx[ia] - sx[ib], dy = sy[ia] - sy[ib], dz = sz[ia] - sz[ib];
            min_image<ORTHO>(c, dx, dy, dz);
            double c1, c2;
            tensor_coeff<DAMP, WF>(k, dx * dx + dy * dy + dz * dz, saux[ia], saux[ib], smol[ia] == smol[ib], c1, c2);
            const double al = sal[ia];
            m[0] = al * (c2 * dx * dx - c1);
            m[1] = al * (c2 * dx * dy);
            m[2] = al * (c2 * dx * dz);
            m[3] = al * (c2 * dy * dy - c1);
            m[4] = al * (c2 * dy * dz);
            m[5] = al * (c2 * dz * dz - c1);
        }
#pragma unroll
        for (int e = 0; e < 6; e++) out[(size_t)idx * 6 + e] = m[e];
    }
}

// (c1, c2) of every pair the RESOLVER applies: rows of block b against columns of block b (the pushes between the
// sub-blocks of a block) and rows of block b+1 against columns of block b (the one push on the critical path), laid
// out [column][row] so that a warp of consecutive rows reads 512 contiguous bytes.  Made once per ordering; the
// resolver then spends ~22 fp64 instructions per pair (minimum image + apply) instead of ~50 (rsqrt, exp, damping).
constexpr int GS_CTAB = 2 * TILE * TILE;  // double2 entries per block
template <bool ORTHO, int DAMP, bool WF>
__global__ void __launch_bounds__(TILE) gs_pair_coeff_kernel(Cell c, SubsetDev p, DipoleConsts k, const double* __restrict__ aux,
                                                             double2* __restrict__ ctab) {
    __shared__ double sx[TILE], sy[TILE], sz[TILE], saux[TILE];
    __shared__ int smol[TILE];
    const int tid = threadIdx.x, b = blockIdx.x, t = blockIdx.y;
    const int j = b * TILE + tid;  // column atom staged by this thread
    sx[tid] = p.x[j];
    sy[tid] = p.y[j];
    sz[tid] = p.z[j];
    saux[tid] = (DAMP == MPMC_DAMP_EXPONENTIAL) ? 0.0 : aux[j];
    smol[tid] = p.mol[j];
    __syncthreads();
    const int i = (b + t) * TILE + tid;  // row atom of this thread
    const bool row_live = i < p.n;       // (i < n_pad always: n_pad is padded past the last live block by construction)
    const double xi = row_live ? p.x[i] : 0.0, yi = row_live ? p.y[i] : 0.0, zi = row_live ? p.z[i] : 0.0;
    const double auxi = (row_live && DAMP != MPMC_DAMP_EXPONENTIAL) ? aux[i] : 0.0;
    const int moli = row_live ? p.mol[i] : -1;
    const int jn = min(TILE, p.n - b * TILE);
    double2* out = ctab + ((size_t)b * 2 + t) * TILE * TILE;
    for (int jj = 0; jj < TILE; jj++) {
        double2 v = make_double2(0.0, 0.0);
        if (row_live && jj < jn && (t == 1 || tid > jj)) {
            double dx = xi - sx[jj], dy = yi - sy[jj], dz = zi - sz[jj];
            min_image<ORTHO>(c, dx, dy, dz);
            tensor_coeff<DAMP, WF>(k, dx * dx + dy * dy + dz * dz, auxi, saux[jj], moli == smol[jj], v.x, v.y);
        }
        out[(size_t)jj * TILE + tid] = v;
    }
}

template <bool ORTHO>
__device__ __forceinline__ void tensor_apply_coef(const Cell& c, double dx, double dy, double dz, double mx, double my,
                                                  double mz, const double2 cf, double& ex, double& ey, double& ez) {
    min_image<ORTHO>(c, dx, dy, dz);
    const double dm = (dx * mx + dy * my + dz * mz) * cf.y;
    ex += dm * dx - cf.x * mx;
    ey += dm * dy - cf.x * my;
    ez += dm * dz - cf.x * mz;
}

// thread 0 of a CTA: wait until *flag == epoch (acquire); false after 2 s
__device__ __forceinline__ bool gsp_wait(const unsigned* flag, unsigned epoch, double* timeout_flag) {
    unsigned long long t0, t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
        unsigned v;
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
        if (v == epoch) return true;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        if (t1 - t0 > 2000000000ull) {
            *timeout_flag = 1.0;
            return false;
        }
    }
}
// thread 0 of a CTA, after a barrier that follows the CTA's global writes: publish them
__device__ __forceinline__ void gsp_post(unsigned* flag, unsigned epoch) {
    __threadfence();
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(flag), "r"(epoch) : "memory");
}

template <bool ORTHO, int DAMP, bool WF>
__global__ void __launch_bounds__(GSP_THREADS, 1) gs_sweep_persistent_kernel(Cell c, DipoleConsts k, GsSweepArgs a) {
    extern __shared__ __align__(16) double sm[];
    __shared__ int s_abort;
    const int tid = threadIdx.x, row = tid & (TILE - 1), q = tid >> 7, lane = tid & 31;
    const int np = a.np;
    if (tid == 0) s_abort = 0;
    __syncthreads();
    if (blockIdx.x == 0) {
        // ------------------------------------------------------------------ resolver
        double* tens0 = sm;                     // [2][GS_TENS]: this block's table and the next one's (cp.async prefetch)
        double* sx = tens0 + 2 * GS_TENS;       // [128] each
        double* sy = sx + TILE;
        double* sz = sy + TILE;
        double* saux = sz + TILE;
        double* sdl = saux + TILE;              // [3][128] finished deltas of this block
        double* part = sdl + 3 * TILE;          // [3][3][128] partial cross pushes of quarters 1..3
        double* nxt = part + 9 * TILE;          // [4][3][128] this CTA's push of block b into the rows of block b+1
        int* smol = reinterpret_cast<int*>(nxt + 12 * TILE);
        auto prefetch = [&](int blk) {          // table of block `blk` -> buffer blk & 1, 16 bytes per copy, asynchronous
            const char* src = reinterpret_cast<const char*>(a.tens + (size_t)blk * GS_TENS);
            const unsigned dst = (unsigned)__cvta_generic_to_shared(tens0 + (size_t)(blk & 1) * GS_TENS);
            for (int ch = tid; ch < GS_TENS / 2; ch += GSP_THREADS)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 16u * ch), "l"(src + 16 * (size_t)ch) : "memory");
            asm volatile("cp.async.commit_group;" ::: "memory");
        };
        prefetch(0);
        for (int b = 0; b < a.nb; b++) {
            const int i = b * TILE + row;
            const int jn = min(TILE, a.n - b * TILE);
            const bool live = row < jn;
            const double* tens = tens0 + (size_t)(b & 1) * GS_TENS;
            GSP_STAMP(0)
            if (b >= 2 && tid == 0)
                if (!gsp_wait(a.pushed + b, a.epoch, a.timeout_flag)) s_abort = 1;
            asm volatile("cp.async.wait_group 0;" ::: "memory");  // this thread's share of the table of block b has landed
            __syncthreads();  // ... and everybody's; nxt of the previous block is complete, its other buffers are free
            if (s_abort) return;
            GSP_STAMP(1)
            if (b + 1 < a.nb) prefetch(b + 1);  // lands while this block's chains run
            const double xi = a.x[i], yi = a.y[i], zi = a.z[i];
            const double auxi = (DAMP == MPMC_DAMP_EXPONENTIAL) ? 0.0 : a.aux[i];
            const int moli = a.mol[i];
            double ex = 0.0, ey = 0.0, ez = 0.0, ox = 0.0, oy = 0.0, oz = 0.0, al = 0.0, esx = 0.0, esy = 0.0, esz = 0.0;
            if (q == 0) {
                sx[row] = xi;
                sy[row] = yi;
                sz[row] = zi;
                saux[row] = auxi;
                smol[row] = moli;
                ex = a.esum[i];
                ey = a.esum[np + i];
                ez = a.esum[2 * np + i];
                if (b >= 2) {
                    ex += __ldcg(a.pushacc + i);
                    ey += __ldcg(a.pushacc + np + i);
                    ez += __ldcg(a.pushacc + 2 * np + i);
                }
                if (b >= 1) {
                    ex += ((nxt[0 * 3 * TILE + row] + nxt[1 * 3 * TILE + row]) + nxt[2 * 3 * TILE + row]) + nxt[3 * 3 * TILE + row];
                    ey += ((nxt[0 * 3 * TILE + TILE + row] + nxt[1 * 3 * TILE + TILE + row]) + nxt[2 * 3 * TILE + TILE + row]) +
                          nxt[3 * 3 * TILE + TILE + row];
                    ez += ((nxt[0 * 3 * TILE + 2 * TILE + row] + nxt[1 * 3 * TILE + 2 * TILE + row]) +
                           nxt[2 * 3 * TILE + 2 * TILE + row]) + nxt[3 * 3 * TILE + 2 * TILE + row];
                }
                ox = a.mu[i];
                oy = a.mu[np + i];
                oz = a.mu[2 * np + i];
                al = live ? a.alpha[i] : 0.0;
                esx = a.es[i];
                esy = a.es[np + i];
                esz = a.es[2 * np + i];
            }
            __syncthreads();
            GSP_STAMP(2)
            for (int s = 0; s < 4; s++) {
                if (q == 0 && (row >> 5) == s) {
                    // Warp s: the 32-step chain (thole_iterative.c:49-53 for 32 atoms in order), carried on the CANDIDATE
                    // delta c = alpha (E_s + E) - mu_old of every lane: lane k's is final at step k, and a finished delta_k
                    // changes a later candidate by M_ak delta_k with M = -alpha_a T_ak tabulated -- per step one
                    // broadcast and a three-deep FMA chain, nothing else on the dependency path.
                    const double* trow = tens + ((size_t)s * GS_SUBPAIRS + (size_t)lane * (lane - 1) / 2) * 6;
                    double cx = fma(al, esx + ex, -ox), cy = fma(al, esy + ey, -oy), cz = fma(al, esz + ez, -oz);
                    // No branch inside a step: every lane reads "its" entry kk (for lanes <= kk that is somebody else's
                    // entry or the first doubles behind the table -- valid shared memory, finite values) and the
                    // result is kept only where lane > kk.  The loads depend on kk alone, so they run ahead of the chain.
#pragma unroll 8
                    for (int kk = 0; kk < GS_SUB; kk++) {
                        const double* t = trow + kk * 6;
                        const double t0 = t[0], t1 = t[1], t2 = t[2], t3 = t[3], t4 = t[4], t5 = t[5];
                        const double bx = __shfl_sync(0xffffffffu, cx, kk), by = __shfl_sync(0xffffffffu, cy, kk),
                                     bz = __shfl_sync(0xffffffffu, cz, kk);
                        const double ncx = fma(t2, bz, fma(t1, by, fma(t0, bx, cx)));
                        const double ncy = fma(t4, bz, fma(t3, by, fma(t1, bx, cy)));
                        const double ncz = fma(t5, bz, fma(t4, by, fma(t2, bx, cz)));
                        const bool upd = lane > kk;
                        cx = upd ? ncx : cx;
                        cy = upd ? ncy : cy;
                        cz = upd ? ncz : cz;
                    }
                    const double nx = ox + cx, ny = oy + cy, nz = oz + cz;  // mu_new = mu_old + delta
                    sdl[row] = live ? cx : 0.0;
                    sdl[TILE + row] = live ? cy : 0.0;
                    sdl[2 * TILE + row] = live ? cz : 0.0;
                    if (live) {
                        a.mu_new[i] = nx;
                        a.mu_new[np + i] = ny;
                        a.mu_new[2 * np + i] = nz;
                        a.mu[i] = nx;
                        a.mu[np + i] = ny;
                        a.mu[2 * np + i] = nz;
                        // the field this atom saw when it was resolved: mu_new = alpha (E_s + E)
                        a.eind[i] = nx / al - esx;
                        a.eind[np + i] = ny / al - esy;
                        a.eind[2 * np + i] = nz / al - esz;
                    }
                    a.delta[i] = live ? cx : 0.0;
                    a.delta[np + i] = live ? cy : 0.0;
                    a.delta[2 * np + i] = live ? cz : 0.0;
                }
                if (s == 3) break;
                __syncthreads();
                // the 32 finished deltas of sub-block s -> rows of the later sub-blocks, 8 columns per quarter
                double px = 0.0, py = 0.0, pz = 0.0;
                const bool later = (row >> 5) > s && live;
                if (later) {
                    const int j0 = s * GS_SUB + q * 8, j1 = min(j0 + 8, jn);
                    const double2* cf = a.ctab + (size_t)b * GS_CTAB + row;
#pragma unroll 4
                    for (int j = j0; j < j1; j++)
                        tensor_apply_coef<ORTHO>(c, xi - sx[j], yi - sy[j], zi - sz[j], sdl[j], sdl[TILE + j], sdl[2 * TILE + j],
                                                 __ldg(cf + (size_t)j * TILE), px, py, pz);
                    if (q > 0) {
                        part[(q - 1) * 3 * TILE + row] = px;
                        part[(q - 1) * 3 * TILE + TILE + row] = py;
                        part[(q - 1) * 3 * TILE + 2 * TILE + row] = pz;
                    }
                }
                __syncthreads();
                if (q == 0 && later) {
                    ex += ((px + part[row]) + part[3 * TILE + row]) + part[6 * TILE + row];
                    ey += ((py + part[TILE + row]) + part[3 * TILE + TILE + row]) + part[6 * TILE + TILE + row];
                    ez += ((pz + part[2 * TILE + row]) + part[3 * TILE + 2 * TILE + row]) + part[6 * TILE + 2 * TILE + row];
                }
            }
            __syncthreads();  // every delta of the block is in sdl and on its way to global memory
            GSP_STAMP(3)
            // publish from the last thread: its fence (~5 us of write-ack latency) then stalls one warp of a
            // throughput-bound phase instead of the thread every stamp and wait of this CTA goes through
            if (tid == GSP_THREADS - 1) gsp_post(a.ready + b, a.epoch);
            GSP_STAMP(4)
            if (b + 1 < a.nb) {  // the one push on the critical path: block b -> rows of block b+1, 32 columns per quarter
                const int i2 = (b + 1) * TILE + row;
                double px = 0.0, py = 0.0, pz = 0.0;
                if (i2 < a.n) {
                    const double x2 = a.x[i2], y2 = a.y[i2], z2 = a.z[i2];
                    const int j0 = q * 32, j1 = min(j0 + 32, jn);
                    const double2* cf = a.ctab + (size_t)b * GS_CTAB + (size_t)TILE * TILE + row;
#pragma unroll 4
                    for (int j = j0; j < j1; j++)
                        tensor_apply_coef<ORTHO>(c, x2 - sx[j], y2 - sy[j], z2 - sz[j], sdl[j], sdl[TILE + j], sdl[2 * TILE + j],
                                                 __ldg(cf + (size_t)j * TILE), px, py, pz);
                }
                nxt[q * 3 * TILE + row] = px;
                nxt[q * 3 * TILE + TILE + row] = py;
                nxt[q * 3 * TILE + 2 * TILE + row] = pz;
            }
            GSP_STAMP(5)
        }
        return;
    }
    // ---------------------------------------------------------------------- pusher
    const int p = (int)blockIdx.x - 1, P = (int)gridDim.x - 1;
    int nown = 0;
    for (int r = 2 + p; r < a.nb; r += P) nown++;
    if (nown == 0) return;
    double* bx = sm;  // block b: positions, delta, aux  [128] each
    double* by = bx + TILE;
    double* bz = by + TILE;
    double* bdx = bz + TILE;
    double* bdy = bdx + TILE;
    double* bdz = bdy + TILE;
    double* baux = bdz + TILE;
    int* bmol = reinterpret_cast<int*>(baux + TILE);
    double* acc = baux + TILE + TILE / 2;  // [nown][4][3][128]
    for (int idx = tid; idx < nown * 12 * TILE; idx += GSP_THREADS) acc[idx] = 0.0;
    const int rmax = 2 + p + (nown - 1) * P;
    for (int b = 0; b + 2 <= rmax; b++) {
        if (tid == 0)
            if (!gsp_wait(a.ready + b, a.epoch, a.timeout_flag)) s_abort = 1;
        __syncthreads();  // also: everybody is done with the previous block's staging buffers
        if (s_abort) return;
        const int jn = min(TILE, a.n - b * TILE);
        if (tid < TILE) {
            const int j = b * TILE + tid;
            bx[tid] = a.x[j];
            by[tid] = a.y[j];
            bz[tid] = a.z[j];
            bdx[tid] = __ldcg(a.delta + j);
            bdy[tid] = __ldcg(a.delta + np + j);
            bdz[tid] = __ldcg(a.delta + 2 * np + j);
            baux[tid] = (DAMP == MPMC_DAMP_EXPONENTIAL) ? 0.0 : a.aux[j];
            bmol[tid] = a.mol[j];
        }
        __syncthreads();
        const int k0 = b > p ? (b - p + P - 1) / P : 0;  // first owned row block r >= b + 2
        for (int kk = k0; kk < nown; kk++) {
            const int r = 2 + p + kk * P;
            const int i = r * TILE + row;
            double* mine = acc + ((size_t)kk * 4 + q) * 3 * TILE;
            if (i < a.n) {
                const double xi = a.x[i], yi = a.y[i], zi = a.z[i];
                const double auxi = (DAMP == MPMC_DAMP_EXPONENTIAL) ? 0.0 : a.aux[i];
                const int moli = a.mol[i];
                double ex = 0.0, ey = 0.0, ez = 0.0;
                const int j0 = q * 32, j1 = min(j0 + 32, jn);
#pragma unroll 2
                for (int j = j0; j < j1; j++)
                    tensor_apply<ORTHO, DAMP, WF>(c, k, xi - bx[j], yi - by[j], zi - bz[j], bdx[j], bdy[j], bdz[j], auxi, baux[j],
                                                  moli == bmol[j], ex, ey, ez);
                mine[row] += ex;
                mine[TILE + row] += ey;
                mine[2 * TILE + row] += ez;
            }
            if (r == b + 2) {  // the deltas of blocks 0..r-2 are all in: hand the sum to the resolver
                __syncthreads();
                if (q == 0) {
                    const double* a0 = acc + (size_t)kk * 12 * TILE;
#pragma unroll
                    for (int d = 0; d < 3; d++)
                        a.pushacc[(size_t)d * np + i] = ((a0[d * TILE + row] + a0[3 * TILE + d * TILE + row]) +
                                                         a0[6 * TILE + d * TILE + row]) + a0[9 * TILE + d * TILE + row];
                }
                __syncthreads();
                if (tid == 0) gsp_post(a.pushed + r, a.epoch);
            }
        }
    }
}

size_t gs_block_tensor_doubles(int n_pad) { return (size_t)(n_pad / TILE) * GS_TENS; }
size_t gs_pair_coeff_doubles(int n_pad) { return (size_t)(n_pad / TILE) * GS_CTAB * 2; }

void launch_gs_block_tensors(const Cell& c, bool ortho, const SubsetDev& pol, const DipoleOpts& o, double* tens, double* ctab,
                             cudaStream_t s) {
    const DipoleConsts k = make_consts(c, o);
    const int nb = (pol.n + TILE - 1) / TILE;
    if (nb <= 0) return;
    const double* aux = pol.aux;
#define LAUNCH(ORTHO, DAMP, WF)                                                        \
    {                                                                                  \
        gs_block_tensor_kernel<ORTHO, DAMP, WF><<<nb, TILE, 0, s>>>(c, pol, k, aux, tens); \
        gs_pair_coeff_kernel<ORTHO, DAMP, WF><<<dim3(nb, 2), TILE, 0, s>>>(c, pol, k, aux, reinterpret_cast<double2*>(ctab)); \
    }
#define DISPATCH(ORTHO)                                              \
    if (o.damp_type == MPMC_DAMP_EXPONENTIAL) {                      \
        if (o.wolf_full) LAUNCH(ORTHO, MPMC_DAMP_EXPONENTIAL, true)  \
        else LAUNCH(ORTHO, MPMC_DAMP_EXPONENTIAL, false)             \
    } else if (o.damp_type == MPMC_DAMP_LINEAR) {                    \
        LAUNCH(ORTHO, MPMC_DAMP_LINEAR, false)                       \
    } else {                                                         \
        if (o.wolf_full) LAUNCH(ORTHO, MPMC_DAMP_OFF, true)          \
        else LAUNCH(ORTHO, MPMC_DAMP_OFF, false)                     \
    }
    if (ortho) {
        DISPATCH(true)
    } else {
        DISPATCH(false)
    }
#undef DISPATCH
#undef LAUNCH
}

// One sweep in one launch.  esum must hold the Jacobi part (launch_dipole_contract_sym with the dipoles as they are at
// the start of the sweep); flags [2 * n_pad / 128] unsigned, zeroed once when allocated; epoch > 0, increasing from call
// to call.  Returns false -- nothing launched -- when the kernel cannot run here (more row blocks than the pushers'
// accumulators hold, no cooperative launch, not one CTA per SM): the caller then takes the per-block launches.
bool launch_gs_sweep_persistent(const Cell& c, bool ortho, const SubsetDev& pol, const DipoleOpts& o, const double* esum,
                                const double* es, double* mu, double* mu_new, double* eind, double* delta, const double* tens,
                                const double* ctab, double* pushacc, unsigned* flags, unsigned epoch, double* timeout_flag,
                                cudaStream_t s, unsigned long long* trace) {
    const int nb = (pol.n + TILE - 1) / TILE;
    if (nb <= 0) return true;
    int dev = 0, sms = 0, coop = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
    if (!coop || sms < 2) return false;
    const int P = sms - 1;
    if (nb - 2 > GSP_MAXOWN * P) return false;
    const size_t smem = sizeof(double) * GSP_SMEM_DOUBLES;
    const DipoleConsts k = make_consts(c, o);
    GsSweepArgs a;
    a.n = pol.n;
    a.np = pol.n_pad;
    a.nb = nb;
    a.x = pol.x;
    a.y = pol.y;
    a.z = pol.z;
    a.aux = pol.aux;
    a.alpha = pol.alpha;
    a.mol = pol.mol;
    a.esum = esum;
    a.es = es;
    a.tens = tens;
    a.ctab = reinterpret_cast<const double2*>(ctab);
    a.mu = mu;
    a.mu_new = mu_new;
    a.eind = eind;
    a.delta = delta;
    a.pushacc = pushacc;
    a.ready = flags;
    a.pushed = flags + pol.n_pad / TILE;
    a.epoch = epoch;
    a.timeout_flag = timeout_flag;
    a.trace = trace;
    Cell cc = c;
    DipoleConsts kk = k;
    void* args[] = {&cc, &kk, &a};
    cudaError_t err = cudaErrorInvalidValue;
#define LAUNCH(ORTHO, DAMP, WF)                                                                                          \
    {                                                                                                                    \
        static unsigned long long seen = 0;                                                                              \
        const void* fn = (const void*)gs_sweep_persistent_kernel<ORTHO, DAMP, WF>;                                       \
        if (first_on_device(seen)) cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);      \
        int per_sm = 0;                                                                                                  \
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, GSP_THREADS, smem);                                   \
        if (per_sm < 1) return false;                                                                                    \
        err = cudaLaunchCooperativeKernel(fn, dim3(sms), dim3(GSP_THREADS), args, smem, s);                              \
    }
#define DISPATCH(ORTHO)                                              \
    if (o.damp_type == MPMC_DAMP_EXPONENTIAL) {                      \
        if (o.wolf_full) LAUNCH(ORTHO, MPMC_DAMP_EXPONENTIAL, true)  \
        else LAUNCH(ORTHO, MPMC_DAMP_EXPONENTIAL, false)             \
    } else if (o.damp_type == MPMC_DAMP_LINEAR) {                    \
        LAUNCH(ORTHO, MPMC_DAMP_LINEAR, false)                       \
    } else {                                                         \
        if (o.wolf_full) LAUNCH(ORTHO, MPMC_DAMP_OFF, true)          \
        else LAUNCH(ORTHO, MPMC_DAMP_OFF, false)                     \
    }
    if (ortho) {
        DISPATCH(true)
    } else {
        DISPATCH(false)
    }
#undef DISPATCH
#undef LAUNCH
    if (err != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return true;
}

size_t gs_block_coeff_doubles(int n_pad) { return (size_t)(n_pad / TILE) * GS_PAIRS * 2; }

void launch_gs_block_coeffs(const Cell& c, bool ortho, const SubsetDev& pol, const DipoleOpts& o, double* coef, cudaStream_t s) {
    const DipoleConsts k = make_consts(c, o);
    const int nb = (pol.n + TILE - 1) / TILE;
    if (nb <= 0) return;
    const double* aux = pol.aux;
    double2* out = reinterpret_cast<double2*>(coef);
#define LAUNCH(ORTHO, DAMP, WF) gs_block_coeff_kernel<ORTHO, DAMP, WF><<<nb, TILE, 0, s>>>(c, pol, k, aux, out);
#define DISPATCH(ORTHO)                                              \
    if (o.damp_type == MPMC_DAMP_EXPONENTIAL) {                      \
        if (o.wolf_full) LAUNCH(ORTHO, MPMC_DAMP_EXPONENTIAL, true)  \
        else LAUNCH(ORTHO, MPMC_DAMP_EXPONENTIAL, false)             \
    } else if (o.damp_type == MPMC_DAMP_LINEAR) {                    \
        LAUNCH(ORTHO, MPMC_DAMP_LINEAR, false)                       \
    } else {                                                         \
        if (o.wolf_full) LAUNCH(ORTHO, MPMC_DAMP_OFF, true)          \
        else LAUNCH(ORTHO, MPMC_DAMP_OFF, false)                     \
    }
    if (ortho) {
        DISPATCH(true)
    } else {
        DISPATCH(false)
    }
#undef DISPATCH
#undef LAUNCH
}

void launch_gs_sweep_delta(const Cell& c, bool ortho, const SubsetDev& pol, const DipoleOpts& o, const double* cbox, int ysplits,
                           double* epartI, double* epartJ, double* esum, const double* es, double* mu, double* mu_new,
                           double* eind, double* delta, const double* coef, cudaStream_t s) {
    static unsigned long long attr_set = 0;
    if (first_on_device(attr_set)) {
        cudaFuncSetAttribute(gs_delta_block_table_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(GS_PAIRS * sizeof(double2)));
        cudaFuncSetAttribute(gs_delta_block_table_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(GS_PAIRS * sizeof(double2)));
    }
    const double2* cf = reinterpret_cast<const double2*>(coef);
    static int warp_form = -1;  // MPMC_B200_GS_BLOCK=table|plain selects the earlier block kernels (A/B)
    if (warp_form < 0) {
        const char* v = getenv("MPMC_B200_GS_BLOCK");
        warp_form = (v && (!strcmp(v, "table") || !strcmp(v, "plain"))) ? 0 : 1;
        if (v && !strcmp(v, "plain")) warp_form = 2;
    }
    if (warp_form == 2) cf = nullptr;
    const size_t wsmem = sizeof(double) * 4 * GS_SUBPAIRS * 6;
    const DipoleConsts k = make_consts(c, o);
    const int ntiles = pol.n_pad / TILE;
    const double* aux = pol.aux;
    const double* alpha = pol.alpha;
    // Jacobi part with the dipoles as they are at the start of the sweep
    launch_dipole_contract_sym(c, ortho, pol, o, mu, cbox, 0, 1, ysplits, epartI, epartJ, esum, s);
    const int nlive_tiles = (pol.n + TILE - 1) / TILE;
    for (int b = 0; b < nlive_tiles; b++) {
        const int row_begin = (b + 1) * TILE;
        const int rows = pol.n - row_begin;
        const int pgrid = rows > 0 ? (rows + PUSH_ROWS - 1) / PUSH_ROWS : 0;
#define LAUNCH(ORTHO, DAMP, WF)                                                                                                \
    {                                                                                                                          \
        if (warp_form == 1) {                                                                                                  \
            static unsigned long long wattr = 0;                                                                                         \
            if (first_on_device(wattr)) {                                                                                                      \
                cudaFuncSetAttribute(gs_delta_block_warp_kernel<ORTHO, DAMP, WF>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                     (int)wsmem);                                                                              \
            }                                                                                                                  \
            gs_delta_block_warp_kernel<ORTHO, DAMP, WF><<<1, TILE, wsmem, s>>>(c, pol, k, aux, alpha, esum, es, mu, mu_new,    \
                                                                             eind, delta, b);                                 \
        } else if (cf)                                                                                                         \
            gs_delta_block_table_kernel<ORTHO><<<1, TILE, GS_PAIRS * sizeof(double2), s>>>(c, pol, alpha, cf, esum, es, mu,     \
                                                                                         mu_new, eind, delta, b);             \
        else                                                                                                                   \
            gs_delta_block_kernel<ORTHO, DAMP, WF><<<1, TILE, 0, s>>>(c, pol, k, aux, alpha, esum, es, mu, mu_new, eind, delta, b); \
        if (pgrid > 0)                                                                                                         \
            gs_push_kernel<ORTHO, DAMP, WF><<<pgrid, 4 * PUSH_ROWS, 0, s>>>(c, pol, k, aux, delta, b, row_begin, esum);       \
    }
#define DISPATCH(ORTHO)                                              \
    if (o.damp_type == MPMC_DAMP_EXPONENTIAL) {                      \
        if (o.wolf_full) LAUNCH(ORTHO, MPMC_DAMP_EXPONENTIAL, true)  \
        else LAUNCH(ORTHO, MPMC_DAMP_EXPONENTIAL, false)             \
    } else if (o.damp_type == MPMC_DAMP_LINEAR) {                    \
        LAUNCH(ORTHO, MPMC_DAMP_LINEAR, false)                       \
    } else {                                                         \
        if (o.wolf_full) LAUNCH(ORTHO, MPMC_DAMP_OFF, true)          \
        else LAUNCH(ORTHO, MPMC_DAMP_OFF, false)                     \
    }
        if (ortho) {
            DISPATCH(true)
        } else {
            DISPATCH(false)
        }
#undef DISPATCH
#undef LAUNCH
    }
    (void)ntiles;
}

// ------------------------------------------------------------------------------------------------
// rank metric of pairs() (src/energy/pairs.c:337-360): rmin = smallest imaged separation between
// polarizable atoms; rank_i = number of polarizable partners whose UN-imaged separation is
// <= 1.5 rmin (the reference mixes the two on purpose or not; reproduced).
// ------------------------------------------------------------------------------------------------
// Rank metric of pairs() (src/energy/pairs.c:337-360): rmin = the smallest minimum-image separation of two polarizable
// atoms; rank_metric_i = the number of polarizable atoms within 1.5 rmin of atom i by the UN-imaged distance.  Both are
// all-pairs tests, so the only things that matter are filling the GPU (source tiles split over gridDim.y) and keeping
// the square root out of the loop.  That is exact: sqrt is monotone and correctly rounded, so min sqrt(r2) = sqrt(min r2),
// and sqrt(r2) <= lim holds exactly when r2 <= t2 with t2 the largest double whose square root is <= lim
// (rank_finish_kernel finds it by stepping through the neighbouring doubles).  Counts are whole numbers far below 2^53:
// adding them with atomicAdd is exact in any order.
constexpr int RANK_MAX_SPLITS = 8;
template <bool ORTHO, int PASS>
__global__ void __launch_bounds__(TILE) rank_kernel(Cell c, SubsetDev p, const double* __restrict__ lim_in, int tiles_per_split,
                                                    double* __restrict__ out) {
    __shared__ double sx[TILE], sy[TILE], sz[TILE];
    __shared__ double red[32];
    const int tid = threadIdx.x, i = blockIdx.x * TILE + tid;
    const double xi = p.x[i], yi = p.y[i], zi = p.z[i];
    const double t2 = PASS == 1 ? lim_in[2] : 0.0;
    double best2 = kMAXVALUE;
    int count = 0;
    const int t0 = blockIdx.y * tiles_per_split, t1 = min(t0 + tiles_per_split, p.n_pad / TILE);
    for (int jt = t0; jt < t1; jt++) {
        __syncthreads();
        sx[tid] = p.x[jt * TILE + tid];
        sy[tid] = p.y[jt * TILE + tid];
        sz[tid] = p.z[jt * TILE + tid];
        __syncthreads();
        const int jn = min(TILE, p.n - jt * TILE);
        const bool own = jt == (int)blockIdx.x;
#pragma unroll 4
        for (int jj = 0; jj < jn; jj++) {
            double dx = xi - sx[jj], dy = yi - sy[jj], dz = zi - sz[jj];
            if (PASS == 0) min_image<ORTHO>(c, dx, dy, dz);
            const double r2 = dx * dx + dy * dy + dz * dz;
            const bool self = own && jj == tid;
            if (PASS == 0)
                best2 = self ? best2 : fmin(best2, r2);
            else
                count += (!self && r2 <= t2) ? 1 : 0;
        }
    }
    if (PASS == 0) {
        if (i >= p.n) best2 = kMAXVALUE;
        best2 = block_reduce<1>(best2, red);
        if (tid == 0) out[blockIdx.y * gridDim.x + blockIdx.x] = best2;
    } else if (i < p.n && count > 0)
        atomicAdd(out + i, (double)count);
}

// out[0] = rmin, out[1] = lim = 1.5 rmin (pairs.c:352), out[2] = the largest r2 with sqrt(r2) <= lim
__global__ void rank_finish_kernel(const double* __restrict__ in, int n, double* __restrict__ out) {
    __shared__ double red[32];
    double best = kMAXVALUE;
    for (int i = threadIdx.x; i < n; i += blockDim.x) best = fmin(best, in[i]);
    best = block_reduce<1>(best, red);
    if (threadIdx.x == 0) {
        const double rmin = sqrt(best);
        const double lim = rmin * 1.5;
        double x = lim * lim;
        for (int it = 0; it < 8 && sqrt(x) > lim; it++) x = __longlong_as_double(__double_as_longlong(x) - 1);
        for (int it = 0; it < 8; it++) {
            const double up = __longlong_as_double(__double_as_longlong(x) + 1);
            if (!(sqrt(up) <= lim)) break;
            x = up;
        }
        out[0] = rmin;
        out[1] = lim;
        out[2] = x;
    }
}

void launch_rank_metric(const Cell& c, bool ortho, const SubsetDev& pol, double* rmin_part, double* rank, cudaStream_t s) {
    const int nt = pol.n_pad / TILE;
    if (nt <= 0) return;
    int sms = 148, dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int js = (8 * sms + nt - 1) / nt;
    js = js < 1 ? 1 : (js > RANK_MAX_SPLITS ? RANK_MAX_SPLITS : js);
    if (js > nt) js = nt;
    const int tps = (nt + js - 1) / js;
    js = (nt + tps - 1) / tps;
    const dim3 grid(nt, js);
    // rmin_part: [js * nt] block minima of r^2, then rmin / lim / t2 (needs RANK_MAX_SPLITS * nt + 8 doubles)
    double* fin = rmin_part + (size_t)js * nt;
    cudaMemsetAsync(rank, 0, sizeof(double) * pol.n_pad, s);
    if (ortho) {
        rank_kernel<true, 0><<<grid, TILE, 0, s>>>(c, pol, nullptr, tps, rmin_part);
        rank_finish_kernel<<<1, 1024, 0, s>>>(rmin_part, js * nt, fin);
        rank_kernel<true, 1><<<grid, TILE, 0, s>>>(c, pol, fin, tps, rank);
    } else {
        rank_kernel<false, 0><<<grid, TILE, 0, s>>>(c, pol, nullptr, tps, rmin_part);
        rank_finish_kernel<<<1, 1024, 0, s>>>(rmin_part, js * nt, fin);
        rank_kernel<false, 1><<<grid, TILE, 0, s>>>(c, pol, fin, tps, rank);
    }
}

// ------------------------------------------------------------------------------------------------
// fp64 FMA peak probe: 8 independent dependent-chains per thread
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) dfma_probe_kernel(double* out, int iters, double a, double b) {
    double v0 = threadIdx.x * 1e-9, v1 = v0 + 1, v2 = v0 + 2, v3 = v0 + 3, v4 = v0 + 4, v5 = v0 + 5, v6 = v0 + 6, v7 = v0 + 7;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 8; u++) {
            v0 = fma(v0, a, b);
            v1 = fma(v1, a, b);
            v2 = fma(v2, a, b);
            v3 = fma(v3, a, b);
            v4 = fma(v4, a, b);
            v5 = fma(v5, a, b);
            v6 = fma(v6, a, b);
            v7 = fma(v7, a, b);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = v0 + v1 + v2 + v3 + v4 + v5 + v6 + v7;
}

double measure_fp64_peak_tflops(cudaStream_t s) {
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int blocks = sms * 8, threads = 256, iters = 4096;
    double* out = nullptr;
    if (cudaMalloc(&out, sizeof(double) * blocks * threads) != cudaSuccess) return -1.0;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    double best = 0.0;
    for (int rep = 0; rep < 5; rep++) {
        cudaEventRecord(e0, s);
        dfma_probe_kernel<<<blocks, threads, 0, s>>>(out, iters, 0.999999, 1e-7);
        cudaEventRecord(e1, s);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        const double flops = 2.0 * 64.0 * iters * (double)blocks * threads;
        const double tf = flops / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(out);
    return best;
}

}  // namespace mpmc
