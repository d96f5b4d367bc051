// recip_kernel.cu -- K2: reciprocal-space structure factor of the Ewald sum, and the
// reciprocal part of the Ewald static field.
//
// Reference loops replaced (relative to the reference tree):
//   coulombic_reciprocal()  src/energy/coulombic.c:42-95   (non-frozen, charged atoms)
//   recip_term()            src/polarization/polar_ewald.c:78-132  (all atoms)
//
// k.r = 2 pi sum_q l_q s_q with the fractional coordinate s_q = sum_p reciprocal_basis[p][q] r_p, so
// e^{i k.r} factorises into three per-axis phases e^{2 pi i l s}.  Each CTA owns a contiguous chunk
// of atoms; per sub-chunk of RC_ATOMS atoms the (kmax+1) phases per axis are tabulated in shared
// memory with sincospi (no recurrence, no accumulated error), then every thread accumulates its
// k-vectors over the sub-chunk.  Per-chunk partial structure factors are added in chunk order by
// recip_finish (deterministic).
#include "kernels.h"

namespace mpmc {

constexpr int RC_ATOMS = 64;     // atoms per shared-memory sub-chunk
constexpr int RC_THREADS = 256;  // threads per CTA
constexpr int RC_MAXL = 16;      // kmax + 1 <= RC_MAXL

__global__ void __launch_bounds__(RC_THREADS) recip_sf_kernel(Cell c, SubsetDev at, KVecDev kv, int chunk_begin,
                                                              int atoms_per_chunk, double* __restrict__ sf_partial) {
    extern __shared__ double smem[];
    // tables: [axis][l][atom] cos and sin, then charge
    const int nl = kv.kmax + 1;
    double* tcos = smem;
    double* tsin = tcos + 3 * nl * RC_ATOMS;
    double* tq = tsin + 3 * nl * RC_ATOMS;

    const int chunk = chunk_begin + blockIdx.x;
    const int a0 = chunk * atoms_per_chunk;
    const int a1 = min(a0 + atoms_per_chunk, at.n);

    // each thread owns k-vectors tid, tid + RC_THREADS, ... (at most 8 => nk <= 2048)
    double sre[8], sim[8];
#pragma unroll
    for (int m = 0; m < 8; m++) sre[m] = sim[m] = 0.0;

    for (int base = a0; base < a1; base += RC_ATOMS) {
        const int cnt = min(RC_ATOMS, a1 - base);
        __syncthreads();
        for (int w = threadIdx.x; w < 3 * nl * RC_ATOMS; w += RC_THREADS) {
            const int aa = w % RC_ATOMS, l = (w / RC_ATOMS) % nl, ax = w / (RC_ATOMS * nl);
            double cs = 1.0, sn = 0.0;
            if (aa < cnt) {
                const int g = base + aa;
                const double x = at.x[g], y = at.y[g], z = at.z[g];
                const double sfrac = c.R[0][ax] * x + c.R[1][ax] * y + c.R[2][ax] * z;
                sincospi(2.0 * (double)l * sfrac, &sn, &cs);
            }
            tcos[w] = cs;
            tsin[w] = sn;
        }
        for (int w = threadIdx.x; w < RC_ATOMS; w += RC_THREADS) tq[w] = (w < cnt) ? at.q[base + w] : 0.0;
        __syncthreads();
#pragma unroll
        for (int m = 0; m < 8; m++) {
            const int ik = threadIdx.x + m * RC_THREADS;
            if (ik >= kv.nk) break;
            const int l0 = kv.l[ik], l1 = kv.l[kv.nk + ik], l2 = kv.l[2 * kv.nk + ik];
            const int o0 = (0 * nl + l0) * RC_ATOMS, o1 = (1 * nl + abs(l1)) * RC_ATOMS, o2 = (2 * nl + abs(l2)) * RC_ATOMS;
            const double sg1 = l1 < 0 ? -1.0 : 1.0, sg2 = l2 < 0 ? -1.0 : 1.0;
            double re = 0.0, im = 0.0;
            for (int aa = 0; aa < RC_ATOMS; aa++) {
                const double c0 = tcos[o0 + aa], s0 = tsin[o0 + aa];
                const double c1 = tcos[o1 + aa], s1 = sg1 * tsin[o1 + aa];
                const double c2 = tcos[o2 + aa], s2 = sg2 * tsin[o2 + aa];
                const double c01 = c0 * c1 - s0 * s1, s01 = s0 * c1 + c0 * s1;
                const double cc = c01 * c2 - s01 * s2, ss = s01 * c2 + c01 * s2;
                const double q = tq[aa];
                re += q * cc;
                im += q * ss;
            }
            sre[m] += re;
            sim[m] += im;
        }
    }
    double* out = sf_partial + (size_t)blockIdx.x * 2 * kv.nk;
#pragma unroll
    for (int m = 0; m < 8; m++) {
        const int ik = threadIdx.x + m * RC_THREADS;
        if (ik >= kv.nk) break;
        out[ik] = sre[m];
        out[kv.nk + ik] = sim[m];
    }
}

// S(k) = sum over chunks.  One CTA per 32 k-vectors: thread (kx, cy) adds the chunks cy, cy + 32, ... of its k-vector
// (independent loads: the serial one-thread-per-k form waited nchunks x one memory latency, 136 us at 300 chunks),
// then the 32 partials of a k-vector are added in cy order.  The CTA behind the last k block, if `mu` is given, sums the
// three components of the dipoles (full-Ewald polarization: the surface term needs sum(mu) in the same pass).
__global__ void __launch_bounds__(1024) chunk_sum_kernel(int nk, const double* __restrict__ sf_partial, int nchunks,
                                                         double* __restrict__ sf, int kblocks, int mu_n, int mu_np,
                                                         const double* __restrict__ mu, double* __restrict__ mu_total) {
    __shared__ double red[2][32][33];
    const int kx = threadIdx.x & 31, cy = threadIdx.x >> 5;
    if ((int)blockIdx.x >= kblocks) {
        double a0 = 0.0, a1 = 0.0, a2 = 0.0;
        for (int i = threadIdx.x; i < mu_n; i += 1024) {
            a0 += mu[i];
            a1 += mu[mu_np + i];
            a2 += mu[2 * (size_t)mu_np + i];
        }
        double* scratch = &red[0][0][0];
        a0 = block_reduce<0>(a0, scratch);
        if (threadIdx.x == 0) mu_total[0] = a0;
        a1 = block_reduce<0>(a1, scratch);
        if (threadIdx.x == 0) mu_total[1] = a1;
        a2 = block_reduce<0>(a2, scratch);
        if (threadIdx.x == 0) mu_total[2] = a2;
        return;
    }
    const int ik = blockIdx.x * 32 + kx;
    double re = 0.0, im = 0.0;
    if (ik < nk)
        for (int ch = cy; ch < nchunks; ch += 32) {
            re += sf_partial[(size_t)ch * 2 * nk + ik];
            im += sf_partial[(size_t)ch * 2 * nk + nk + ik];
        }
    red[0][cy][kx] = re;
    red[1][cy][kx] = im;
    __syncthreads();
    if (cy < 2 && ik < nk) {  // warp 0: real parts, warp 1: imaginary parts
        double v = 0.0;
        for (int q = 0; q < 32; q++) v += red[cy][q][kx];
        sf[cy * nk + ik] = v;
    }
}

// E = scale * sum_k w_k |S(k)|^2 over this shard's k-vectors: k in list order per thread, fixed-order block reduction
__global__ void __launch_bounds__(1024) recip_energy_kernel(KVecDev kv, const double* __restrict__ sf, double scale,
                                                            double* __restrict__ energy_out, int k_begin, int k_end) {
    __shared__ double red[32];
    double acc = 0.0;
    for (int ik = k_begin + threadIdx.x; ik < k_end; ik += blockDim.x) {
        const double re = sf[ik], im = sf[kv.nk + ik];
        acc += kv.weight[ik] * (re * re + im * im);
    }
    double v = block_reduce<0>(acc, red);
    if (threadIdx.x == 0) *energy_out = v * scale;
}

// Ewald static field, reciprocal part: for every atom i (frozen and uncharged included)
//   E_i += scale * sum_k w_k k [ sin(k.r_i) Re S(k) - cos(k.r_i) Im S(k) ]
// (polar_ewald.c:117-125, 8 pi / V applied at :127-131 => `scale`).
__global__ void __launch_bounds__(TILE) recip_field_kernel(Cell c, AtomsDev a, KVecDev kv, const double* __restrict__ kvec,
                                                           const double* __restrict__ sf, double scale, int row_begin,
                                                           int row_end, double* __restrict__ ef) {
    extern __shared__ double smem[];
    const int nl = kv.kmax + 1;
    double* tcos = smem;                   // [axis][l][thread]
    double* tsin = tcos + 3 * nl * TILE;
    const int i = row_begin + blockIdx.x * TILE + threadIdx.x;
    const bool live = i < row_end && (a.flags[i] & F_VALID);
    const double x = live ? a.x[i] : 0.0, y = live ? a.y[i] : 0.0, z = live ? a.z[i] : 0.0;
    for (int ax = 0; ax < 3; ax++) {
        const double sfrac = c.R[0][ax] * x + c.R[1][ax] * y + c.R[2][ax] * z;
        for (int l = 0; l < nl; l++) {
            double sn, cs;
            sincospi(2.0 * (double)l * sfrac, &sn, &cs);
            tcos[(ax * nl + l) * TILE + threadIdx.x] = cs;
            tsin[(ax * nl + l) * TILE + threadIdx.x] = sn;
        }
    }
    double ex = 0.0, ey = 0.0, ez = 0.0;
    for (int ik = 0; ik < kv.nk; ik++) {
        const int l0 = kv.l[ik], l1 = kv.l[kv.nk + ik], l2 = kv.l[2 * kv.nk + ik];
        const int o0 = (0 * nl + l0) * TILE + threadIdx.x, o1 = (1 * nl + abs(l1)) * TILE + threadIdx.x,
                  o2 = (2 * nl + abs(l2)) * TILE + threadIdx.x;
        const double c0 = tcos[o0], s0 = tsin[o0];
        const double c1 = tcos[o1], s1 = (l1 < 0 ? -1.0 : 1.0) * tsin[o1];
        const double c2 = tcos[o2], s2 = (l2 < 0 ? -1.0 : 1.0) * tsin[o2];
        const double c01 = c0 * c1 - s0 * s1, s01 = s0 * c1 + c0 * s1;
        const double cc = c01 * c2 - s01 * s2, ss = s01 * c2 + c01 * s2;
        const double t = kv.weight[ik] * (ss * sf[ik] - cc * sf[kv.nk + ik]);
        ex += kvec[ik] * t;
        ey += kvec[kv.nk + ik] * t;
        ez += kvec[2 * kv.nk + ik] * t;
    }
    if (live) {
        ef[0 * a.n_pad + i] = ex * scale;
        ef[1 * a.n_pad + i] = ey * scale;
        ef[2 * a.n_pad + i] = ez * scale;
    }
}

void launch_recip_sf(const Cell& c, const SubsetDev& atoms, const KVecDev& kv, int chunk_begin, int chunk_end,
                     int atoms_per_chunk, double* sf_partial, cudaStream_t s) {
    const int n = chunk_end - chunk_begin;
    if (n <= 0) return;
    const size_t smem = sizeof(double) * (6 * (kv.kmax + 1) * RC_ATOMS + RC_ATOMS);
    static unsigned long long attr_set = 0;
    if (first_on_device(attr_set)) {
        cudaFuncSetAttribute(recip_sf_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    }
    recip_sf_kernel<<<n, RC_THREADS, smem, s>>>(c, atoms, kv, chunk_begin, atoms_per_chunk, sf_partial);
}

int launch_recip_finish(const KVecDev& kv, const double* sf_partial, int nchunks, double scale, double* sf,
                        double* energy_out, int k_begin, int k_end, cudaStream_t s, int mu_n, int mu_np, const double* mu,
                        double* mu_total) {
    const int kblocks = (kv.nk + 31) / 32;
    int launches = 0;
    if (kblocks + (mu ? 1 : 0) > 0) {
        chunk_sum_kernel<<<kblocks + (mu ? 1 : 0), 1024, 0, s>>>(kv.nk, sf_partial, nchunks, sf, kblocks, mu_n, mu_np, mu, mu_total);
        launches++;
    }
    if (energy_out) {
        recip_energy_kernel<<<1, 1024, 0, s>>>(kv, sf, scale, energy_out, k_begin, k_end);
        launches++;
    }
    return launches;
}

// Incremental structure factor: one thread per k-vector adds the change of the moved atoms' terms.
__global__ void __launch_bounds__(128) recip_delta_kernel(int nk, const double* __restrict__ kvec,
                                                          const MovedRec* __restrict__ moved, int nmoved,
                                                          double* __restrict__ sf) {
    __shared__ MovedRec smv[INC_MAX_MOVED];
    for (int m = threadIdx.x; m < nmoved; m += blockDim.x) smv[m] = moved[m];
    __syncthreads();
    const int ik = blockIdx.x * blockDim.x + threadIdx.x;
    if (ik >= nk) return;
    const double kx = kvec[ik], ky = kvec[nk + ik], kz = kvec[2 * nk + ik];
    double re = 0.0, im = 0.0;
    for (int m = 0; m < nmoved; m++) {
        const MovedRec& r = smv[m];
        if (r.q == 0.0 || (r.fl & F_FROZEN)) continue;  // coulombic.c:74: charged, mobile atoms only
        double sn, cn, so, co;
        sincos(kx * r.xn + ky * r.yn + kz * r.zn, &sn, &cn);
        sincos(kx * r.xo + ky * r.yo + kz * r.zo, &so, &co);
        re += r.q * (cn - co);
        im += r.q * (sn - so);
    }
    sf[ik] += re;
    sf[nk + ik] += im;
}

void launch_recip_delta(const KVecDev& kv, const double* kvec, const MovedRec* moved, int nmoved, double* sf, double scale,
                        double* energy_out, cudaStream_t s) {
    if (kv.nk > 0 && nmoved > 0) recip_delta_kernel<<<(kv.nk + 127) / 128, 128, 0, s>>>(kv.nk, kvec, moved, nmoved, sf);
    recip_energy_kernel<<<1, 1024, 0, s>>>(kv, sf, scale, energy_out, 0, kv.nk);
}

void launch_recip_field(const Cell& c, const AtomsDev& a, const KVecDev& kv, const double* kvec, const double* sf,
                        double scale, int row_begin, int row_end, double* ef, cudaStream_t s) {
    const int rows = row_end - row_begin;
    if (rows <= 0) return;
    const size_t smem = sizeof(double) * 6 * (kv.kmax + 1) * TILE;
    static unsigned long long attr_set = 0;
    if (first_on_device(attr_set)) {
        cudaFuncSetAttribute(recip_field_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    }
    recip_field_kernel<<<(rows + TILE - 1) / TILE, TILE, smem, s>>>(c, a, kv, kvec, sf, scale, row_begin, row_end, ef);
}

}  // namespace mpmc
