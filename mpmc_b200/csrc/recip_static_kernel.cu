// recip_static_kernel.cu -- mpmc_b200_recip_energy_static(): the reciprocal-space Ewald energy (K2) for callers
// that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).
//
// No counterpart in the reference.  The quantity is the reference's
// coulombic_reciprocal() (src/energy/coulombic.c:42-95), the same one mpmc_b200_energy() reports as es_recip.
//
// A variant of recip_sf_kernel / chunk_sum_kernel / recip_energy_kernel (recip_kernel.cu) whose scratch -- the
// per-chunk partial structure factors and S(k) -- lives in ONE file-scope device array instead of buffers the context
// allocates, and whose k-vector list is not a buffer either: the hemisphere is addressed by a dense index over
// (l0, l1, l2) in the reference's loop order, points the reference skips get weight 0.  The entry point calls no
// cudaMalloc / cudaFree / allocating routine; the caller passes device pointers for inputs and the output.
//
// The array is held by every GPU that loads the library until the process exits (RS_SCRATCH_DOUBLES * 8 bytes =
// 4.9 MB).  All calls share it and it keeps the last call's contents: every slot a call reads is written by the same
// call first (nothing relies on zeros), and two calls that overlap on the GPU (different streams, or graphs replayed
// concurrently) race on it -- serialise them.
#include <cstdio>
#include <mutex>

#include "kernels.h"

namespace mpmc {

constexpr int RS_ATOMS = 64;      // atoms per shared-memory sub-chunk (as RC_ATOMS)
constexpr int RS_THREADS = 256;   // threads per CTA
constexpr int RS_PER_THREAD = 8;  // dense k indices per thread => at most 2048
constexpr int RS_MAX_KMAX = 7;    // (kmax+1)(2kmax+1)^2 = 1800 <= 2048 (the reference's default ewald_kmax is 7)
constexpr int RS_MAX_NKD = RS_THREADS * RS_PER_THREAD;
constexpr int RS_MAX_CHUNKS = 148;  // one CTA per SM of a B200
constexpr int RS_MAX_ATOMS = 1 << 20;
constexpr size_t RS_SCRATCH_DOUBLES = (size_t)RS_MAX_CHUNKS * 2 * RS_MAX_NKD + 2 * RS_MAX_NKD;

// [chunks][2][nkd] partial structure factors, then [2][nkd] S(k).  In HOST code this name is a host-side stand-in, not
// a device address: the launcher below resolves the address once per device with cudaGetSymbolAddress().
static __device__ __align__(256) double g_recip_scratch[RS_SCRATCH_DOUBLES];

struct RsCell {
    double R[3][3];  // reciprocal_basis[p][q]
    double alpha, scale;
    int kmax, nkd;
};

// dense index -> (l0, l1, l2), the loops of coulombic.c:56-58 with their skipped points marked invalid
__device__ __forceinline__ bool rs_decode(int ik, int kmax, int& l0, int& l1, int& l2) {
    const int w = 2 * kmax + 1;
    l2 = ik % w - kmax;
    l1 = (ik / w) % w - kmax;
    l0 = ik / (w * w);
    if (l0 > kmax) return false;
    if (l0 == 0 && (l1 < 0 || (l1 == 0 && l2 < 1))) return false;  // hemisphere, face not counted twice
    return l0 * l0 + l1 * l1 + l2 * l2 <= kmax * kmax;             // coulombic.c:60
}

__global__ void __launch_bounds__(RS_THREADS) recip_sf_static_kernel(const __grid_constant__ RsCell c, int n, const double* __restrict__ pos,
                                                                     const double* __restrict__ charge,
                                                                     const int* __restrict__ frozen, int atoms_per_chunk,
                                                                     double* __restrict__ sf_partial) {
    extern __shared__ double smem[];
    const int nl = c.kmax + 1;
    double* tcos = smem;  // [axis][l][atom]
    double* tsin = tcos + 3 * nl * RS_ATOMS;
    double* tq = tsin + 3 * nl * RS_ATOMS;

    const int a0 = blockIdx.x * atoms_per_chunk;
    const int a1 = min(a0 + atoms_per_chunk, n);

    int o0[RS_PER_THREAD], o1[RS_PER_THREAD], o2[RS_PER_THREAD];  // table offsets; o0 < 0: not a k-vector
    double sre[RS_PER_THREAD], sim[RS_PER_THREAD];
#pragma unroll
    for (int m = 0; m < RS_PER_THREAD; m++) {
        sre[m] = sim[m] = 0.0;
        int l0, l1, l2;
        const int ik = threadIdx.x + m * RS_THREADS;
        const bool ok = ik < c.nkd && rs_decode(ik, c.kmax, l0, l1, l2);
        o0[m] = ok ? (0 * nl + l0) * RS_ATOMS : -1;
        // the sign of l1 / l2 rides in bit 30 of the offset
        o1[m] = ok ? ((1 * nl + abs(l1)) * RS_ATOMS) | (l1 < 0 ? 1 << 30 : 0) : 0;
        o2[m] = ok ? ((2 * nl + abs(l2)) * RS_ATOMS) | (l2 < 0 ? 1 << 30 : 0) : 0;
    }

    for (int base = a0; base < a1; base += RS_ATOMS) {
        const int cnt = min(RS_ATOMS, a1 - base);
        __syncthreads();
        for (int w = threadIdx.x; w < 3 * nl * RS_ATOMS; w += RS_THREADS) {
            const int aa = w % RS_ATOMS, l = (w / RS_ATOMS) % nl, ax = w / (RS_ATOMS * nl);
            double cs = 1.0, sn = 0.0;
            if (aa < cnt) {
                const double* r = pos + 3 * (size_t)(base + aa);
                const double sfrac = c.R[0][ax] * r[0] + c.R[1][ax] * r[1] + c.R[2][ax] * r[2];
                sincospi(2.0 * (double)l * sfrac, &sn, &cs);
            }
            tcos[w] = cs;
            tsin[w] = sn;
        }
        for (int w = threadIdx.x; w < RS_ATOMS; w += RS_THREADS)  // coulombic.c:73-74: mobile, charged atoms only
            tq[w] = (w < cnt && !(frozen && frozen[base + w])) ? charge[base + w] : 0.0;
        __syncthreads();
#pragma unroll
        for (int m = 0; m < RS_PER_THREAD; m++) {
            if (o0[m] < 0) continue;  // uniform per thread across sub-chunks
            const int p0 = o0[m], p1 = o1[m] & ~(1 << 30), p2 = o2[m] & ~(1 << 30);
            const double sg1 = (o1[m] >> 30) ? -1.0 : 1.0, sg2 = (o2[m] >> 30) ? -1.0 : 1.0;
            double re = 0.0, im = 0.0;
            for (int aa = 0; aa < RS_ATOMS; aa++) {
                const double c0 = tcos[p0 + aa], s0 = tsin[p0 + aa];
                const double c1 = tcos[p1 + aa], s1 = sg1 * tsin[p1 + aa];
                const double c2 = tcos[p2 + aa], s2 = sg2 * tsin[p2 + aa];
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
    // every slot of this chunk is written (zeros where the index is not a k-vector): the array keeps old contents
    double* out = sf_partial + (size_t)blockIdx.x * 2 * c.nkd;
#pragma unroll
    for (int m = 0; m < RS_PER_THREAD; m++) {
        const int ik = threadIdx.x + m * RS_THREADS;
        if (ik >= c.nkd) break;
        out[ik] = sre[m];
        out[c.nkd + ik] = sim[m];
    }
}

// S(k) = sum over chunks, in chunk order per 32-way split as chunk_sum_kernel (recip_kernel.cu) does
__global__ void __launch_bounds__(1024) chunk_sum_static_kernel(int nkd, const double* __restrict__ sf_partial, int nchunks,
                                                                double* __restrict__ sf) {
    __shared__ double red[2][32][33];
    const int kx = threadIdx.x & 31, cy = threadIdx.x >> 5;
    const int ik = blockIdx.x * 32 + kx;
    double re = 0.0, im = 0.0;
    if (ik < nkd)
        for (int ch = cy; ch < nchunks; ch += 32) {
            re += sf_partial[(size_t)ch * 2 * nkd + ik];
            im += sf_partial[(size_t)ch * 2 * nkd + nkd + ik];
        }
    red[0][cy][kx] = re;
    red[1][cy][kx] = im;
    __syncthreads();
    if (cy < 2 && ik < nkd) {
        double v = 0.0;
        for (int q = 0; q < 32; q++) v += red[cy][q][kx];
        sf[cy * nkd + ik] = v;
    }
}

// E = 4 pi / V * sum_k exp(-k^2 / 4 alpha^2) / k^2 |S(k)|^2 (coulombic.c:61-66,86,92); fixed-order block reduction
__global__ void __launch_bounds__(1024) recip_energy_static_kernel(const __grid_constant__ RsCell c, const double* __restrict__ sf,
                                                                   double* __restrict__ energy_out) {
    __shared__ double red[32];
    double acc = 0.0;
    for (int ik = threadIdx.x; ik < c.nkd; ik += blockDim.x) {
        int l[3];
        if (!rs_decode(ik, c.kmax, l[0], l[1], l[2])) continue;
        double k2 = 0.0;
        for (int p = 0; p < 3; p++) {
            double kp = 0.0;
            for (int q = 0; q < 3; q++) kp += 2.0 * M_PI * c.R[p][q] * l[q];
            k2 += kp * kp;
        }
        const double re = sf[ik], im = sf[c.nkd + ik];
        acc += exp(-k2 / (4.0 * c.alpha * c.alpha)) / k2 * (re * re + im * im);
    }
    const double v = block_reduce<0>(acc, red);
    if (threadIdx.x == 0) *energy_out = v * c.scale;
}

// Device address of the scratch array on the CURRENT device, resolved once per device.
static int recip_static_scratch(double** out) {
    static std::mutex mtx;
    static double* addr[64] = {nullptr};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= 64) {
        set_last_error("mpmc_b200_recip_energy_static: no current CUDA device");
        return 10;
    }
    std::lock_guard<std::mutex> lock(mtx);
    if (!addr[dev]) {
        double* p = nullptr;
        e = cudaGetSymbolAddress((void**)&p, g_recip_scratch);
        if (e != cudaSuccess || !p) {
            char b[256];
            snprintf(b, sizeof(b), "mpmc_b200_recip_energy_static: cudaGetSymbolAddress failed on device %d (%s)", dev,
                     cudaGetErrorString(e));
            set_last_error(b);
            return 10;
        }
        addr[dev] = p;
    }
    *out = addr[dev];
    return 0;
}

}  // namespace mpmc

using namespace mpmc;

extern "C" int mpmc_b200_recip_energy_static(int n, const double* d_pos, const double* d_charge, const int* d_frozen,
                                             const double* reciprocal_basis, double volume, double ewald_alpha,
                                             int ewald_kmax, double* d_energy, void* cuda_stream) {
    char b[256];
    if (n < 0 || n > RS_MAX_ATOMS) {
        snprintf(b, sizeof(b), "mpmc_b200_recip_energy_static: n = %d is outside 0..%d (the static scratch is sized for that)",
                 n, RS_MAX_ATOMS);
        set_last_error(b);
        return 3;
    }
    if (ewald_kmax < 1 || ewald_kmax > RS_MAX_KMAX) {
        snprintf(b, sizeof(b), "mpmc_b200_recip_energy_static: ewald_kmax = %d is outside 1..%d (the static scratch holds %d k slots)",
                 ewald_kmax, RS_MAX_KMAX, RS_MAX_NKD);
        set_last_error(b);
        return 3;
    }
    if (!d_energy || !reciprocal_basis || (n > 0 && (!d_pos || !d_charge)) || !(volume > 0.0) || !(ewald_alpha > 0.0)) {
        set_last_error("mpmc_b200_recip_energy_static: null pointer, or volume / ewald_alpha not positive");
        return 3;
    }
    double* scratch = nullptr;
    if (int rc = recip_static_scratch(&scratch)) return rc;

    RsCell c;
    for (int p = 0; p < 3; p++)
        for (int q = 0; q < 3; q++) c.R[p][q] = reciprocal_basis[3 * p + q];
    c.alpha = ewald_alpha;
    c.scale = 4.0 * M_PI / volume;
    c.kmax = ewald_kmax;
    c.nkd = (ewald_kmax + 1) * (2 * ewald_kmax + 1) * (2 * ewald_kmax + 1);

    // chunks of whole sub-chunks, at most one CTA per SM
    const int subs = (n + RS_ATOMS - 1) / RS_ATOMS;
    const int nchunks = subs < 1 ? 1 : (subs < RS_MAX_CHUNKS ? subs : RS_MAX_CHUNKS);
    const int atoms_per_chunk = ((subs + nchunks - 1) / nchunks < 1 ? 1 : (subs + nchunks - 1) / nchunks) * RS_ATOMS;
    double* sf_partial = scratch;
    double* sf = scratch + (size_t)RS_MAX_CHUNKS * 2 * RS_MAX_NKD;
    cudaStream_t s = (cudaStream_t)cuda_stream;
    const size_t smem = sizeof(double) * (6 * (c.kmax + 1) * RS_ATOMS + RS_ATOMS);  // <= 25 KB: no attribute needed

    recip_sf_static_kernel<<<nchunks, RS_THREADS, smem, s>>>(c, n, d_pos, d_charge, d_frozen, atoms_per_chunk, sf_partial);
    chunk_sum_static_kernel<<<(c.nkd + 31) / 32, 1024, 0, s>>>(c.nkd, sf_partial, nchunks, sf);
    recip_energy_static_kernel<<<1, 1024, 0, s>>>(c, sf, d_energy);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        snprintf(b, sizeof(b), "mpmc_b200_recip_energy_static: launch failed (%s)", cudaGetErrorString(e));
        set_last_error(b);
        return 10;
    }
    return 0;
}
