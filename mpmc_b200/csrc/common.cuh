// common.cuh -- shared device helpers of the mpmc_b200 energy engine (sm_100a, fp64).
//
// Geometry follows the reference's minimum_image() (src/energy/pairs.c:230-290):
//   img[p] = rint( sum_q reciprocal_basis[q][p] * d[q] ),  dimg[p] = d[p] - sum_q basis[q][p] * img[q]
// evaluated in registers for every pair, every time (the reference's pair_t cache is not kept).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "mpmc_b200.h"

namespace mpmc {

constexpr int TILE = 128;  // atoms per tile == threads per CTA in the pair / field / dipole kernels

// reference src/include/defines.h
constexpr double kHBAR2 = 1.11211999e-68;
constexpr double kHBAR4 = 1.23681087e-136;
constexpr double kKB = 1.3806503e-23;
constexpr double kKB2 = 1.90619525e-46;
constexpr double kM2A2 = 1.0e20;
constexpr double kM2A4 = 1.0e40;
constexpr double kAMU2KG = 1.66053873e-27;
constexpr double kDEBYE2SKA = 85.10597636;
constexpr double kMAXVALUE = 1.0e40;
constexpr double kSMALL_dR = 1.0e-12;
constexpr int kMAX_ITERATION_COUNT = 128;

// cudaFuncSetAttribute() applies to the current device only, and one process may drive several GPUs
// (mpmc_b200_create_group): true the first time a launcher runs with the current device (`seen` is its static bit set).
inline bool first_on_device(unsigned long long& seen) {
    int d = 0;
    cudaGetDevice(&d);
    const unsigned long long bit = 1ull << (d & 63);
    if (seen & bit) return false;
    seen |= bit;
    return true;
}

// Scalars every kernel needs, passed by value (lands in the constant bank).
struct Cell {
    double B[3][3];  // basis[q][p]
    double R[3][3];  // reciprocal_basis[q][p]
    double cutoff;
    double volume;
};

__device__ __forceinline__ double rint_magic(double x) {
#ifdef MPMC_RINT_FRND
    return rint(x);  // FRND.F64 on the conversion pipe: same value, off the fp64 pipe (A/B: make EXTRA=-DMPMC_RINT_FRND)
#else
    // round-half-even like rint() for |x| < 2^51; two DADDs instead of a conversion-pipe op
    const double m = 6755399441055744.0;  // 1.5 * 2^52
    return __dadd_rn(__dadd_rn(x, m), -m);
#endif
}

template <bool ORTHO>
__device__ __forceinline__ void min_image(const Cell& c, double& dx, double& dy, double& dz) {
    if (ORTHO) {
        dx -= c.B[0][0] * rint_magic(c.R[0][0] * dx);
        dy -= c.B[1][1] * rint_magic(c.R[1][1] * dy);
        dz -= c.B[2][2] * rint_magic(c.R[2][2] * dz);
    } else {
        double i0 = rint_magic(c.R[0][0] * dx + c.R[1][0] * dy + c.R[2][0] * dz);
        double i1 = rint_magic(c.R[0][1] * dx + c.R[1][1] * dy + c.R[2][1] * dz);
        double i2 = rint_magic(c.R[0][2] * dx + c.R[1][2] * dy + c.R[2][2] * dz);
        dx -= c.B[0][0] * i0 + c.B[1][0] * i1 + c.B[2][0] * i2;
        dy -= c.B[0][1] * i0 + c.B[1][1] * i1 + c.B[2][1] * i2;
        dz -= c.B[0][2] * i0 + c.B[1][2] * i1 + c.B[2][2] * i2;
    }
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_min(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_down_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_down_sync(0xffffffffu, v, o));
    return v;
}

// Fixed-order block reduction (deterministic run to run): warp shuffles, then warp 0 adds the
// per-warp partials in warp order.  Result valid in thread 0.  `scratch` holds >= 32 doubles.
template <int OP>  // 0 sum, 1 min, 2 max
__device__ __forceinline__ double block_reduce(double v, double* scratch) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = OP == 0 ? warp_sum(v) : OP == 1 ? warp_min(v) : warp_max(v);
    __syncthreads();
    if (lane == 0) scratch[w] = v;
    __syncthreads();
    if (w == 0) {
        double ident = OP == 0 ? 0.0 : OP == 1 ? 1.0e300 : -1.0e300;
        v = lane < nw ? scratch[lane] : ident;
        v = OP == 0 ? warp_sum(v) : OP == 1 ? warp_min(v) : warp_max(v);
    }
    return v;
}

// Orthorhombic box lengths for the cluster-level distance bounds.
struct BoxLen {
    double Lx, Ly, Lz, iLx, iLy, iLz;
};
__host__ __device__ inline BoxLen box_len(const Cell& c) {
    BoxLen b;
    b.Lx = c.B[0][0];
    b.Ly = c.B[1][1];
    b.Lz = c.B[2][2];
    b.iLx = c.R[0][0];
    b.iLy = c.R[1][1];
    b.iLz = c.R[2][2];
    return b;
}

// Lower bound of the minimum-image distance^2 between any atom of cluster ca (boxes `a`) and any atom of
// cluster cb (boxes `b`), orthorhombic box: per axis |minimg(centre_a - centre_b)| - (half_a + half_b), clamped
// at 0.  Box arrays are [6][nclus]: centre xyz, half extent xyz (cluster_box_kernel).
__device__ __forceinline__ double cluster_gap2(const BoxLen& k, const double* __restrict__ a, int na, int ca,
                                               const double* __restrict__ b, int nb, int cb) {
    double g2 = 0.0;
    {
        double d = a[0 * na + ca] - b[0 * nb + cb];
        d = fabs(fma(-k.Lx, rint(d * k.iLx), d)) - (a[3 * na + ca] + b[3 * nb + cb]);
        d = fmax(d, 0.0);
        g2 = fma(d, d, g2);
    }
    {
        double d = a[1 * na + ca] - b[1 * nb + cb];
        d = fabs(fma(-k.Ly, rint(d * k.iLy), d)) - (a[4 * na + ca] + b[4 * nb + cb]);
        d = fmax(d, 0.0);
        g2 = fma(d, d, g2);
    }
    {
        double d = a[2 * na + ca] - b[2 * nb + cb];
        d = fabs(fma(-k.Lz, rint(d * k.iLz), d)) - (a[5 * na + ca] + b[5 * nb + cb]);
        d = fmax(d, 0.0);
        g2 = fma(d, d, g2);
    }
    return g2;
}

// Upper bound of the same distance^2 (for the image the centres pick): per axis |minimg(centre_a - centre_b)| +
// (half_a + half_b).  A cluster pair whose span is inside the cutoff has EVERY atom pair inside it.
__device__ __forceinline__ double cluster_span2(const BoxLen& k, const double* __restrict__ a, int na, int ca,
                                                const double* __restrict__ b, int nb, int cb) {
    double s2 = 0.0;
    {
        double d = a[0 * na + ca] - b[0 * nb + cb];
        d = fabs(fma(-k.Lx, rint(d * k.iLx), d)) + (a[3 * na + ca] + b[3 * nb + cb]);
        s2 = fma(d, d, s2);
    }
    {
        double d = a[1 * na + ca] - b[1 * nb + cb];
        d = fabs(fma(-k.Ly, rint(d * k.iLy), d)) + (a[4 * na + ca] + b[4 * nb + cb]);
        s2 = fma(d, d, s2);
    }
    {
        double d = a[2 * na + ca] - b[2 * nb + cb];
        d = fabs(fma(-k.Lz, rint(d * k.iLz), d)) + (a[5 * na + ca] + b[5 * nb + cb]);
        s2 = fma(d, d, s2);
    }
    return s2;
}

// ---- fast fp64 helpers shared by the dipole kernels (K: a by-value kernel parameter struct with `double ec[12]` = 1/k!)
// 1/sqrt(a): MUFU.RSQ64H seed (2^-22.9) + one third-order step, no range checks (a is a squared distance
// between distinct atoms).  5 fp64 instructions like the library's fast path, without its slow-path test.
__device__ __forceinline__ double rsqrt_fast(double a) {
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(a));
    const double e = fma(a, -(y * y), 1.0);
    const double p = fma(e, 0.375, 0.5);
    return fma(p, y * e, y);
}

// e^{-x} for 0 <= x < 700: Cody-Waite reduction by ln2 with the magic-number rint, degree-11 Taylor
// polynomial on |r| <= ln2/2 (truncation error 6.4e-15 relative), exponent inserted with one integer add.
// The coefficients arrive through the constant bank (kernel parameter), so every DFMA takes its constant
// as an operand instead of materialising it with two moves.
template <class K>
__device__ __forceinline__ double exp_neg_fast(const K& k, double x) {
    const double magic = 6755399441055744.0;
    const double t = fma(x, -1.4426950408889634, magic);
    const double nd = t - magic;
    double r = fma(nd, -6.93147180369123816490e-01, -x);
    r = fma(nd, -1.90821492927058770002e-10, r);
    // even / odd split: two independent Horner chains in r^2 (depth 7 instead of 11; with two pairs in flight
    // that is four chains, enough to keep the half-rate fp64 pipe fed from one warp)
    const double r2 = r * r;
    double pe = k.ec[10], po = k.ec[11];
    pe = fma(pe, r2, k.ec[8]);
    po = fma(po, r2, k.ec[9]);
    pe = fma(pe, r2, k.ec[6]);
    po = fma(po, r2, k.ec[7]);
    pe = fma(pe, r2, k.ec[4]);
    po = fma(po, r2, k.ec[5]);
    pe = fma(pe, r2, k.ec[2]);
    po = fma(po, r2, k.ec[3]);
    pe = fma(pe, r2, 1.0);
    po = fma(po, r2, 1.0);
    const double p = fma(po, r, pe);
    const int n = __double2loint(t);
    return __hiloint2double(__double2hiint(p) + (n << 20), __double2loint(p));
}


// host side: ec[q] = 1/q!
inline void fill_exp_coeffs(double* ec) {
    double f = 1.0;
    for (int q = 0; q < 12; q++) {
        ec[q] = 1.0 / f;
        f *= (double)(q + 1);
    }
}

// Cluster-pair culling for the one-CTA-per-i-cluster cutoff kernels (K3, K7): the four warps test interleaved
// batches of 32 j-clusters against cluster ci and leave one pass mask per batch in shared memory; afterwards every
// warp walks the masks in order and takes every fourth passing cluster (position % 4 == warp), so the work is
// spread evenly over the warps whatever the spatial order, and the assignment -- hence every partial sum -- is a
// pure function of the geometry (bit-reproducible).  Every thread of the CTA must call.  b0/b1: batch range.
constexpr int CULL_MAX_BATCHES = 2048;  // 65,536 clusters = 2 M atoms
__device__ __forceinline__ void cull_masks(const BoxLen& box, const double* __restrict__ cbox, int nclus, int ci, double cull2,
                                           int b0, int b1, unsigned* __restrict__ smask) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int b = b0 + w; b < b1; b += nw) {
        const int cj = b * 32 + lane;
        bool pass = cj < nclus;
        if (pass && cull2 > 0.0) pass = cluster_gap2(box, cbox, nclus, ci, cbox, nclus, cj) <= cull2;
        const unsigned m = __ballot_sync(0xffffffffu, pass);
        if (lane == 0) smask[b - b0] = m;
    }
    __syncthreads();
}

}  // namespace mpmc
