// ewald_full_kernel.cu -- K7: full-Ewald polarization.
//
// Reference loops replaced (relative to the reference tree, src/polarization/polar_ewald.c):
//   induced_real_term()   :176-230   cutoff pairs of polarizable atoms, erfc-screened damped tensor
//   induced_recip_term()  :232-294   k-sum of the dipole structure factor, all atoms
//   induced_corr_term()   :296-317   surface term -4 pi/(3V) sum(mu) and self term 4 a^3/(3 sqrt(pi)) mu_i
//   init_dipoles_ewald / new_dipoles / ewald_palmo_contraction  :135-148, :319-371
//   calc_dipole_rrms / are_we_done_yet (thole_iterative.c:61-117) as ewald_full() :375-418 calls them
//
// Layout: the polarizable subset in spatial order (the same packed image the symmetric K4 uses), solver
// vectors [3][n_pad].  Nothing is accumulated with atomics: the real term is row-owned (one warp per
// 32-atom i-cluster, j-clusters culled by their bounding boxes against the cutoff, j range split into
// `jsplits` slices written to separate partial rows), the dipole structure factor is summed per chunk
// and then in chunk order, the per-atom k-sum runs in k order.
//
// Two reference quirks are reproduced because parity is against the reference as it is:
//   * polar_ewald.c:265-266 overwrites `kweight` in a loop over p, so only k_z survives and the same
//     scalar is added to all three components of the reciprocal induced field (:282-286);
//   * SqrtPi is the truncated 1.77245385091 (:7).
#include "kernels.h"

namespace mpmc {

namespace {

constexpr double kOneOverSqrtPi = 0.5641895835477562869480794515607725858440506293289988;
constexpr double kSqrtPiRef = 1.77245385091;

__device__ __forceinline__ double flip_sign_if(double v, bool neg) {
    return __hiloint2double(__double2hiint(v) ^ (neg ? 0x80000000 : 0), __double2loint(v));
}

struct JRec {
    double x, y, z, mx, my, mz;
    int valid, pad;
};

// E_i += T_ij mu_j with T = 3 d d^T s2 / r^5 - s1 / r^3 (polar_ewald.c:196-226)
template <bool ORTHO>
__device__ __forceinline__ void induced_real_pair(const Cell& c, const EwaldFullOpts& k, double xi, double yi, double zi,
                                                  const JRec& r, double& ex, double& ey, double& ez) {
    double dx = xi - r.x, dy = yi - r.y, dz = zi - r.z;
    min_image<ORTHO>(c, dx, dy, dz);
    const double r2 = dx * dx + dy * dy + dz * dz;
    const double rr = sqrt(r2);
    if (rr > c.cutoff) return;  // :188 (r == cutoff is kept)
    const double ir = 1.0 / rr, ir3 = ir * ir * ir, ir5 = ir * ir * ir3;
    const double a = k.a;
    const double erfcar = erfc(a * rr), ex2 = exp(-a * a * r2);
    const double t = k.damp * rr, et = exp(-t);
    const double df2 = (1.0 + t + 0.5 * t * t) * et;            // damp_factor(t, 2)
    const double df3 = (1.0 + t + 0.5 * t * t + t * t * t / 6.0) * et;  // damp_factor(t, 3)
    const double g = 2.0 * a * rr * kOneOverSqrtPi * ex2;
    const double s1 = erfcar + g - df2;
    const double s2 = erfcar + g + 4.0 * a * a * a * rr * r2 / 3.0 * kOneOverSqrtPi * ex2 - df3;
    const double dmu = dx * r.mx + dy * r.my + dz * r.mz;
    const double w2 = 3.0 * s2 * ir5 * dmu, w1 = s1 * ir3;
    ex += w2 * dx - w1 * r.mx;
    ey += w2 * dy - w1 * r.my;
    ez += w2 * dz - w1 * r.mz;
}

template <bool ORTHO>
__global__ void __launch_bounds__(TILE) induced_real_cluster_kernel(Cell c, SubsetDev a, EwaldFullOpts k, BoxLen box,
                                                                    const double* __restrict__ cbox, double cull2,
                                                                    const double* __restrict__ mu, int batches_per_slice,
                                                                    double* __restrict__ part) {
    // One CTA per i-cluster: the four warps scan interleaved batches of 32 j-clusters (a warp's serial scan is the
    // critical path; ncu of the one-warp-per-cluster form: 13.7 % achieved occupancy, 3.58 ms at 46k atoms) and
    // their partial fields are added in warp order.
    __shared__ JRec rec[4][32];
    __shared__ double sred[4][3][32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int nclus = a.n_pad >> 5, np = a.n_pad;
    const int ci = blockIdx.x;
    if (ci >= nclus) return;  // whole CTA
    const int i = ci * 32 + lane;
    const double xi = a.x[i], yi = a.y[i], zi = a.z[i];
    double ex = 0.0, ey = 0.0, ez = 0.0;
    __shared__ unsigned smask[CULL_MAX_BATCHES];
    const int nb_all = (nclus + 31) >> 5;
    const int b0 = blockIdx.y * batches_per_slice, b1 = min(nb_all, b0 + batches_per_slice);
    cull_masks(box, cbox, nclus, ci, cull2, b0, b1, smask);
    int pos = 0;
    for (int b = b0; b < b1; b++) {
        unsigned mask = smask[b - b0];
        while (mask) {
            const int bit = __ffs(mask) - 1;
            mask &= mask - 1;
            if (((pos++) & 3) != w) continue;  // warp-uniform
            const int cj = b * 32 + bit;
            __syncwarp();
            {
                const int j = cj * 32 + lane;
                JRec r;
                r.x = a.x[j];
                r.y = a.y[j];
                r.z = a.z[j];
                r.mx = mu[j];
                r.my = mu[np + j];
                r.mz = mu[2 * np + j];
                r.valid = j < a.n;
                r.pad = 0;
                rec[w][lane] = r;
            }
            __syncwarp();
            const bool self_cluster = (cj == ci);
            for (int s = 0; s < 32; s++) {
                const JRec& r = rec[w][s];
                if (!r.valid) continue;
                if (self_cluster && s == lane) continue;
                induced_real_pair<ORTHO>(c, k, xi, yi, zi, r, ex, ey, ez);
            }
        }
    }
    sred[w][0][lane] = ex;
    sred[w][1][lane] = ey;
    sred[w][2][lane] = ez;
    __syncthreads();
    if (w != 0) return;
    double* out = part + (size_t)blockIdx.y * 3 * np;
    out[i] = ((sred[0][0][lane] + sred[1][0][lane]) + sred[2][0][lane]) + sred[3][0][lane];
    out[np + i] = ((sred[0][1][lane] + sred[1][1][lane]) + sred[2][1][lane]) + sred[3][1][lane];
    out[2 * np + i] = ((sred[0][2][lane] + sred[1][2][lane]) + sred[2][2][lane]) + sred[3][2][lane];
}

// Dipole structure factor: P(k) = sum_j (k . mu_j) e^{i k.r_j}  (polar_ewald.c:268-276); same phase
// tables as recip_sf_kernel, the weight depends on k.
constexpr int DS_ATOMS = 64;
constexpr int DS_THREADS = 256;

__global__ void __launch_bounds__(DS_THREADS) dipole_sf_kernel(Cell c, SubsetDev at, KVecDev kv,
                                                              const double* __restrict__ kvec,
                                                              const double* __restrict__ mu, int atoms_per_chunk,
                                                              double* __restrict__ sf_partial) {
    extern __shared__ double smem[];
    const int nl = kv.kmax + 1, np = at.n_pad;
    double* tcos = smem;
    double* tsin = tcos + 3 * nl * DS_ATOMS;
    double* tm = tsin + 3 * nl * DS_ATOMS;  // [3][DS_ATOMS]
    const int a0 = blockIdx.x * atoms_per_chunk;
    const int a1 = min(a0 + atoms_per_chunk, at.n);
    double sre[8], sim[8];
#pragma unroll
    for (int m = 0; m < 8; m++) sre[m] = sim[m] = 0.0;
    for (int base = a0; base < a1; base += DS_ATOMS) {
        const int cnt = min(DS_ATOMS, a1 - base);
        __syncthreads();
        for (int w = threadIdx.x; w < 3 * nl * DS_ATOMS; w += DS_THREADS) {
            const int aa = w % DS_ATOMS, l = (w / DS_ATOMS) % nl, ax = w / (DS_ATOMS * nl);
            double cs = 1.0, sn = 0.0;
            if (aa < cnt) {
                const int g = base + aa;
                const double sfrac = c.R[0][ax] * at.x[g] + c.R[1][ax] * at.y[g] + c.R[2][ax] * at.z[g];
                sincospi(2.0 * (double)l * sfrac, &sn, &cs);
            }
            tcos[w] = cs;
            tsin[w] = sn;
        }
        for (int w = threadIdx.x; w < 3 * DS_ATOMS; w += DS_THREADS) {
            const int aa = w % DS_ATOMS, d = w / DS_ATOMS;
            tm[w] = (aa < cnt) ? mu[d * np + base + aa] : 0.0;
        }
        __syncthreads();
#pragma unroll
        for (int m = 0; m < 8; m++) {
            const int ik = threadIdx.x + m * DS_THREADS;
            if (ik >= kv.nk) break;
            const int l0 = kv.l[ik], l1 = kv.l[kv.nk + ik], l2 = kv.l[2 * kv.nk + ik];
            const double kx = kvec[ik], ky = kvec[kv.nk + ik], kz = kvec[2 * kv.nk + ik];
            const int o0 = (0 * nl + l0) * DS_ATOMS, o1 = (1 * nl + abs(l1)) * DS_ATOMS, o2 = (2 * nl + abs(l2)) * DS_ATOMS;
            const double sg1 = l1 < 0 ? -1.0 : 1.0, sg2 = l2 < 0 ? -1.0 : 1.0;
            double re = 0.0, im = 0.0;
            for (int aa = 0; aa < DS_ATOMS; aa++) {
                const double c0 = tcos[o0 + aa], s0 = tsin[o0 + aa];
                const double c1 = tcos[o1 + aa], s1 = sg1 * tsin[o1 + aa];
                const double c2 = tcos[o2 + aa], s2 = sg2 * tsin[o2 + aa];
                const double c01 = c0 * c1 - s0 * s1, s01 = s0 * c1 + c0 * s1;
                const double cc = c01 * c2 - s01 * s2, ss = s01 * c2 + c01 * s2;
                const double km = kx * tm[aa] + ky * tm[DS_ATOMS + aa] + kz * tm[2 * DS_ATOMS + aa];
                re += km * cc;
                im += km * ss;
            }
            sre[m] += re;
            sim[m] += im;
        }
    }
    double* out = sf_partial + (size_t)blockIdx.x * 2 * kv.nk;
#pragma unroll
    for (int m = 0; m < 8; m++) {
        const int ik = threadIdx.x + m * DS_THREADS;
        if (ik >= kv.nk) break;
        out[ik] = sre[m];
        out[kv.nk + ik] = sim[m];
    }
}

// Grouped form of the same sum (kmax <= 7): one thread per (l0, l1) GROUP of the lexicographic k list -- its k-vectors
// differ only in l2 (a contiguous run l2 = d0 .. d0 + count - 1, count <= 15) -- so the phase product of the first two
// axes and l0 M0 + l1 M1 (k . mu = sum_q l_q M_q with M_q = 2 pi sum_p R[p][q] mu_p) are formed once per (group, atom):
// 7 executed fp64 instructions per (k, atom) + 6 per (group, atom) against 17 in dipole_sf_kernel.  Groups arrive
// sorted by count (engine.cu:build_kvectors) so the lanes of a warp run the same number of l2 steps; SG_SUB threads
// share a group (atoms aa = sub, sub + SG_SUB, ... of every 64-atom batch) and are added in sub order at the end.
// Phase tables: [atom][axis][l] (conflict-free for lanes with different l), e^{i l theta} by recurrence from l = 1.
constexpr int SG_MAXD = 15;
constexpr int SG_ATOMS = 64;
constexpr int SG_SUB = 4;

__global__ void __launch_bounds__(512) dipole_sf_grouped_kernel(Cell c, SubsetDev at, int nk, int nl, int ngroups, int gthreads,
                                                                const int4* __restrict__ groups,
                                                                const double* __restrict__ mu, int atoms_per_chunk,
                                                                double* __restrict__ sf_partial) {
    extern __shared__ double smem[];
    const int ST = 3 * nl, np = at.n_pad;
    double* tcos = smem;                    // [SG_ATOMS][3 nl]
    double* tsin = tcos + SG_ATOMS * ST;
    double* tm = tsin + SG_ATOMS * ST;      // [SG_ATOMS][3]: M_q
    const int gi = threadIdx.x % gthreads, sub = threadIdx.x / gthreads;
    const bool has = gi < ngroups;
    const int4 g = has ? groups[gi] : make_int4(0, 0, 0, 0);
    const int l0 = g.y, l1 = g.z, d0 = (g.w & 0xff) - 64, count = has ? (g.w >> 8) : 0;
    const int wmax = __reduce_max_sync(0xffffffffu, count);
    const int al1 = abs(l1);
    const double l0d = (double)l0, l1d = (double)l1;
    double re[SG_MAXD], im[SG_MAXD];
#pragma unroll
    for (int q = 0; q < SG_MAXD; q++) re[q] = im[q] = 0.0;
    const int a0 = blockIdx.x * atoms_per_chunk;
    const int a1 = min(a0 + atoms_per_chunk, at.n);
    for (int base = a0; base < a1; base += SG_ATOMS) {
        const int cnt = min(SG_ATOMS, a1 - base);
        __syncthreads();
        for (int w = threadIdx.x; w < 3 * SG_ATOMS; w += blockDim.x) {
            const int aa = w % SG_ATOMS, ax = w / SG_ATOMS;
            double c1 = 1.0, s1 = 0.0, m = 0.0;
            if (aa < cnt) {
                const int gidx = base + aa;
                const double sfrac = c.R[0][ax] * at.x[gidx] + c.R[1][ax] * at.y[gidx] + c.R[2][ax] * at.z[gidx];
                sincospi(2.0 * sfrac, &s1, &c1);
                m = 2.0 * M_PI * (c.R[0][ax] * mu[gidx] + c.R[1][ax] * mu[np + gidx] + c.R[2][ax] * mu[2 * (size_t)np + gidx]);
            }
            double cl = 1.0, sl = 0.0;
            double* tc = tcos + aa * ST + ax * nl;
            double* ts = tsin + aa * ST + ax * nl;
            for (int l = 0; l < nl; l++) {
                tc[l] = cl;
                ts[l] = sl;
                const double cn = cl * c1 - sl * s1;
                sl = sl * c1 + cl * s1;
                cl = cn;
            }
            tm[aa * 3 + ax] = m;
        }
        __syncthreads();
        if (wmax > 0)
            for (int aa = sub; aa < cnt; aa += SG_SUB) {
                const double* tc = tcos + aa * ST;
                const double* ts = tsin + aa * ST;
                const double c0 = tc[l0], s0 = ts[l0];
                const double c1 = tc[nl + al1], s1 = flip_sign_if(ts[nl + al1], l1 < 0);
                const double c01 = c0 * c1 - s0 * s1, s01 = s0 * c1 + c0 * s1;
                const double m2 = tm[aa * 3 + 2];
                const double km01 = l0d * tm[aa * 3] + l1d * tm[aa * 3 + 1];
#pragma unroll
                for (int q = 0; q < SG_MAXD; q++) {
                    if (q < wmax) {  // warp-uniform
                        const int d = d0 + q, ad = min(abs(d), nl - 1);
                        const double c2 = tc[2 * nl + ad], s2 = flip_sign_if(ts[2 * nl + ad], d < 0);
                        const double cc = c01 * c2 - s01 * s2, ss = s01 * c2 + c01 * s2;
                        const double km = km01 + (double)d * m2;
                        re[q] += km * cc;
                        im[q] += km * ss;
                    }
                }
            }
    }
    // add the SG_SUB partial sums of a group in sub order (staged through the table memory)
    double* stage = smem;  // [2 SG_MAXD][gthreads]
    for (int r = 1; r < SG_SUB; r++) {
        __syncthreads();
        if (sub == r) {
#pragma unroll
            for (int q = 0; q < SG_MAXD; q++) {
                stage[(2 * q) * gthreads + gi] = re[q];
                stage[(2 * q + 1) * gthreads + gi] = im[q];
            }
        }
        __syncthreads();
        if (sub == 0) {
#pragma unroll
            for (int q = 0; q < SG_MAXD; q++) {
                re[q] += stage[(2 * q) * gthreads + gi];
                im[q] += stage[(2 * q + 1) * gthreads + gi];
            }
        }
    }
    if (sub == 0 && has) {
        double* out = sf_partial + (size_t)blockIdx.x * 2 * nk;
#pragma unroll
        for (int q = 0; q < SG_MAXD; q++)
            if (q < count) {
                out[g.x + q] = re[q];
                out[nk + g.x + q] = im[q];
            }
    }
}

// sum of the dipoles (polar_ewald.c:303-307), one CTA, fixed order
__global__ void __launch_bounds__(1024) total_mu_kernel(int n, int np, const double* __restrict__ mu, double* __restrict__ out) {
    __shared__ double red[32];
    for (int d = 0; d < 3; d++) {
        double acc = 0.0;
        for (int i = threadIdx.x; i < n; i += blockDim.x) acc += mu[d * np + i];
        acc = block_reduce<0>(acc, red);
        if (threadIdx.x == 0) out[d] = acc;
        __syncthreads();
    }
}

// ef_induced of the rows of `rows`: real-space partial rows (if any) + the reciprocal term
//   g_i = sum_k w_k ( -sin(k.r_i) Im P(k) - cos(k.r_i) Re P(k) ),  w_k = scale exp(-k^2/4a^2)/k^2 k_z
// added to all three components (:278-287) + surface and self terms (:310-314).
//
// One thread per atom, 64-thread CTAs (726 CTAs at 46k atoms: the 128-thread form left 148 SMs with 2 or 3 CTAs each).
// The k list is lexicographic in (l0, l1, l2) (engine.cu:build_kvectors), so the phase product of the first two axes is
// kept while l2 runs; w_k Re P / w_k Im P are formed once per CTA; sin(-x) is a sign-bit flip.  Executed fp64
// instructions per (k, atom): 6 + 4 per (l0, l1) group (~1 amortised), against 17 in the first form.
constexpr int FIN_TILE = 64;

__global__ void __launch_bounds__(FIN_TILE) induced_finish_kernel(Cell c, SubsetDev rows, KVecDev kv, const double* __restrict__ kvec,
                                                                  const double* __restrict__ sf, double scale,
                                                                  const double* __restrict__ part, int jsplits, int part_np,
                                                                  const double* __restrict__ total_mu, double surf, double selfc,
                                                                  const double* __restrict__ mu, double* __restrict__ out) {
    extern __shared__ double smem[];
    const int nl = kv.kmax + 1, np = rows.n_pad, nk = kv.nk;
    double* tcos = smem;  // [axis][l][thread]
    double* tsin = tcos + 3 * nl * FIN_TILE;
    double* kA = tsin + 3 * nl * FIN_TILE;  // w_k Re P(k)
    double* kB = kA + nk;                   // w_k Im P(k)
    int* lpack = reinterpret_cast<int*>(kB + nk);
    const int i = blockIdx.x * FIN_TILE + threadIdx.x;
    const bool live = i < rows.n;
    const double x = live ? rows.x[i] : 0.0, y = live ? rows.y[i] : 0.0, z = live ? rows.z[i] : 0.0;
    for (int ik = threadIdx.x; ik < nk; ik += FIN_TILE) {
        const double wk = scale * kv.weight[ik] * kvec[2 * nk + ik];
        kA[ik] = wk * sf[ik];
        kB[ik] = wk * sf[nk + ik];
        lpack[ik] = kv.l[ik] | ((kv.l[nk + ik] + 64) << 8) | ((kv.l[2 * nk + ik] + 64) << 16);
    }
    for (int ax = 0; ax < 3; ax++) {  // e^{i l theta}, l = 0 .. kmax, by recurrence from one sincospi
        const double sfrac = c.R[0][ax] * x + c.R[1][ax] * y + c.R[2][ax] * z;
        double s1, c1, cl = 1.0, sl = 0.0;
        sincospi(2.0 * sfrac, &s1, &c1);
        for (int l = 0; l < nl; l++) {
            tcos[(ax * nl + l) * FIN_TILE + threadIdx.x] = cl;
            tsin[(ax * nl + l) * FIN_TILE + threadIdx.x] = sl;
            const double cn = cl * c1 - sl * s1;
            sl = sl * c1 + cl * s1;
            cl = cn;
        }
    }
    __syncthreads();
    double g = 0.0;
    int p01 = -1;
    double c01 = 1.0, s01 = 0.0;
    for (int ik = 0; ik < nk; ik++) {
        const int lp = lpack[ik];
        if ((lp & 0xffff) != p01) {  // CTA-uniform
            p01 = lp & 0xffff;
            const int l0 = lp & 0xff, l1 = ((lp >> 8) & 0xff) - 64;
            const int o0 = (0 * nl + l0) * FIN_TILE + threadIdx.x, o1 = (1 * nl + abs(l1)) * FIN_TILE + threadIdx.x;
            const double c0 = tcos[o0], s0 = tsin[o0];
            const double c1 = tcos[o1], s1 = flip_sign_if(tsin[o1], l1 < 0);
            c01 = c0 * c1 - s0 * s1;
            s01 = s0 * c1 + c0 * s1;
        }
        const int l2 = ((lp >> 16) & 0xff) - 64;
        const int o2 = (2 * nl + abs(l2)) * FIN_TILE + threadIdx.x;
        const double c2 = tcos[o2], s2 = flip_sign_if(tsin[o2], l2 < 0);
        const double cc = c01 * c2 - s01 * s2, ss = s01 * c2 + c01 * s2;
        g -= ss * kB[ik] + cc * kA[ik];
    }
    if (i >= np) return;
    for (int d = 0; d < 3; d++) {
        double e = 0.0;
        if (live) {
            for (int js = 0; js < jsplits; js++) e += part[((size_t)js * 3 + d) * part_np + i];
            e += g;
            e += surf * total_mu[d] + (mu ? selfc * mu[d * np + i] : 0.0);
        }
        out[d * np + i] = e;
    }
}

// calc_dipole_rrms (previous pass' new_mu / old_mu, as ewald_full() orders it) + new_dipoles (:319-345)
// + the are_we_done_yet residual
__global__ void ewald_update_kernel(int n, int np, const double* __restrict__ alpha, UpdateOpts o,
                                    const double* __restrict__ es, const double* __restrict__ eind,
                                    double* __restrict__ mu, double* __restrict__ mu_old, double* __restrict__ mu_new,
                                    double* __restrict__ rrms, unsigned long long* __restrict__ max_err2_bits) {
    __shared__ double red[32];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    double worst = 0.0;
    if (i < np) {
        const double al = i < n ? alpha[i] : 0.0;
        if (o.do_rrms) {
            double num = 0.0, den = 0.0;
            for (int d = 0; d < 3; d++) {
                const double nw = mu_new[d * np + i], df = nw - mu_old[d * np + i];
                num += df * df;
                den += nw * nw;
            }
            double v = sqrt(num / den);
            if (!isfinite(v)) v = 0.0;
            rrms[i] = v;
        }
        const double wold = o.esor ? exp(-o.gamma * (double)(o.iteration + 1)) : 0.0;
        for (int d = 0; d < 3; d++) {
            const double old = mu[d * np + i];
            double nw = al * (es[d * np + i] + eind[d * np + i]);
            if (o.sor)
                nw = o.gamma * nw + (1.0 - o.gamma) * old;
            else if (o.esor)
                nw = (1.0 - wold) * nw + wold * old;
            mu_old[d * np + i] = old;
            mu_new[d * np + i] = nw;
            mu[d * np + i] = nw;
            const double df = nw - old;
            worst = fmax(worst, df * df);
        }
    }
    worst = block_reduce<2>(worst, red);
    if (threadIdx.x == 0 && max_err2_bits) atomicMax(max_err2_bits, (unsigned long long)__double_as_longlong(worst));
}

// ewald_palmo_contraction (:347-371): eic = ef_induced - (new_mu / alpha - ef_static), alpha != 0
__global__ void ewald_palmo_finish_kernel(int n, int np, const double* __restrict__ alpha, const double* __restrict__ es,
                                          const double* __restrict__ eind, const double* __restrict__ mu_new,
                                          double* __restrict__ eic) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double al = alpha[i];
    if (al == 0.0) return;
    for (int d = 0; d < 3; d++) eic[d * np + i] = eind[d * np + i] - (mu_new[d * np + i] / al - es[d * np + i]);
}


// ---- pair list of the real-space term ---------------------------------------------------------------------------------
// The tensor of an in-cutoff pair depends on the positions only, so erfc / exp / rsqrt are evaluated ONCE per energy():
// a CSR list (row = atom i, four consecutive segments = the four warps of its cluster's CTA) holds, per in-cutoff pair,
// the minimum-image vector, the two coefficients w1 = s1 / r^3, w2 = 3 s2 / r^5 and the partner index (48 bytes).  Every
// solver pass then streams the list (HBM / L2 bound, ~25 fp64 instructions per pair) instead of re-evaluating ~400.
// Traversal, and therefore the order of every sum, is the deterministic pass-mask walk of induced_real_cluster_kernel.
struct PairEntry {
    double dx, dy, dz, w1, w2;
    long long j;
};

// `sqrt(r2) > cutoff` (polar_ewald.c:188) without the square root except in a 1e-12-wide shell around the cutoff
struct CutTest {
    double cut, lo, hi;
};
__device__ __forceinline__ bool inside_cutoff(const CutTest& t, double r2) {
    if (r2 < t.lo) return true;
    if (r2 > t.hi) return false;
    return !(sqrt(r2) > t.cut);
}

struct QueuedPair {
    double dx, dy, dz;
    unsigned long long dest;
    int j, pad;
};

// erfc / exp evaluated for ONE pair per lane (dense: the queue below holds in-cutoff pairs only)
__device__ __forceinline__ void list_emit(const EwaldFullOpts& k, const QueuedPair& q, PairEntry* __restrict__ list) {
    const double r2 = q.dx * q.dx + q.dy * q.dy + q.dz * q.dz;
    const double rr = sqrt(r2);
    const double ir = 1.0 / rr, ir3 = ir * ir * ir, ir5 = ir * ir * ir3;
    const double aa = k.a;
    const double erfcar = erfc(aa * rr), ex2 = exp(-aa * aa * r2);
    const double t = k.damp * rr, et = exp(-t);
    const double df2 = (1.0 + t + 0.5 * t * t) * et;
    const double df3 = (1.0 + t + 0.5 * t * t + t * t * t / 6.0) * et;
    const double g = 2.0 * aa * rr * kOneOverSqrtPi * ex2;
    const double s1 = erfcar + g - df2;
    const double s2 = erfcar + g + 4.0 * aa * aa * aa * rr * r2 / 3.0 * kOneOverSqrtPi * ex2 - df3;
    PairEntry e;
    e.dx = q.dx;
    e.dy = q.dy;
    e.dz = q.dz;
    e.w1 = s1 * ir3;
    e.w2 = 3.0 * s2 * ir5;
    e.j = q.j;
    list[q.dest] = e;
}

// FILL = false: counts per (atom, warp slot).  FILL = true: the same walk; the lanes of a warp own the 32 atoms of the
// i-cluster and meet an in-cutoff partner in ~1 of 6 tests, so the in-cutoff pairs are first compacted (ballot + prefix)
// into a 64-entry ring per warp and the erfc / exp body runs on 32 queued pairs at a time, one per lane.
template <bool ORTHO, bool FILL>
__global__ void __launch_bounds__(TILE) induced_list_kernel(Cell c, SubsetDev a, EwaldFullOpts k, BoxLen box,
                                                            const double* __restrict__ cbox, double cull2,
                                                            const unsigned long long* __restrict__ offs,
                                                            unsigned int* __restrict__ cnt, PairEntry* __restrict__ list) {
    __shared__ double jx[4][32], jy[4][32], jz[4][32];
    __shared__ unsigned smask[CULL_MAX_BATCHES];
    __shared__ QueuedPair ring[FILL ? 4 : 1][FILL ? 64 : 1];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int nclus = a.n_pad >> 5;
    const int ci = blockIdx.x;
    if (ci >= nclus) return;
    const int i = ci * 32 + lane;
    const bool live = i < a.n;
    const double xi = a.x[i], yi = a.y[i], zi = a.z[i];
    const int nb = min((nclus + 31) >> 5, CULL_MAX_BATCHES);
    cull_masks(box, cbox, nclus, ci, cull2, 0, nb, smask);
    CutTest ct;
    ct.cut = c.cutoff;
    ct.lo = c.cutoff * c.cutoff * (1.0 - 1e-12);
    ct.hi = c.cutoff * c.cutoff * (1.0 + 1e-12);
    unsigned int n = 0;
    unsigned long long at = FILL ? offs[(size_t)i * 4 + w] : 0ull;
    int head = 0, qn = 0;  // warp-uniform
    const unsigned lt = (1u << lane) - 1u;
    int pos = 0;
    for (int b = 0; b < nb; b++) {
        unsigned mask = smask[b];
        while (mask) {
            const int bit = __ffs(mask) - 1;
            mask &= mask - 1;
            if (((pos++) & 3) != w) continue;
            const int cj = b * 32 + bit;
            __syncwarp();
            jx[w][lane] = a.x[cj * 32 + lane];
            jy[w][lane] = a.y[cj * 32 + lane];
            jz[w][lane] = a.z[cj * 32 + lane];
            __syncwarp();
            const int jn = min(32, a.n - cj * 32);
            for (int s = 0; s < jn; s++) {
                double dx = xi - jx[w][s], dy = yi - jy[w][s], dz = zi - jz[w][s];
                min_image<ORTHO>(c, dx, dy, dz);
                const double r2 = dx * dx + dy * dy + dz * dz;
                const bool in = live && !(cj == ci && s == lane) && inside_cutoff(ct, r2);
                if (!FILL) {
                    n += in ? 1u : 0u;
                } else {
                    const unsigned m = __ballot_sync(0xffffffffu, in);
                    if (m == 0u) continue;
                    if (in) {
                        QueuedPair& q = ring[w][(head + qn + __popc(m & lt)) & 63];
                        q.dx = dx;
                        q.dy = dy;
                        q.dz = dz;
                        q.dest = at++;
                        q.j = cj * 32 + s;
                    }
                    qn += __popc(m);
                    if (qn >= 32) {
                        __syncwarp();
                        list_emit(k, ring[w][(head + lane) & 63], list);
                        __syncwarp();
                        head = (head + 32) & 63;
                        qn -= 32;
                    }
                }
            }
        }
    }
    if (FILL) {
        __syncwarp();
        if (lane < qn) list_emit(k, ring[w][(head + lane) & 63], list);
    } else
        cnt[(size_t)i * 4 + w] = n;
}

// exclusive scan of m counters (one CTA; m = 4 n_pad ~ 2e5): offs[q] = sum_{p<q} cnt[p], offs[m] = total
__global__ void __launch_bounds__(1024) list_scan_kernel(const unsigned int* __restrict__ cnt, int m,
                                                         unsigned long long* __restrict__ offs) {
    __shared__ unsigned long long part[1024];
    const int t = threadIdx.x;
    const int chunk = (m + 1023) / 1024;
    const int b = t * chunk, e = min(b + chunk, m);
    unsigned long long sum = 0;
    for (int q = b; q < e; q++) sum += cnt[q];
    part[t] = sum;
    __syncthreads();
    if (t == 0) {
        unsigned long long run = 0;
        for (int q = 0; q < 1024; q++) {
            const unsigned long long v = part[q];
            part[q] = run;
            run += v;
        }
        offs[m] = run;
    }
    __syncthreads();
    unsigned long long run = part[t];
    for (int q = b; q < e; q++) {
        offs[q] = run;
        run += cnt[q];
    }
}

// one warp per atom: E_i = sum over its list entries of w2 (d . mu_j) d - w1 mu_j; lanes stride the entries, fixed-order
// butterfly reduction
__global__ void __launch_bounds__(256) induced_list_apply_kernel(int n, int np, const unsigned long long* __restrict__ offs,
                                                                 const PairEntry* __restrict__ list,
                                                                 const double* __restrict__ mu, double* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (i >= np) return;
    double ex = 0.0, ey = 0.0, ez = 0.0;
    if (i < n) {
        const unsigned long long b = offs[(size_t)i * 4], e = offs[(size_t)i * 4 + 4];
        for (unsigned long long q = b + lane; q < e; q += 32) {
            const PairEntry p = list[q];
            const double mx = mu[p.j], my = mu[np + p.j], mz = mu[2 * (size_t)np + p.j];
            const double dm = (p.dx * mx + p.dy * my + p.dz * mz) * p.w2;
            ex += dm * p.dx - p.w1 * mx;
            ey += dm * p.dy - p.w1 * my;
            ez += dm * p.dz - p.w1 * mz;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        ex += __shfl_xor_sync(0xffffffffu, ex, o);
        ey += __shfl_xor_sync(0xffffffffu, ey, o);
        ez += __shfl_xor_sync(0xffffffffu, ez, o);
    }
    if (lane == 0) {
        out[i] = ex;
        out[np + i] = ey;
        out[2 * (size_t)np + i] = ez;
    }
}

}  // namespace

int launch_induced_real_clusters(const Cell& c, bool ortho, const SubsetDev& pol, const EwaldFullOpts& o, const double* cbox,
                                 const double* mu, int jsplits, double* part, cudaStream_t s) {
    const int nclus = pol.n_pad / 32;
    if (nclus <= 0) return 0;
    const double cull = c.cutoff + 1e-6;
    const double cull2 = ortho ? cull * cull : 0.0;
    const int batches = (nclus + 31) / 32;
    if (jsplits < 1) jsplits = 1;
    if (jsplits > batches) jsplits = batches;
    const int bps = (batches + jsplits - 1) / jsplits;
    jsplits = (batches + bps - 1) / bps;
    dim3 grid(nclus, jsplits);
    if (ortho)
        induced_real_cluster_kernel<true><<<grid, TILE, 0, s>>>(c, pol, o, box_len(c), cbox, cull2, mu, bps, part);
    else
        induced_real_cluster_kernel<false><<<grid, TILE, 0, s>>>(c, pol, o, box_len(c), cbox, cull2, mu, bps, part);
    return jsplits;
}

int launch_dipole_sf(const Cell& c, const SubsetDev& pol, const KVecDev& kv, const double* kvec, const double* mu,
                     int atoms_per_chunk, double* sf_partial, cudaStream_t s, const KGroupsDev* kg) {
    const int nchunks = (pol.n + atoms_per_chunk - 1) / atoms_per_chunk;
    if (nchunks <= 0) return 0;
    if (kg && kg->ngroups > 0 && kg->max_count <= SG_MAXD) {
        const int nl = kv.kmax + 1;
        const int gthreads = (kg->ngroups + 31) / 32 * 32;
        if (gthreads * SG_SUB <= 512) {
            const size_t tables = sizeof(double) * (2 * SG_ATOMS * 3 * nl + 3 * SG_ATOMS);
            const size_t stage = sizeof(double) * 2 * SG_MAXD * gthreads;
            const size_t smem = tables > stage ? tables : stage;
            static unsigned long long attr_set_g = 0;
            if (first_on_device(attr_set_g)) {
                cudaFuncSetAttribute(dipole_sf_grouped_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
            }
            dipole_sf_grouped_kernel<<<nchunks, gthreads * SG_SUB, smem, s>>>(c, pol, kv.nk, nl, kg->ngroups, gthreads, kg->groups, mu,
                                                                             atoms_per_chunk, sf_partial);
            return nchunks;
        }
    }
    const size_t smem = sizeof(double) * (6 * (kv.kmax + 1) * DS_ATOMS + 3 * DS_ATOMS);
    static unsigned long long attr_set = 0;
    if (first_on_device(attr_set)) {
        cudaFuncSetAttribute(dipole_sf_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    }
    dipole_sf_kernel<<<nchunks, DS_THREADS, smem, s>>>(c, pol, kv, kvec, mu, atoms_per_chunk, sf_partial);
    return nchunks;
}

void launch_total_mu(const SubsetDev& pol, const double* mu, double* out3, cudaStream_t s) {
    total_mu_kernel<<<1, 1024, 0, s>>>(pol.n, pol.n_pad, mu, out3);
}

void launch_induced_finish(const Cell& c, const SubsetDev& rows, const KVecDev& kv, const double* kvec, const double* sf,
                           const EwaldFullOpts& o, const double* part, int jsplits, const double* total_mu, const double* mu,
                           double* out, cudaStream_t s) {
    if (rows.n_pad <= 0) return;
    const size_t smem = sizeof(double) * (6 * (kv.kmax + 1) * FIN_TILE + 2 * (size_t)kv.nk) + sizeof(int) * kv.nk;
    static unsigned long long attr_set = 0;
    if (first_on_device(attr_set)) {
        cudaFuncSetAttribute(induced_finish_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    }
    const double scale = 8.0 * M_PI / c.volume;                            // :266
    const double surf = -4.0 * M_PI / (3.0 * c.volume);                    // :314
    const double selfc = 4.0 * o.a * o.a * o.a / (3.0 * kSqrtPiRef);       // :314
    induced_finish_kernel<<<(rows.n_pad + FIN_TILE - 1) / FIN_TILE, FIN_TILE, smem, s>>>(c, rows, kv, kvec, sf, scale, part, part ? jsplits : 0,
                                                               rows.n_pad, total_mu, surf, selfc, mu, out);
}

void launch_ewald_update(const SubsetDev& pol, const UpdateOpts& o, const double* es, const double* eind, double* mu,
                         double* mu_old, double* mu_new, double* rrms, unsigned long long* max_err2_bits, cudaStream_t s) {
    if (pol.n_pad <= 0) return;
    ewald_update_kernel<<<(pol.n_pad + 255) / 256, 256, 0, s>>>(pol.n, pol.n_pad, pol.alpha, o, es, eind, mu, mu_old, mu_new,
                                                              rrms, max_err2_bits);
}

void launch_ewald_palmo_finish(const SubsetDev& pol, const double* es, const double* eind, const double* mu_new, double* eic,
                               cudaStream_t s) {
    if (pol.n <= 0) return;
    ewald_palmo_finish_kernel<<<(pol.n + 255) / 256, 256, 0, s>>>(pol.n, pol.n_pad, pol.alpha, es, eind, mu_new, eic);
}


// ---- pair-list form of the real-space term (see PairEntry)
size_t induced_list_entry_bytes() { return sizeof(PairEntry); }

// counts per (atom, warp slot) and their exclusive scan: cnt [4 n_pad] uints, offs [4 n_pad + 1] u64 (offs[4 n_pad] = total)
void launch_induced_list_count(const Cell& c, bool ortho, const SubsetDev& pol, const EwaldFullOpts& o, const double* cbox,
                               unsigned int* cnt, unsigned long long* offs, cudaStream_t s) {
    const int nclus = pol.n_pad / 32;
    if (nclus <= 0) return;
    const double cull = c.cutoff + 1e-6;
    const double cull2 = ortho ? cull * cull : 0.0;
    if (ortho)
        induced_list_kernel<true, false><<<nclus, TILE, 0, s>>>(c, pol, o, box_len(c), cbox, cull2, offs, cnt, nullptr);
    else
        induced_list_kernel<false, false><<<nclus, TILE, 0, s>>>(c, pol, o, box_len(c), cbox, cull2, offs, cnt, nullptr);
    list_scan_kernel<<<1, 1024, 0, s>>>(cnt, 4 * pol.n_pad, offs);
}

void launch_induced_list_fill(const Cell& c, bool ortho, const SubsetDev& pol, const EwaldFullOpts& o, const double* cbox,
                              const unsigned long long* offs, void* list, cudaStream_t s) {
    const int nclus = pol.n_pad / 32;
    if (nclus <= 0) return;
    const double cull = c.cutoff + 1e-6;
    const double cull2 = ortho ? cull * cull : 0.0;
    PairEntry* l = reinterpret_cast<PairEntry*>(list);
    if (ortho)
        induced_list_kernel<true, true><<<nclus, TILE, 0, s>>>(c, pol, o, box_len(c), cbox, cull2, offs, nullptr, l);
    else
        induced_list_kernel<false, true><<<nclus, TILE, 0, s>>>(c, pol, o, box_len(c), cbox, cull2, offs, nullptr, l);
}

void launch_induced_list_apply(const SubsetDev& pol, const unsigned long long* offs, const void* list, const double* mu,
                               double* part, cudaStream_t s) {
    if (pol.n_pad <= 0) return;
    const long long threads = (long long)pol.n_pad * 32;
    induced_list_apply_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, s>>>(pol.n, pol.n_pad, offs,
                                                                              reinterpret_cast<const PairEntry*>(list), mu, part);
}

}  // namespace mpmc
