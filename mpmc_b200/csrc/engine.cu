// engine.cu -- host side of libmpmc_b200: device-resident system image, evaluation pipeline and
// the C ABI of include/mpmc_b200.h.
//
// The pipeline mirrors energy() (reference src/energy/energy.c:67-226):
//   pairs()+lj()+coulombic_real()  -> K1 pair tiles            (pair_kernel.cu)
//   coulombic_reciprocal()         -> K2 structure factor      (recip_kernel.cu)
//   coulombic_self(), LRC          -> O(N) / O(types^2) on the host, identical formulas
//   polar(): thole_field()         -> K3 static field          (field_kernel.cu)
//            thole_iterative()     -> K4 contraction + K5 update, or the blocked GS sweep
//                                                              (dipole_kernel.cu)
// The O(N^2) pair_t lists and the 3N x 3N A matrix are never built: device memory is O(N).
#include <algorithm>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include <cstdio>
#include <cstring>
#include <map>
#include <numeric>
#include <string>
#include <utility>
#include <vector>

#include "kernels.h"

using namespace mpmc;

static thread_local std::string g_err = "";
// for entry points that live in other translation units (recip_static_kernel.cu)
namespace mpmc {
void set_last_error(const char* msg) { g_err = msg ? msg : ""; }
}  // namespace mpmc

#define CU(call)                                                                                      \
    do {                                                                                              \
        cudaError_t e__ = (call);                                                                     \
        if (e__ != cudaSuccess) {                                                                     \
            char b__[512];                                                                            \
            snprintf(b__, sizeof(b__), "CUDA error %s at %s:%d (%s)", cudaGetErrorString(e__), __FILE__, __LINE__, #call); \
            g_err = b__;                                                                              \
            return 10;                                                                                \
        }                                                                                             \
    } while (0)

#define FAIL(code, msg) \
    do {                \
        g_err = (msg);  \
        return (code);  \
    } while (0)

namespace {

template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t cap = 0;
    int ensure(size_t n) {
        if (n <= cap) return 0;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        size_t want = n + n / 4 + 256;
        if (cudaMalloc(&p, want * sizeof(T)) != cudaSuccess) return 1;
        cap = want;
        return 0;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

struct Subset {
    std::vector<int> idx;  // atom indices, ascending
    int n = 0, n_pad = 0;
    DevBuf<double> x, y, z, q, alpha, aux;
    DevBuf<double> eps, sig, mm, cbox;  // only the sorted image of all atoms (cutoff kernels)
    DevBuf<int> mol, flags, didx;
    SubsetDev dev() const {
        SubsetDev d;
        d.n = n;
        d.n_pad = n_pad;
        d.x = x.p;
        d.y = y.p;
        d.z = z.p;
        d.q = q.p;
        d.alpha = alpha.p;
        d.aux = aux.p;
        d.mol = mol.p;
        d.flags = flags.p;
        d.idx = didx.p;
        d.eps = eps.p;
        d.sig = sig.p;
        d.molmass = mm.p;
        d.cbox = cbox.p;
        return d;
    }
    void release() {
        x.release(), y.release(), z.release(), q.release(), alpha.release(), aux.release();
        mol.release(), flags.release(), didx.release();
        eps.release(), sig.release(), mm.release(), cbox.release();
    }
};

enum Phase { PH_PAIR = 0, PH_RECIP, PH_FIELD, PH_DIPOLE, PH_TOTAL, PH_COUNT };

// device scalar slots
enum Slot {
    S_PAIR = 0,    // 8 doubles: rd, es, intra, min_r, n_tested, n_lj, n_es
    S_RECIP = 8,   // 1
    S_POLAR = 9,   // 3: sum mu.Es, sum mu.eic, sum rrms
    S_ERR = 12,    // max err^2 bits (as unsigned long long)
    S_FIELDCNT = 13,
    S_TOTMU = 16,  // 3: sum of the dipoles (full-Ewald polarization)
    S_PEER_TIMEOUT = 20,  // set to 1.0 by a peer-wait kernel whose peer never signalled
    S_GS_TIMEOUT = 21,
    S_FUSED_OUT = 26,     // 2: iterations done / iter_success of the fused small-system solve (decided on the device)
    S_FUSED_ERR = 28,     // 2 words: its residual, alternating by iteration parity    // set to 1.0 by the persistent Gauss-Seidel sweep when one of its CTAs waited 2 s for another
    S_CRYSTAL = 22,       // 2: rd_crystal image sums of the pairs, pairs counted
    S_COUNT = 24
};

}  // namespace

// One host thread per extra rank of a device group: the ranks' launches are issued side by side (one thread issuing
// ~60 launches per evaluation to each of 8 devices in turn put 2.4 ms of launch latency on an 6 ms evaluation).
struct RankWorker {
    std::thread th;
    std::mutex m;
    std::condition_variable cv;
    std::function<void()> job;
    bool has_job = false, done = false, quit = false;
    void loop() {
        std::unique_lock<std::mutex> lk(m);
        for (;;) {
            cv.wait(lk, [&] { return has_job || quit; });
            if (quit) return;
            lk.unlock();
            job();
            lk.lock();
            has_job = false;
            done = true;
            cv.notify_all();
        }
    }
    void submit(std::function<void()> f) {
        std::lock_guard<std::mutex> lk(m);
        job = std::move(f);
        has_job = true;
        done = false;
        cv.notify_all();
    }
    void wait() {
        std::unique_lock<std::mutex> lk(m);
        cv.wait(lk, [&] { return done; });
    }
    void stop() {
        {
            std::lock_guard<std::mutex> lk(m);
            quit = true;
            cv.notify_all();
        }
        if (th.joinable()) th.join();
    }
};

struct mpmc_b200_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    mpmc_b200_params P{};
    bool have_params = false, have_atoms = false;
    int rank = 0, world = 1;

    // host copy (list order)
    int n = 0, n_pad = 0;
    std::vector<double> hx, hy, hz, hq, halpha, heps, hsig, hmm;
    std::vector<int> hmol, hfrozen;

    // device image
    DevBuf<double> x, y, z, q, eps, sig, molmass, alpha;
    DevBuf<int> mol, flags;
    Subset charged, recip, pol, gs;  // gs: polarizable atoms permuted into Gauss-Seidel order
    Subset nonpol;                   // alpha == 0 atoms: rows of the palmo field that are only written back
    DevBuf<double> eic_np_part;      // [source slices][3][nonpol.n_pad] partial rows of the same
    DevBuf<double> eic_np;           // [3][nonpol.n_pad]
    bool eic_np_valid = false;
    DevBuf<double> eind_np;          // [3][nonpol.n_pad]: ef_induced of the alpha == 0 atoms (full-Ewald polarization only)
    // full-Ewald polarization: pair list of the real-space term, rebuilt by every energy() (ewald_full_kernel.cu PairEntry)
    DevBuf<unsigned int> k7_cnt;
    DevBuf<unsigned long long> k7_offs;
    DevBuf<char> k7_list;
    bool k7_list_on = false;
    bool k7_grouped_sf = true;       // MPMC_B200_K7_SF=plain: one thread per k-vector
    double k7_list_cap_gb = 16.0;    // MPMC_B200_K7_LIST_GB; 0 = always the on-the-fly kernel
    bool eind_np_valid = false;
    DevBuf<int2> tiles;
    int ntiles = 0;
    DevBuf<double> pair_part;
    int pair_stride = 0;
    // rd_crystal (lj.c:175-216): all tile pairs of the list-ordered atoms, image vectors, partial sums
    DevBuf<int2> crystal_tiles;
    int crystal_ntiles = 0, crystal_stride = 0, crystal_nimg = 0;
    DevBuf<double> crystal_part, crystal_img;
    std::vector<double> h_crystal_img;
    bool crystal_img_dirty = true;
    double crystal_self = 0;  // rd_crystal_self() summed over the atoms (host, compute_analytic)

    // reciprocal space
    DevBuf<int> kl;
    DevBuf<int4> kgroups;  // (l0, l1) runs of the k list, see build_kvectors
    int nkgroups = 0, kgroup_max = 0;
    DevBuf<double> kweight, kweight_field, kvec, sf_partial, sf;
    int nk = 0;
    bool kdirty = true;
    int recip_chunks = 0, recip_apc = 0;

    // polarization
    DevBuf<double> ef_static;  // [3][n_pad]
    DevBuf<double> epart;      // max(field, dipole) partials
    DevBuf<double> es, mu, mu_old, mu_new, eind, eic, rrms, rank_packed, rmin_part, field_cnt;
    DevBuf<double> g_es, g_mu, g_mu_old, g_mu_new, g_eind, g_eic, g_rrms, g_delta, g_coef;  // GS space
    bool gs_delta = true;  // delta form of the sweep (symmetric Jacobi part + block pushes); MPMC_B200_GS=panels: panel form
    bool gs_persistent = true;  // one cooperative launch per sweep (resolver + pushers); MPMC_B200_GS=blocks: two launches per block
    DevBuf<double> g_tens, g_pushacc, g_ctab;
    // Gauss-Seidel: the Jacobi part of a sweep (and the Palmo contraction) does not care about the processing order, so it
    // runs on a spatially ordered image of the polarizable atoms -- tight cluster boxes, the far-cluster damping skip --
    // and its result is gathered back into processing order (4.6 -> 3.7 ms per contraction at 63k atoms)
    Subset gsp;
    std::vector<int> gsp_packed;        // packed (list-order) index of the s-th atom of the spatial order
    DevBuf<int> map_s2g, map_g2s;       // spatial index -> processing index and back (rebuilt with the processing order)
    DevBuf<double> g_mu_sp, g_esum_sp, cbox_sp;
    bool gs_spatial = true;             // MPMC_B200_GS_JACOBI=order: the Jacobi part in processing order (A/B)
    DevBuf<unsigned> gs_flags;
    unsigned gs_epoch = 0;
    DevBuf<unsigned long long> gs_trace;  // MPMC_B200_GS_TRACE=1: resolver time stamps of the last sweep, summarised on stderr
    DevBuf<int> perm;
    std::vector<int> hperm;
    DevBuf<double> wolf_table;
    int wolf_table_max = 0;
    double wolf_table_alpha = -1.0, wolf_table_cut = -1.0;
    int field_jsplits = 1, dipole_jsplits = 1;
    // symmetric K4 (dipole_sym_kernel.cu): cluster boxes, I-side / J-side partial rows, reduced field
    bool use_sym = true;
    bool pol_sorted = false, resort_pending = false;
    // cluster form of the cutoff kernels K1 / K3: all atoms in spatial order
    bool use_clusters = true;
    Subset sorted;
    DevBuf<int> follow;        // list order: later atoms of the same, not entirely frozen, molecule
    DevBuf<double> intra_part;
    int pair_kslices = 1;
    bool boxes_fresh = false;  // sorted.cbox matches the current positions
    double pairs_tested_analytic = 0.0;
    std::vector<double> hx_sort, hy_sort, hz_sort;  // positions at the time of the last spatial sort
    // uVT: an insertion or a removal of a molecule changes the atom set by a few atoms.  The spatial orders (a matter of
    // speed, never of results) are then carried over instead of being rebuilt: surviving atoms keep their place, new atoms
    // go to the end; a full sort again after enough edits or when an atom has drifted (resort_pending)
    std::vector<int> edit_map;     // set_atoms -> repack: old atom index -> new one, -1 = gone; empty = no carry-over
    std::vector<int> edit_fresh;   // the atoms to place: new ones (set_atoms) or drifted ones (refresh_subsets), ascending
    // atoms that moved further than the drift limit from where they were when the orders were made (set_positions, the end
    // of a resident batch): refresh_subsets() takes them out of the orders and puts them back behind their nearest
    // neighbour instead of sorting everything again -- with the reference's default move_factor of 1.0 nearly every trial
    // move is such a jump (a full sort per step: 19 ms at 63k atoms, 25 -> 17 MC steps/s)
    std::vector<int> drifted;
    bool drift_tracked = false;    // `drifted` names every atom that has to be relocated
    int order_edits = 0;           // atoms added + removed since the last full sort
    bool orders_valid = false;     // pol / sorted / gsp hold spatial orders made under the current options
    DevBuf<double> cbox, sym_partI, sym_partJ, esum;
    int sym_ysplits = 1;
    int fb = 0, fe = 0, pb = 0, pe = 0;  // this shard's static-field rows (atom space) and dipole rows (packed space)

    DevBuf<double> scalars;
    double* h_scalars = nullptr;  // pinned
    double* h_stage = nullptr;    // pinned staging for positions
    size_t h_stage_cap = 0;
    DevBuf<double> wb;            // write-back image (writeback_kernel), 14 n doubles
    double* h_wb = nullptr;       // pinned
    size_t h_wb_cap = 0;

    // warm-started dipole solve (opt-in: MPMC_B200_WARM_START=1 / mpmc_b200_set_warm_start): in tolerance
    // mode the iteration starts from the previous evaluation's converged dipoles instead of gamma alpha E_s
    bool warm_enabled = false, warm_valid = false;
    DevBuf<double> mu_keep;  // [3][n_pad], atom space (survives re-sorting of the polarizable subset)

    // incremental evaluation after small moves (opt-in: MPMC_B200_INCREMENTAL=1 / mpmc_b200_set_incremental)
    bool inc_enabled = false;
    bool inc_valid = false;          // inc_base / sf describe the positions the device held at the last evaluation
    int inc_since_full = 0;          // incremental evaluations since the last full one (rounding drift bound)
    mpmc_b200_energies inc_base;     // energies of the last evaluation
    mpmc_b200_counters inc_cnt;      // its counters
    std::vector<MovedRec> inc_moved; // atoms that moved since then, with their positions then and now
    std::vector<int> inc_slot;       // per atom: index in inc_moved + 1, 0 = not moved
    DevBuf<MovedRec> inc_moved_dev;
    DevBuf<double> inc_part, inc_out;

    // host-side analytic terms
    double lrc_pair = 0, lrc_self = 0, es_self = 0;
    bool analytic_dirty = true;
    bool subsets_stale = false;

    // peer-memory exchange of the induced-field partial sums (sharded runs; see PeerSignal / PeerWait)
    // One exported allocation per rank (layout: peer_* helpers below): the induced-field partial sums of the symmetric
    // contraction (double buffered), this rank's static-field rows, its device scalars (double buffered) and one flag
    // array per exchange channel.  Peers map it through CUDA IPC (one process per GPU) or directly (one process
    // driving several GPUs: mpmc_b200_create_group).
    unsigned char* peer_local = nullptr;
    size_t peer_cap = 0, peer_cap_atoms = 0;  // padded polarizable / atom rows the allocation was sized for
    void* peer_base[PEER_MAX] = {nullptr};    // every rank's allocation as seen from this process
    bool peer_ready = false, peer_ipc = false;
    unsigned long long peer_seq[3] = {0, 0, 0};  // per channel: induced field, static field, scalars
    bool field_via_peer = false;    // polar_field() leaves this rank's rows in the exported buffer (sharded_eval)
    bool scalars_complete = false;  // the device scalars of the pending evaluation are already summed over the ranks
    std::vector<mpmc_b200_ctx*> group;  // leader only (mpmc_b200_create_group): every rank's context, this one first
    bool group_member = false;          // owned by a leader: destroyed with it
    std::vector<RankWorker*> workers;   // leader only: host threads that issue ranks 1.. (MPMC_B200_GROUP_THREADS=0: none)
    bool layout_dirty = true;           // buffers may still be (re)allocated by the next evaluation: see sharded_eval
    cudaStream_t side = nullptr;  // K1 / K2 overlap with the polarization solve
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    bool side_active = false;
    cudaEvent_t ev[PH_COUNT][2];
    bool ev_used[PH_COUNT] = {false, false, false, false, false};
    float phase_ms[PH_COUNT] = {0, 0, 0, 0, 0};

    // per-launch timing of the dominant kernel (K4 contraction): event pairs, created on demand
    std::vector<cudaEvent_t> kev;
    int kev_used = 0;
    float contract_ms_sum = 0.f;
    int contract_launches = 0;

    // resident MC steps (mpmc_b200_mc_steps): the batch's moves and per-step records on the device, and a
    // ring of per-step write-back images (14 n doubles each) with its pinned host copy
    DevBuf<mpmc_b200_mc_move> mc_moves;
    DevBuf<McStepDev> mc_res;
    DevBuf<double> mc_ecur, mc_ring;
    DevBuf<int> mc_ctr;             // step counter on the device (read by the replayed graph of a step)
    int mc_graph = -1;              // MPMC_B200_MC_GRAPH: 1 = steps 2.. of a batch replay a CUDA graph captured from step 1
    McStepDev* h_mc_res = nullptr;  // pinned
    double* h_mc_ring = nullptr;    // pinned
    size_t h_mc_res_cap = 0, h_mc_ring_cap = 0;
    int mc_last_steps = 0;
    bool mc_ring_valid = false;

    mpmc_b200_counters cnt{};
    int launches = 0;
    int pending_iterations = 0, pending_iter_success = 0;
    bool iter_from_device = false;  // the fused solve counted its iterations itself: read them with the scalars
    bool pending = false;
    bool last_had_polar = false;
};

typedef mpmc_b200_ctx Ctx;

// ------------------------------------------------------------------------------------------------
// small device helpers defined here
// ------------------------------------------------------------------------------------------------
__global__ void gather_positions_kernel(const double* __restrict__ x, const double* __restrict__ y,
                                        const double* __restrict__ z, const int* __restrict__ idx, int n,
                                        double* __restrict__ sx, double* __restrict__ sy, double* __restrict__ sz) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const int g = idx[k];
    sx[k] = x[g];
    sy[k] = y[g];
    sz[k] = z[g];
}

// A subset's own SoA copy, filled on the device from the list-order atom arrays and the subset's index list (one
// upload per subset instead of nine to twelve): entries k < n take atom idx[k]; padding takes zeros, molecule -1, no flags
// (aux 1.0 under linear damping, where the kernels form its powers).
__global__ void subset_gather_kernel(int n, int n_pad, const int* __restrict__ idx, const double* __restrict__ ax,
                                     const double* __restrict__ ay, const double* __restrict__ az,
                                     const double* __restrict__ aq, const double* __restrict__ aalpha,
                                     const double* __restrict__ aeps, const double* __restrict__ asig,
                                     const double* __restrict__ amm, const int* __restrict__ amol,
                                     const int* __restrict__ aflags, int with_aux, int damp_linear, int with_lj,
                                     double* __restrict__ x, double* __restrict__ y, double* __restrict__ z,
                                     double* __restrict__ q, double* __restrict__ alpha, double* __restrict__ aux,
                                     double* __restrict__ eps, double* __restrict__ sig, double* __restrict__ mm,
                                     int* __restrict__ mol, int* __restrict__ flags) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_pad) return;
    if (k < n) {
        const int g = idx[k];
        const double qq = aq[g], al = aalpha[g];
        x[k] = ax[g];
        y[k] = ay[g];
        z[k] = az[g];
        q[k] = qq;
        alpha[k] = al;
        aux[k] = with_aux ? (damp_linear ? pow(al, 1.0 / 6.0) : qq) : 0.0;
        mol[k] = amol[g];
        flags[k] = aflags[g];
        if (with_lj) {
            eps[k] = aeps[g];
            sig[k] = asig[g];
            mm[k] = amm[g];
        }
    } else {
        x[k] = y[k] = z[k] = q[k] = alpha[k] = 0.0;
        aux[k] = (with_aux && damp_linear) ? 1.0 : 0.0;
        mol[k] = -1;
        flags[k] = 0;
        if (with_lj) eps[k] = sig[k] = mm[k] = 0.0;
    }
}

// dst[c][k] = src[c][perm[k]] (forward) or dst[c][perm[k]] = src[c][k] (inverse), k < n
__global__ void permute_kernel(const double* __restrict__ src, double* __restrict__ dst, const int* __restrict__ perm, int n,
                               int np, int ncomp, int inverse) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const int pk = perm[k];
    for (int c = 0; c < ncomp; c++) {
        if (inverse)
            dst[c * np + pk] = src[c * np + k];
        else
            dst[c * np + k] = src[c * np + pk];
    }
}
__global__ void permute_int_kernel(const int* __restrict__ src, int* __restrict__ dst, const int* __restrict__ perm, int n) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) dst[k] = src[perm[k]];
}

// GS convergence / rrms on (mu_new - mu_old), same arithmetic as dipole_update_kernel
__global__ void gs_check_kernel(int n, int np, const double* __restrict__ mu_new, const double* __restrict__ mu_old,
                                int do_rrms, double* __restrict__ rrms, unsigned long long* __restrict__ max_err2_bits) {
    __shared__ double red[32];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    double worst = 0.0;
    if (i < n) {
        double num = 0.0, den = 0.0;
        for (int c = 0; c < 3; c++) {
            const double nw = mu_new[c * np + i], d = nw - mu_old[c * np + i];
            worst = fmax(worst, d * d);
            num += d * d;
            den += nw * nw;
        }
        if (do_rrms) {
            double v = sqrt(num / den);
            if (!isfinite(v)) v = 0.0;
            rrms[i] = v;
        }
    }
    worst = block_reduce<2>(worst, red);
    if (threadIdx.x == 0) atomicMax(max_err2_bits, (unsigned long long)__double_as_longlong(worst));
}

__global__ void __launch_bounds__(1024) sum_reduce_kernel(const double* __restrict__ in, int n, double* __restrict__ out) {
    __shared__ double red[32];
    double acc = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) acc += in[i];
    acc = block_reduce<0>(acc, red);
    if (threadIdx.x == 0) out[0] = acc;
}

// failure path of thole_iterative (src/polarization/thole_iterative.c:199-210): mu = alpha E_s, eic = 0
__global__ void reset_dipoles_kernel(int n, int np, const double* __restrict__ alpha, const double* __restrict__ es,
                                     double* __restrict__ mu, double* __restrict__ eic) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    for (int c = 0; c < 3; c++) {
        mu[c * np + i] = alpha[i] * es[c * np + i];
        eic[c * np + i] = 0.0;
    }
}

// What polar() leaves in atom_t, assembled on the device in the caller's layout (list order, [n][3]):
// out = mu | ef_static | ef_induced | ef_induced_change | dipole_rrms | rank_metric  (14 n doubles, zero filled
// beforehand).  Packed rows [own_b, own_e) carry the row-owned arrays; mu (and everything when own_b..own_e is
// the whole range) is complete.  One D2H copy of the result replaces six staged ones + host scatter loops.
__global__ void writeback_kernel(int n, int np, const double* __restrict__ ef_static, int n_pol, int pp,
                                 const int* __restrict__ pol_idx, const double* __restrict__ mu,
                                 const double* __restrict__ eind, const double* __restrict__ eic,
                                 const double* __restrict__ rrms, const double* __restrict__ rank_metric, int own_b,
                                 int own_e, int n_np, int qp, const int* __restrict__ np_idx,
                                 const double* __restrict__ eind_np, const double* __restrict__ eic_np,
                                 double* __restrict__ out) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    double* o_mu = out;
    double* o_es = out + 3 * (size_t)n;
    double* o_ei = out + 6 * (size_t)n;
    double* o_eic = out + 9 * (size_t)n;
    double* o_rr = out + 12 * (size_t)n;
    double* o_rk = out + 13 * (size_t)n;
    if (t < n)
        for (int d = 0; d < 3; d++) o_es[3 * (size_t)t + d] = ef_static[(size_t)d * np + t];
    if (t < n_pol) {
        const size_t g = pol_idx[t];
        const bool own = t >= own_b && t < own_e;
        for (int d = 0; d < 3; d++) {
            o_mu[3 * g + d] = mu[(size_t)d * pp + t];
            if (own) {
                o_ei[3 * g + d] = eind[(size_t)d * pp + t];
                o_eic[3 * g + d] = eic[(size_t)d * pp + t];
            }
        }
        if (own) {
            o_rr[g] = rrms[t];
            o_rk[g] = rank_metric[t];
        }
    }
    if (t < n_np && (eind_np || eic_np)) {
        const size_t g = np_idx[t];
        for (int d = 0; d < 3; d++) {
            if (eind_np) o_ei[3 * g + d] = eind_np[(size_t)d * qp + t];
            if (eic_np) o_eic[3 * g + d] = eic_np[(size_t)d * qp + t];
        }
    }
}

// ------------------------------------------------------------------------------------------------
// host helpers
// ------------------------------------------------------------------------------------------------
static int round_up(int n, int m) { return ((n + m - 1) / m) * m; }

static bool is_ortho(const mpmc_b200_params& P) {
    for (int a = 0; a < 3; a++)
        for (int b = 0; b < 3; b++)
            if (a != b && (P.basis[a][b] != 0.0 || P.reciprocal_basis[a][b] != 0.0)) return false;
    return true;
}

static Cell make_cell(const mpmc_b200_params& P) {
    Cell c;
    for (int a = 0; a < 3; a++)
        for (int b = 0; b < 3; b++) {
            c.B[a][b] = P.basis[a][b];
            c.R[a][b] = P.reciprocal_basis[a][b];
        }
    c.cutoff = P.cutoff;
    c.volume = P.volume;
    return c;
}

template <typename T>
static int upload(DevBuf<T>& b, const std::vector<T>& h, cudaStream_t s) {
    if (b.ensure(h.size() ? h.size() : 1)) FAIL(11, "cudaMalloc failed");
    if (!h.empty()) CU(cudaMemcpyAsync(b.p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, s));
    return 0;
}

// mixed LJ parameters on the host for the analytic long-range correction (pairs.c:84-211)
static void host_mix(int mixing, double ei, double si, double ej, double sj, double& eps, double& sigma) {
    eps = 0.0;
    sigma = 0.0;
    switch (mixing) {
        case MPMC_MIX_WALDMAN_HAGLER: {
            double si3 = si * si * si, si6 = si3 * si3, sj3 = sj * sj * sj, sj6 = sj3 * sj3;
            if (si < 0.0 || sj < 0.0)
                sigma = pow(0.5 * (si6 + sj6), 1. / 6.);
            else if (si == 0 || sj == 0)
                eps = sqrt(ei * ej);
            else {
                sigma = pow(0.5 * (si6 + sj6), 1. / 6.);
                eps = sqrt(ei * ej) * 2.0 * si3 * sj3 / (si6 + sj6);
            }
        } break;
        case MPMC_MIX_HALGREN:
            if (si > 0.0 && sj > 0.0) sigma = (si * si * si + sj * sj * sj) / (si * si + sj * sj);
            if (ei > 0.0 && ej > 0.0) eps = 4 * ei * ej / pow(sqrt(ei) + sqrt(ej), 2);
            break;
        case MPMC_MIX_C6:
            sigma = 0.5 * (si + sj);
            if (sigma != 0.0) eps = 64.0 * sqrt(ei * ej) * pow(si, 3.0) * pow(sj, 3.0) / pow(si + sj, 6.0);
            break;
        default:
            if (si < 0.0 || sj < 0.0)
                sigma = 0.5 * (fabs(si) + fabs(sj));
            else if (si == 0 || sj == 0)
                eps = sqrt(ei * ej);
            else {
                sigma = 0.5 * (si + sj);
                eps = sqrt(ei * ej);
            }
    }
}

// lj(): src/energy/lj.c:175-178 -- with rd_crystal the LJ cutoff (pair test, image sums, LRC) is wider
static double rd_cutoff(const mpmc_b200_params& P) {
    return P.rd_crystal ? 2.0 * P.cutoff * ((double)P.rd_crystal_order - 0.5) : P.cutoff;
}

static double lrc_value(double eps, double sigma, double cutoff, double volume) {
    // lj_lrc_corr / lj_lrc_self (src/energy/lj.c:69-80, 92-103)
    double sig_cut = fabs(sigma) / cutoff;
    double sig3 = fabs(sigma);
    sig3 *= sig3 * sig3;
    double sig_cut3 = sig_cut * sig_cut * sig_cut;
    double sig_cut9 = sig_cut3 * sig_cut3 * sig_cut3;
    return ((16.0 / 3.0) * M_PI * eps * sig3) * ((1.0 / 3.0) * sig_cut9 - sig_cut3) / volume;
}

// LRC and Ewald self term.  The reference adds one LRC value per unordered pair that is not
// frozen-frozen and has non-zero mixed parameters -- intramolecular and beyond-cutoff pairs
// included (lj.c:62-65,188,256).  The value depends only on the two (eps, sigma) types, so the sum is
// pair-count(type a, type b) * value(a, b): O(N) + O(types^2) instead of O(N^2).
static void compute_analytic(Ctx* c) {
    const mpmc_b200_params& P = c->P;
    c->lrc_pair = c->lrc_self = c->es_self = 0.0;
    if (P.rd_lrc) {
        struct TypeCount {
            double eps, sig;
            double total = 0, frozen = 0;
        };
        std::map<std::pair<double, double>, int> index;
        std::vector<TypeCount> types;
        for (int i = 0; i < c->n; i++) {
            auto key = std::make_pair(c->heps[i], c->hsig[i]);
            auto it = index.find(key);
            int t;
            if (it == index.end()) {
                t = (int)types.size();
                index[key] = t;
                TypeCount tc;
                tc.eps = c->heps[i];
                tc.sig = c->hsig[i];
                types.push_back(tc);
            } else
                t = it->second;
            types[t].total += 1;
            if (c->hfrozen[i]) types[t].frozen += 1;
        }
        double pair_sum = 0.0, self_sum = 0.0;
        for (size_t a = 0; a < types.size(); a++) {
            for (size_t b = a; b < types.size(); b++) {
                double eps, sigma;
                host_mix(P.mixing, types[a].eps, types[a].sig, types[b].eps, types[b].sig, eps, sigma);
                if (!(eps != 0 && sigma != 0)) continue;
                double npairs;
                if (a == b)
                    npairs = 0.5 * types[a].total * (types[a].total - 1) - 0.5 * types[a].frozen * (types[a].frozen - 1);
                else
                    npairs = types[a].total * types[b].total - types[a].frozen * types[b].frozen;
                if (npairs > 0) pair_sum += npairs * lrc_value(eps, sigma, rd_cutoff(P), P.volume);
            }
            if (types[a].sig != 0 && types[a].eps != 0) {
                double mobile = types[a].total - types[a].frozen;
                if (mobile > 0) self_sum += mobile * lrc_value(types[a].eps, types[a].sig, rd_cutoff(P), P.volume);
            }
        }
        c->lrc_pair = pair_sum;
        c->lrc_self = self_sum;
    }
    c->crystal_self = 0.0;
    c->h_crystal_img.clear();
    if (P.rd_crystal) {
        // image vectors a[p] = sum_q basis[q][p] * l[q] in the loop order of lj.c:199-211
        const int m = P.rd_crystal_order - 1;
        const double cut = rd_cutoff(P);
        for (int l0 = -m; l0 <= m; l0++)
            for (int l1 = -m; l1 <= m; l1++)
                for (int l2 = -m; l2 <= m; l2++) {
                    const int l[3] = {l0, l1, l2};
                    for (int p = 0; p < 3; p++) {
                        double a = 0;
                        for (int q = 0; q < 3; q++) a += P.basis[q][p] * l[q];
                        c->h_crystal_img.push_back(a);
                    }
                }
        c->crystal_nimg = (int)c->h_crystal_img.size() / 3;
        c->crystal_img_dirty = true;
        // rd_crystal_self() (lj.c:109-162; spectre / polarvdw / cdvdw off): every atom -- frozen ones included -- with
        // its own images, half weight; added per atom in list order as lj.c:264-267 does
        const int home = (c->crystal_nimg - 1) / 2;
        for (int i = 0; i < c->n; i++) {
            const double sig = c->hsig[i], eps = c->heps[i];
            if (sig == 0 && eps == 0) continue;
            double s6 = 0, s12 = 0;
            for (int k = 0; k < c->crystal_nimg; k++) {
                if (k == home) continue;
                const double* a = &c->h_crystal_img[3 * k];
                const double r = sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]);
                if (r > cut) continue;
                const double sor = fabs(sig) / r;
                s6 += 0.5 * pow(sor, 6);
                s12 += 0.5 * pow(sor, 12);
            }
            c->crystal_self += 4.0 * eps * ((sig < 0.0 ? 0.0 : s12) - s6);
        }
    }
    if (!P.rd_only) {  // coulombic_self (src/energy/coulombic.c:97-112), same accumulation order
        double self_potential = 0.0;
        const double alpha = P.ewald_alpha;
        for (int i = 0; i < c->n; i++) {
            if (c->hfrozen[i]) continue;
            self_potential -= alpha * c->hq[i] * c->hq[i] / sqrt(M_PI);
        }
        c->es_self = self_potential;
    }
    c->analytic_dirty = false;
}

// hemisphere of reciprocal vectors, reference enumeration order (coulombic.c:56-60)
static int build_kvectors(Ctx* c) {
    const mpmc_b200_params& P = c->P;
    const int kmax = P.ewald_kmax;
    if (kmax + 1 > 16) FAIL(3, "ewald_kmax > 15 not supported");
    std::vector<int> l0, l1, l2;
    for (int a = 0; a <= kmax; a++)
        for (int b = (!a ? 0 : -kmax); b <= kmax; b++)
            for (int d = ((!a && !b) ? 1 : -kmax); d <= kmax; d++) {
                if (a * a + b * b + d * d > kmax * kmax) continue;
                l0.push_back(a);
                l1.push_back(b);
                l2.push_back(d);
            }
    const int nk = (int)l0.size();
    if (nk > 2048) FAIL(3, "too many reciprocal vectors (nk > 2048)");
    std::vector<int> l(3 * (size_t)std::max(nk, 1));
    std::vector<double> w(std::max(nk, 1)), wf(std::max(nk, 1)), kv(3 * (size_t)std::max(nk, 1));
    for (int ik = 0; ik < nk; ik++) {
        int ll[3] = {l0[ik], l1[ik], l2[ik]};
        double k[3];
        for (int p = 0; p < 3; p++) {
            k[p] = 0;
            for (int q = 0; q < 3; q++) k[p] += 2.0 * M_PI * P.reciprocal_basis[p][q] * ll[q];
        }
        const double k2 = k[0] * k[0] + k[1] * k[1] + k[2] * k[2];
        w[ik] = exp(-k2 / (4.0 * P.ewald_alpha * P.ewald_alpha)) / k2;                  // coulombic.c:86
        wf[ik] = exp(-k2 / (4.0 * P.polar_ewald_alpha * P.polar_ewald_alpha)) / k2;     // polar_ewald.c:101-103
        for (int p = 0; p < 3; p++) {
            l[p * nk + ik] = ll[p];
            kv[p * nk + ik] = k[p];
        }
    }
    // (l0, l1) groups of the list (runs of consecutive l2), largest first: dipole_sf_grouped_kernel
    {
        std::vector<int4> groups;
        int maxc = 0;
        for (int ik = 0; ik < nk;) {
            int e = ik + 1;
            while (e < nk && l0[e] == l0[ik] && l1[e] == l1[ik] && l2[e] == l2[e - 1] + 1) e++;
            groups.push_back(make_int4(ik, l0[ik], l1[ik], (l2[ik] + 64) | ((e - ik) << 8)));
            maxc = std::max(maxc, e - ik);
            ik = e;
        }
        std::stable_sort(groups.begin(), groups.end(), [](const int4& a, const int4& b) { return (a.w >> 8) > (b.w >> 8); });
        c->nkgroups = (int)groups.size();
        c->kgroup_max = maxc;
        if (!groups.empty() && upload(c->kgroups, groups, c->stream)) return 11;
    }
    c->nk = nk;
    if (upload(c->kl, l, c->stream) || upload(c->kweight, w, c->stream) || upload(c->kweight_field, wf, c->stream) ||
        upload(c->kvec, kv, c->stream))
        return 11;
    if (c->sf.ensure(2 * (size_t)std::max(nk, 1))) FAIL(11, "cudaMalloc failed");
    c->kdirty = false;
    return 0;
}

// Spatial order of a set of atoms for the symmetric K4: recursive bisection along the longest axis of
// the (wrapped) bounding box, split points on multiples of 32, so that every 32-atom cluster the kernel
// classifies by its bounding box is a compact brick.  Any order gives the same sums up to rounding (the
// reference's contract_dipoles adds j in list order, thole_iterative.c:38-46); the order is a pure function
// of the positions, so results stay reproducible run to run.
static void spatial_order(const Ctx* c, std::vector<int>& idx) {
    const int n = (int)idx.size();
    if (n <= 32) return;
    const mpmc_b200_params& P = c->P;
    // wrapped fractional coordinates scaled by the lattice vector lengths
    std::vector<double> w(3 * (size_t)n);
    double len[3];
    for (int q = 0; q < 3; q++) len[q] = sqrt(P.basis[q][0] * P.basis[q][0] + P.basis[q][1] * P.basis[q][1] + P.basis[q][2] * P.basis[q][2]);
    for (int k = 0; k < n; k++) {
        const int g = idx[k];
        const double r[3] = {c->hx[g], c->hy[g], c->hz[g]};
        for (int p = 0; p < 3; p++) {
            double f = P.reciprocal_basis[0][p] * r[0] + P.reciprocal_basis[1][p] * r[1] + P.reciprocal_basis[2][p] * r[2];
            f -= floor(f);
            w[3 * (size_t)k + p] = f * len[p];
        }
    }
    std::vector<int> ord(n);
    std::iota(ord.begin(), ord.end(), 0);
    struct Range { int b, e; };
    std::vector<Range> stack;
    stack.push_back({0, n});
    while (!stack.empty()) {
        const Range r = stack.back();
        stack.pop_back();
        const int sz = r.e - r.b;
        if (sz <= 32) {
            std::sort(ord.begin() + r.b, ord.begin() + r.e);  // fixed order inside a cluster: by atom index
            continue;
        }
        double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
        for (int k = r.b; k < r.e; k++)
            for (int p = 0; p < 3; p++) {
                const double v = w[3 * (size_t)ord[k] + p];
                lo[p] = std::min(lo[p], v);
                hi[p] = std::max(hi[p], v);
            }
        int ax = 0;
        if (hi[1] - lo[1] > hi[ax] - lo[ax]) ax = 1;
        if (hi[2] - lo[2] > hi[ax] - lo[ax]) ax = 2;
        const int clusters = (sz + 31) / 32;
        const int left = (clusters / 2) * 32;
        const int mid = r.b + left;
        std::nth_element(ord.begin() + r.b, ord.begin() + mid, ord.begin() + r.e, [&](int u, int v) {
            const double a = w[3 * (size_t)u + ax], b = w[3 * (size_t)v + ax];
            return a < b || (a == b && u < v);
        });
        stack.push_back({mid, r.e});
        stack.push_back({r.b, mid});
    }
    std::vector<int> out(n);
    for (int k = 0; k < n; k++) out[k] = idx[ord[k]];
    idx.swap(out);
}

static int pack_subset(Ctx* c, Subset& s, const std::vector<int>& idx, int pad_multiple, bool with_aux, bool with_lj = false) {
    const mpmc_b200_params& P = c->P;
    s.idx = idx;
    s.n = (int)idx.size();
    s.n_pad = std::max(round_up(s.n, pad_multiple), pad_multiple);
    // the index list goes up; the subset's arrays are gathered on the device from the list-order atom arrays, which
    // repack() has uploaded before (subset_gather_kernel)
    std::vector<int> di(s.n_pad, 0);
    std::copy(idx.begin(), idx.end(), di.begin());
    cudaStream_t st = c->stream;
    const size_t np = (size_t)s.n_pad;
    if (upload(s.didx, di, st)) return 11;
    if (s.x.ensure(np) || s.y.ensure(np) || s.z.ensure(np) || s.q.ensure(np) || s.alpha.ensure(np) || s.aux.ensure(np) ||
        s.mol.ensure(np) || s.flags.ensure(np))
        FAIL(11, "cudaMalloc failed");
    if (with_lj && (s.eps.ensure(np) || s.sig.ensure(np) || s.mm.ensure(np) || s.cbox.ensure((size_t)7 * (s.n_pad / 32))))
        FAIL(11, "cudaMalloc failed");
    subset_gather_kernel<<<(s.n_pad + 255) / 256, 256, 0, st>>>(
        s.n, s.n_pad, s.didx.p, c->x.p, c->y.p, c->z.p, c->q.p, c->alpha.p, c->eps.p, c->sig.p, c->molmass.p, c->mol.p, c->flags.p,
        with_aux ? 1 : 0, P.damp_type == MPMC_DAMP_LINEAR ? 1 : 0, with_lj ? 1 : 0, s.x.p, s.y.p, s.z.p, s.q.p, s.alpha.p, s.aux.p,
        with_lj ? s.eps.p : nullptr, with_lj ? s.sig.p : nullptr, with_lj ? s.mm.p : nullptr, s.mol.p, s.flags.p);
    CU(cudaGetLastError());
    return 0;
}

// the order `old_idx` of a set of atoms after a small edit of the atom set (edit_map): survivors keep their place, the new
// atoms that belong to the set (`member`) go in right behind the survivor nearest to them (minimum image) -- at the end
// of the order they would make one cluster with a box as large as the cell, and the CTA that owns it in the row-owned
// cutoff kernels would walk every other cluster (a straggler on the critical path: +1.4 ms at 63k atoms)
template <class Member>
static std::vector<int> carried_order(const Ctx* c, const std::vector<int>& old_idx, Member member) {
    const mpmc_b200_params& P = c->P;
    const int nmap = (int)c->edit_map.size();
    // the atoms to place, in runs of consecutive indices (a molecule's atoms are consecutive): one anchor per run
    std::vector<int> fresh;
    for (int i : c->edit_fresh)
        if (member(i)) fresh.push_back(i);
    struct Run { int b, e, anchor; };
    std::vector<Run> runs;
    for (size_t k = 0; k < fresh.size();) {
        size_t e = k + 1;
        while (e < fresh.size() && fresh[e] == fresh[e - 1] + 1 && c->hmol[fresh[e]] == c->hmol[fresh[k]]) e++;
        runs.push_back({(int)k, (int)e, -1});
        k = e;
    }
    const bool ortho = is_ortho(P);
    const double Lx = P.basis[0][0], Ly = P.basis[1][1], Lz = P.basis[2][2];
    const double ix = 1.0 / Lx, iy = 1.0 / Ly, iz = 1.0 / Lz;
    for (Run& r : runs) {  // nearest survivor of the set (minimum image)
        const int a = fresh[r.b];
        const double xa = c->hx[a], ya = c->hy[a], za = c->hz[a];
        double best = 1e300;
        for (int g : old_idx) {
            const int m = (g >= 0 && g < nmap) ? c->edit_map[g] : -1;
            if (m < 0) continue;
            double dx = c->hx[m] - xa, dy = c->hy[m] - ya, dz = c->hz[m] - za, r2;
            if (ortho) {  // the common case: per-axis wrap, no matrix products
                dx -= Lx * rint(dx * ix);
                dy -= Ly * rint(dy * iy);
                dz -= Lz * rint(dz * iz);
                r2 = dx * dx + dy * dy + dz * dz;
            } else {
                double f[3], w[3] = {0.0, 0.0, 0.0};
                for (int p = 0; p < 3; p++) {
                    f[p] = P.reciprocal_basis[0][p] * dx + P.reciprocal_basis[1][p] * dy + P.reciprocal_basis[2][p] * dz;
                    f[p] -= rint(f[p]);
                }
                for (int p = 0; p < 3; p++)
                    for (int q = 0; q < 3; q++) w[q] += P.basis[p][q] * f[p];
                r2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2];
            }
            if (r2 < best) {
                best = r2;
                r.anchor = m;
            }
        }
    }
    // runs by anchor: head[anchor] -> first run, next[run] -> the following one with the same anchor
    std::vector<int> head(c->n > 0 ? c->n : 1, -1), next(runs.size(), -1);
    std::vector<int> loose;
    for (int k = (int)runs.size() - 1; k >= 0; k--) {
        if (runs[k].anchor < 0) {
            loose.push_back(k);
            continue;
        }
        next[k] = head[runs[k].anchor];
        head[runs[k].anchor] = k;
    }
    std::vector<int> out;
    out.reserve(old_idx.size() + fresh.size());
    for (int g : old_idx) {
        const int m = (g >= 0 && g < nmap) ? c->edit_map[g] : -1;
        if (m < 0) continue;
        out.push_back(m);
        for (int k = head[m]; k >= 0; k = next[k]) out.insert(out.end(), fresh.begin() + runs[k].b, fresh.begin() + runs[k].e);
    }
    for (int k : loose) out.insert(out.end(), fresh.begin() + runs[k].b, fresh.begin() + runs[k].e);
    return out;
}

static int repack(Ctx* c) {
    const int n = c->n;
    // small edit of the atom set: the spatial orders are carried over (see edit_map); anything else sorts afresh
    const bool carry = !c->edit_map.empty() && c->orders_valid;
    std::vector<double> sx_old, sy_old, sz_old;
    if (carry && c->hx_sort.size() == c->edit_map.size()) {
        sx_old.swap(c->hx_sort), sy_old.swap(c->hy_sort), sz_old.swap(c->hz_sort);
    }
    auto sort_positions = [&]() {  // the positions the drift check of set_positions compares with
        c->hx_sort = c->hx, c->hy_sort = c->hy, c->hz_sort = c->hz;
        if (carry && !sx_old.empty())
            for (size_t g = 0; g < c->edit_map.size(); g++) {
                const int m = c->edit_map[g];
                if (m >= 0) c->hx_sort[m] = sx_old[g], c->hy_sort[m] = sy_old[g], c->hz_sort[m] = sz_old[g];
            }
    };
    static const bool timing = getenv("MPMC_B200_REPACK_TIMING") != nullptr;
    const auto t_begin = std::chrono::steady_clock::now();
    double t_sort = 0.0;
    auto timed_sort = [&](std::vector<int>& idx) {
        const auto t0 = std::chrono::steady_clock::now();
        spatial_order(c, idx);
        t_sort += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    };
    c->layout_dirty = true;
    c->inc_valid = false;
    c->inc_moved.clear();
    c->inc_slot.assign(n, 0);
    c->warm_valid = false;
    cudaStream_t st = c->stream;
    c->n_pad = std::max(round_up(n, TILE * std::max(c->world, 1)), TILE);
    const int np = c->n_pad;
    auto padded = [&](const std::vector<double>& h) {
        std::vector<double> v(np, 0.0);
        std::copy(h.begin(), h.end(), v.begin());
        return v;
    };
    std::vector<int> mol(np, -1), fl(np, 0);
    for (int i = 0; i < n; i++) {
        mol[i] = c->hmol[i];
        fl[i] = F_VALID | (c->hfrozen[i] ? F_FROZEN : 0);
    }
    if (upload(c->x, padded(c->hx), st) || upload(c->y, padded(c->hy), st) || upload(c->z, padded(c->hz), st) ||
        upload(c->q, padded(c->hq), st) || upload(c->eps, padded(c->heps), st) || upload(c->sig, padded(c->hsig), st) ||
        upload(c->molmass, padded(c->hmm), st) || upload(c->alpha, padded(c->halpha), st) || upload(c->mol, mol, st) ||
        upload(c->flags, fl, st))
        return 11;

    // subsets
    std::vector<int> charged, recip, pol, nonpol;
    for (int i = 0; i < n; i++) {
        if (c->halpha[i] == 0.0) nonpol.push_back(i);
        if (c->hq[i] != 0.0) {
            charged.push_back(i);
            if (!c->hfrozen[i]) recip.push_back(i);  // coulombic.c:74-75
        }
        if (c->halpha[i] != 0.0) pol.push_back(i);
    }
    // Jacobi family: the polarizable list is kept in spatial order (symmetric K4); Gauss-Seidel processes
    // atoms in list / rank order and keeps the list order
    c->pol_sorted = false;
    if (c->use_sym && (!(c->P.polar_gs || c->P.polar_gs_ranked) || c->P.polar_ewald_full) && c->have_params) {
        if (carry && !c->pol.idx.empty())
            pol = carried_order(c, c->pol.idx, [&](int i) { return c->halpha[i] != 0.0; });
        else
            timed_sort(pol);
        c->pol_sorted = true;
        sort_positions();
    }
    const int polpad = dipole_sym_tile() * std::max(c->world, 1);  // super tiles of the symmetric K4, equal shards
    if (pack_subset(c, c->charged, charged, TILE, false) || pack_subset(c, c->recip, recip, TILE, false) ||
        pack_subset(c, c->pol, pol, polpad, true) || pack_subset(c, c->nonpol, nonpol, polpad, true))
        return 11;
    if (c->eic_np.ensure((size_t)3 * c->nonpol.n_pad) || c->eind_np.ensure((size_t)3 * c->nonpol.n_pad))
        FAIL(11, "cudaMalloc failed");
    c->gsp.n = 0;
    if (c->gs_spatial && c->gs_delta && c->have_params && (c->P.polar_gs || c->P.polar_gs_ranked) && !c->P.polar_ewald_full &&
        c->pol.n > 0) {
        std::vector<int> sp = c->pol.idx;  // list order
        if (carry && !c->gsp.idx.empty())
            sp = carried_order(c, c->gsp.idx, [&](int i) { return c->halpha[i] != 0.0; });
        else
            timed_sort(sp);
        if (pack_subset(c, c->gsp, sp, polpad, true)) return 11;
        std::vector<int> packed_of_atom(n, -1);
        for (int k = 0; k < c->pol.n; k++) packed_of_atom[c->pol.idx[k]] = k;
        c->gsp_packed.assign(c->gsp.n_pad, 0);
        for (int k = 0; k < c->gsp.n_pad; k++) c->gsp_packed[k] = k < c->gsp.n ? packed_of_atom[sp[k]] : k;
        const size_t gp = (size_t)c->gsp.n_pad;
        if (c->map_s2g.ensure(gp) || c->map_g2s.ensure(gp) || c->g_mu_sp.ensure(3 * gp) || c->g_esum_sp.ensure(3 * gp) ||
            c->cbox_sp.ensure(7 * (gp / 32)))
            FAIL(11, "cudaMalloc failed");
        if (!c->pol_sorted && !c->use_clusters) sort_positions();  // the drift check of set_positions needs the sort-time positions
    }
    c->hperm.assign(c->pol.n_pad, 0);
    std::iota(c->hperm.begin(), c->hperm.end(), 0);

    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device);
    if (c->use_clusters) {
        // spatially sorted image of ALL atoms for the cutoff kernels (K1, K3)
        std::vector<int> all(n);
        std::iota(all.begin(), all.end(), 0);
        if (carry && !c->sorted.idx.empty())
            all = carried_order(c, c->sorted.idx, [](int) { return true; });
        else
            timed_sort(all);
        if (pack_subset(c, c->sorted, all, 32, false, true)) return 11;
        if (!c->pol_sorted) sort_positions();  // the drift check of set_positions needs the sort-time positions in either case
        // intramolecular erf term: list order, molecules are contiguous (pairs.c:28-34)
        std::vector<int> follow(np, 0);
        double same_mol_pairs = 0.0;  // pairs of one molecule that are not frozen-frozen
        for (int b = 0; b < n;) {
            int e = b, nfz = 0;
            while (e < n && c->hmol[e] == c->hmol[b]) nfz += c->hfrozen[e++];
            const int m = e - b;
            if (nfz < m)
                for (int i = b; i < e; i++) follow[i] = e - 1 - i;
            same_mol_pairs += 0.5 * m * (m - 1.0) - 0.5 * nfz * (nfz - 1.0);
            b = e;
        }
        double nfrozen = 0.0;
        for (int i = 0; i < n; i++) nfrozen += c->hfrozen[i];
        // pairs the reference evaluates geometry for in lj()/coulombic_real(): not frozen-frozen, not same molecule
        c->pairs_tested_analytic = 0.5 * n * (n - 1.0) - 0.5 * nfrozen * (nfrozen - 1.0) - same_mol_pairs;
        if (upload(c->follow, follow, st)) return 11;
        if (c->intra_part.ensure((size_t)(np + 255) / 256 + 8)) FAIL(11, "cudaMalloc failed");
        const int nclus = c->sorted.n_pad / 32;
        const int own = std::max(1, nclus / std::max(c->world, 1) + 1);
        const int ctas_x = (own + 3) / 4;
        // Without polarization K1 is the critical path: slice down to one candidate offset per warp (small systems are
        // bound by a warp's serial walk).  With polarization K1 runs beside the solve on the side stream and only has to
        // finish before it does: whole batches of 32 offsets per slice, so that it takes few issue slots from K3 / K4.
        // (small systems: the fused solve is shorter than a warp's walk over whole batches -- K1 0.41-0.90 ms at 1k-4k atoms)
        const bool k1_critical = !(c->P.polarization && !c->P.rd_only) || n <= 8192;
        const int batches = k1_critical ? nclus / 2 + 1 : (nclus / 2 + 1 + 31) / 32;
        // CTAs aimed at: ~8 per SM beside the solve; ~28 per SM where K1 is the critical path -- the grid is then many waves
        // of short CTAs instead of 1.4 waves of long ones (co2_fluid_12k: 13 -> 38 slices, K1 0.676 -> 0.585 ms; optimum 39-78)
        const int per_sm = k1_critical ? 28 : 8;
        c->pair_kslices = std::max(1, std::min(batches, (per_sm * sms + ctas_x - 1) / ctas_x));
        if (const char* v = getenv("MPMC_B200_K1_SLICES"))  // A/B: slices of the candidate offsets per i-cluster
            if (atoi(v) > 0) c->pair_kslices = std::min(atoi(v), nclus / 2 + 1);
        const int cols = pair_cluster_columns(c->sorted.n_pad, 0, own, c->pair_kslices);
        c->ntiles = 0;
        c->pair_stride = round_up(std::max(cols, 1), 32);
        if (c->pair_part.ensure((size_t)PAIR_NPART * c->pair_stride)) FAIL(11, "cudaMalloc failed");
    } else {
    // tile pairs of the upper triangle, skipping tile pairs in which every atom is frozen
        const int nt = np / TILE;
        std::vector<char> all_frozen(nt, 1);
        for (int t = 0; t < nt; t++)
            for (int k = t * TILE; k < std::min(n, (t + 1) * TILE); k++)
                if (!c->hfrozen[k]) all_frozen[t] = 0;
        std::vector<int2> tiles;
        tiles.reserve((size_t)nt * (nt + 1) / 2);
        // interleave short (near-diagonal-first) ordering is irrelevant: every tile pair costs the same
        for (int I = 0; I < nt; I++)
            for (int J = I; J < nt; J++) {
                if (all_frozen[I] && all_frozen[J]) continue;
                tiles.push_back(make_int2(I, J));
            }
        c->ntiles = (int)tiles.size();
        if (tiles.empty()) tiles.push_back(make_int2(0, 0));
        if (upload(c->tiles, tiles, st)) return 11;
        c->pair_stride = round_up(std::max(c->ntiles, 1), 32);
        if (c->pair_part.ensure((size_t)PAIR_NPART * c->pair_stride)) FAIL(11, "cudaMalloc failed");

    }

    c->crystal_ntiles = 0;
    if (c->P.rd_crystal) {  // all tile pairs of the upper triangle except those in which every atom is frozen
        const int nt = np / TILE;
        std::vector<char> all_frozen(nt, 1);
        for (int t = 0; t < nt; t++)
            for (int k = t * TILE; k < std::min(n, (t + 1) * TILE); k++)
                if (!c->hfrozen[k]) all_frozen[t] = 0;
        std::vector<int2> ct;
        for (int I = 0; I < nt; I++)
            for (int J = I; J < nt; J++)
                if (!(all_frozen[I] && all_frozen[J])) ct.push_back(make_int2(I, J));
        c->crystal_ntiles = (int)ct.size();
        if (ct.empty()) ct.push_back(make_int2(0, 0));
        if (upload(c->crystal_tiles, ct, st)) return 11;
        c->crystal_stride = round_up(std::max(c->crystal_ntiles, 1), 32);
        if (c->crystal_part.ensure((size_t)2 * c->crystal_stride)) FAIL(11, "cudaMalloc failed");
    }

    // reciprocal chunks: about two CTAs per SM
    const int target = 2 * sms;
    c->recip_apc = std::max(64, round_up((std::max(c->charged.n, 1) + target - 1) / target, 64));
    // (chunking is computed per use from the subset actually summed)

    // polarization buffers
    const int pp = c->pol.n_pad;
    const int ptiles = pp / TILE;
    // j-splits so that (row tiles / world) x splits is a few waves of CTAs
    int rows_tiles = std::max(1, ptiles / std::max(c->world, 1));
    int js = std::max(1, std::min(ptiles, (4 * sms + rows_tiles - 1) / rows_tiles));
    c->dipole_jsplits = js;
    const int ctiles = c->charged.n_pad / TILE;
    int frows = std::max(1, (np / TILE) / std::max(c->world, 1));
    c->field_jsplits = std::max(1, std::min(ctiles, (4 * sms + frows - 1) / frows));
    const size_t gs_splits = (size_t)std::min(ptiles, 148);
    size_t epart_need = std::max({(size_t)c->field_jsplits * 3 * np, (size_t)c->dipole_jsplits * 3 * pp, gs_splits * 3 * pp});
    if (c->ef_static.ensure((size_t)3 * np) || c->epart.ensure(epart_need) || c->es.ensure((size_t)3 * pp) ||
        c->mu.ensure((size_t)3 * pp) || c->mu_old.ensure((size_t)3 * pp) || c->mu_new.ensure((size_t)3 * pp) ||
        c->eind.ensure((size_t)3 * pp) || c->eic.ensure((size_t)3 * pp) || c->rrms.ensure(pp) ||
        c->rank_packed.ensure(pp) || c->rmin_part.ensure((size_t)8 * ptiles + 8) ||
        c->field_cnt.ensure((size_t)std::max(c->field_jsplits * (np / TILE), np / 32) + 40))
        FAIL(11, "cudaMalloc failed");
    {
        // symmetric K4: CTAs = I tiles x offset slices; a few waves of 4 CTAs per SM
        const int stiles = pp / dipole_sym_tile();
        const int own_tiles = std::max(1, stiles / std::max(c->world, 1));
        const int K1 = stiles / 2 + 1;
        int per_sm = 48;  // measured on M4: 16 -> 48 CTAs per SM removes most of the last-wave tail
        if (const char* v = getenv("MPMC_B200_SYM_CTAS_PER_SM")) per_sm = std::max(1, atoi(v));
        c->sym_ysplits = std::max(1, std::min(K1, (per_sm * sms + own_tiles - 1) / own_tiles));
        if (c->cbox.ensure((size_t)7 * (pp / 32)) || c->sym_partI.ensure((size_t)std::max(2, dipole_sym_hsplit(pp, c->sym_ysplits, c->world)) * c->sym_ysplits * 3 * pp) ||
            c->sym_partJ.ensure((size_t)dipole_sym_partial_rows(pp, c->sym_ysplits) * 3 * pp) || c->esum.ensure((size_t)3 * pp))
            FAIL(11, "cudaMalloc failed");
    }
    CU(cudaMemsetAsync(c->ef_static.p, 0, sizeof(double) * 3 * np, st));
    for (DevBuf<double>* b : {&c->es, &c->mu, &c->mu_old, &c->mu_new, &c->eind, &c->eic})
        CU(cudaMemsetAsync(b->p, 0, sizeof(double) * 3 * pp, st));
    CU(cudaMemsetAsync(c->rrms.p, 0, sizeof(double) * pp, st));
    CU(cudaMemsetAsync(c->rank_packed.p, 0, sizeof(double) * pp, st));
    c->analytic_dirty = true;
    c->subsets_stale = false;
    c->have_atoms = true;
    if (!carry) c->order_edits = 0;
    c->orders_valid = c->have_params;
    c->edit_map.clear();
    if (timing)
        fprintf(stderr, "MPMC_B200_REPACK_TIMING: n %d, %s, spatial sorts %.3f ms, whole repack %.3f ms (host work + enqueued uploads)\n", n,
                carry ? "orders carried over" : "full sort", t_sort,
                std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count());
    return 0;
}

static void shard_range(int total_units, int rank, int world, int& b, int& e) {
    b = (int)((long long)total_units * rank / world);
    e = (int)((long long)total_units * (rank + 1) / world);
}

static void phase_begin(Ctx* c, Phase p, cudaStream_t st = nullptr) {
    cudaEventRecord(c->ev[p][0], st ? st : c->stream);
    c->ev_used[p] = true;
}
static void phase_end(Ctx* c, Phase p, cudaStream_t st = nullptr) { cudaEventRecord(c->ev[p][1], st ? st : c->stream); }

// The pair (K1) and reciprocal (K2) energies do not feed the polarization solve: they run on a side stream
// forked from the engine's stream and are joined before the evaluation's results are read.
static void join_side(Ctx* c) {
    if (!c->side_active) return;
    cudaStreamWaitEvent(c->stream, c->ev_join, 0);
    c->side_active = false;
}

// bracket one K4 launch with its own event pair (at most 160 per evaluation are timed)
static void kev_mark(Ctx* c) {
    if (c->kev_used >= 320) return;
    if ((int)c->kev.size() <= c->kev_used) {
        cudaEvent_t e;
        cudaEventCreate(&e);
        c->kev.push_back(e);
    }
    cudaEventRecord(c->kev[c->kev_used++], c->stream);
}

static int refresh_subsets(Ctx* c) {
    // a few atoms jumped (a large displacement move, or its restore): they leave their place in the orders and go in
    // again behind their nearest neighbour; everything else stays -- O(n) host work and one index upload + gather per
    // subset instead of the k-d sorts (16 ms at 63k atoms)
    static const bool reuse = [] { const char* v = getenv("MPMC_B200_REUSE_ORDER"); return !(v && v[0] == '0'); }();
    if (c->resort_pending && reuse && c->drift_tracked && c->orders_valid && !c->drifted.empty() && c->n >= 1024 &&
        c->hx_sort.size() == (size_t)c->n) {
        std::sort(c->drifted.begin(), c->drifted.end());
        c->drifted.erase(std::unique(c->drifted.begin(), c->drifted.end()), c->drifted.end());
        const int nd = (int)c->drifted.size();
        if (nd <= 256 && c->order_edits + nd <= std::max(64, c->n / 128)) {
            c->edit_map.resize(c->n);
            std::iota(c->edit_map.begin(), c->edit_map.end(), 0);
            for (int g : c->drifted) c->edit_map[g] = -1;
            c->edit_fresh = c->drifted;
            const int polpad = dipole_sym_tile() * std::max(c->world, 1);
            int rc = 0;
            if (c->use_clusters && !c->sorted.idx.empty())
                rc = pack_subset(c, c->sorted, carried_order(c, c->sorted.idx, [](int) { return true; }), 32, false, true);
            if (!rc && c->pol_sorted && !c->pol.idx.empty())
                rc = pack_subset(c, c->pol, carried_order(c, c->pol.idx, [&](int i) { return c->halpha[i] != 0.0; }), polpad, true);
            if (!rc && c->gsp.n > 0) {
                const std::vector<int> sp = carried_order(c, c->gsp.idx, [&](int i) { return c->halpha[i] != 0.0; });
                rc = pack_subset(c, c->gsp, sp, polpad, true);
                std::vector<int> packed_of_atom(c->n, -1);
                for (int k = 0; k < c->pol.n; k++) packed_of_atom[c->pol.idx[k]] = k;
                for (int k = 0; k < c->gsp.n_pad; k++) c->gsp_packed[k] = k < c->gsp.n ? packed_of_atom[sp[k]] : k;
            }
            c->edit_map.clear();
            if (rc) return 11;
            for (int g : c->drifted) c->hx_sort[g] = c->hx[g], c->hy_sort[g] = c->hy[g], c->hz_sort[g] = c->hz[g];
            c->order_edits += nd;
            c->resort_pending = false;
            c->subsets_stale = true;  // the other subsets still hold the old coordinates
        }
    }
    c->drifted.clear();
    c->drift_tracked = false;
    if (c->resort_pending && c->use_clusters) {
        std::vector<int> all(c->n);
        std::iota(all.begin(), all.end(), 0);
        spatial_order(c, all);
        if (pack_subset(c, c->sorted, all, 32, false, true)) return 11;
        if (!c->pol_sorted) {
            c->hx_sort = c->hx;
            c->hy_sort = c->hy;
            c->hz_sort = c->hz;
        }
    }
    if (c->resort_pending && c->pol_sorted) {
        std::vector<int> pol = c->pol.idx;
        std::sort(pol.begin(), pol.end());
        spatial_order(c, pol);
        if (pack_subset(c, c->pol, pol, dipole_sym_tile() * std::max(c->world, 1), true)) return 11;
        c->hx_sort = c->hx;
        c->hy_sort = c->hy;
        c->hz_sort = c->hz;
    }
    if (c->resort_pending && c->gsp.n > 0) {  // Gauss-Seidel: the spatial image of the Jacobi part follows the atoms too
        std::vector<int> sp = c->pol.idx;
        spatial_order(c, sp);
        if (pack_subset(c, c->gsp, sp, dipole_sym_tile() * std::max(c->world, 1), true)) return 11;
        std::vector<int> packed_of_atom(c->n, -1);
        for (int k = 0; k < c->pol.n; k++) packed_of_atom[c->pol.idx[k]] = k;
        for (int k = 0; k < c->gsp.n_pad; k++) c->gsp_packed[k] = k < c->gsp.n ? packed_of_atom[sp[k]] : k;
    }
    if (c->resort_pending) c->order_edits = 0;
    c->resort_pending = false;
    if (!c->subsets_stale) return 0;
    for (Subset* s : {&c->charged, &c->recip, &c->pol, &c->nonpol, &c->sorted, &c->gsp}) {
        if (s->n <= 0) continue;
        gather_positions_kernel<<<(s->n + 255) / 256, 256, 0, c->stream>>>(c->x.p, c->y.p, c->z.p, s->didx.p, s->n, s->x.p,
                                                                          s->y.p, s->z.p);
        c->launches++;
    }
    c->subsets_stale = false;
    return 0;
}

static AtomsDev atoms_dev(Ctx* c) {
    AtomsDev a;
    a.n = c->n;
    a.n_pad = c->n_pad;
    a.x = c->x.p;
    a.y = c->y.p;
    a.z = c->z.p;
    a.q = c->q.p;
    a.eps = c->eps.p;
    a.sig = c->sig.p;
    a.molmass = c->molmass.p;
    a.alpha = c->alpha.p;
    a.mol = c->mol.p;
    a.flags = c->flags.p;
    return a;
}

static int check_supported(const mpmc_b200_params& P) {
    if (P.polarization && !P.rd_only) {
        if (!P.polar_iterative && !P.polar_zodid && !P.polar_ewald_full)
            FAIL(3, "matrix-inversion polarization (polar_iterative off) is out of scope: O(N^3) dense LU");
        if (!P.polar_zodid && !P.polar_ewald_full && P.polar_precision == 0.0 && P.polar_max_iter <= 0)
            FAIL(3, "polar_max_iter must be > 0 when polar_precision is 0");
        if (P.damp_type < 0 || P.damp_type > 2) FAIL(3, "unknown damp_type");
    }
    if (P.feynman_hibbs && !(P.temperature > 0.0)) FAIL(3, "feynman_hibbs needs temperature > 0");
    // coulombic_wolf(): src/energy/coulombic.c:293-297 -- the reference dies on this combination
    if (P.es_wolf && !P.rd_only && P.feynman_hibbs) FAIL(3, "COULOMBIC: FH + es_wolf is not implemented");
    if (P.rd_crystal) {  // check_input.c:1258-1262
        if (P.rd_crystal_order <= 0) FAIL(3, "rd crystal order must be positive");
        const int side = 2 * P.rd_crystal_order - 1;
        if (side * side * side > crystal_max_images()) FAIL(3, "rd_crystal_order > 4 is not supported");
    }
    return 0;
}

// ---- reciprocal sum over a subset; leaves S(k) in c->sf and (optionally) the energy in a scalar slot
static int run_structure_factor(Ctx* c, const Subset& sub, const DevBuf<double>& weight, double scale, double* energy_slot,
                                bool sharded, cudaStream_t st = nullptr) {
    if (!st) st = c->stream;
    KVecDev kv;
    kv.nk = c->nk;
    kv.kmax = c->P.ewald_kmax;
    kv.l = c->kl.p;
    kv.weight = weight.p;
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device);
    const int target = 2 * sms;
    const int apc = std::max(64, round_up((std::max(sub.n, 1) + target - 1) / target, 64));
    const int nchunks = std::max(1, (sub.n + apc - 1) / apc);
    // Every rank accumulates the complete S(k) (0.1-0.2 ms at 64k atoms; |S|^2 is not linear in
    // atom partials); the SUM over k-vectors is what is split across ranks.
    const int cb = 0, ce = nchunks;
    const int mine = nchunks;
    int kb = 0, ke = c->nk;
    if (sharded) shard_range(c->nk, c->rank, c->world, kb, ke);
    if (c->sf_partial.ensure((size_t)std::max(mine, 1) * 2 * std::max(c->nk, 1))) FAIL(11, "cudaMalloc failed");
    if (c->nk > 0 && sub.n > 0 && mine > 0) {
        launch_recip_sf(make_cell(c->P), sub.dev(), kv, cb, ce, apc, c->sf_partial.p, st);
        c->launches++;
    }
    c->launches += launch_recip_finish(kv, c->sf_partial.p, (c->nk > 0 && sub.n > 0) ? mine : 0, scale, c->sf.p, energy_slot, kb, ke, st);
    CU(cudaGetLastError());
    return 0;
}

// ---- peer-memory exchange: layout of one rank's exported allocation
// doubles: ind [2][3][cap] | field [3][cap_atoms] | scal [2][64]; then CH_COUNT x PEER_MAX flags
enum PeerChannel { CH_IND = 0, CH_FIELD = 1, CH_SCAL = 2, CH_COUNT = 3 };
static size_t peer_doubles(size_t cap, size_t cap_atoms) { return 2 * 3 * cap + 3 * cap_atoms + 2 * 64; }
static size_t peer_bytes(size_t cap, size_t cap_atoms) {
    return sizeof(double) * peer_doubles(cap, cap_atoms) + sizeof(unsigned long long) * CH_COUNT * PEER_MAX + 64;
}
static double* peer_buf(void* base, size_t cap, int parity) { return reinterpret_cast<double*>(base) + (size_t)parity * 3 * cap; }
static double* peer_field(void* base, size_t cap) { return reinterpret_cast<double*>(base) + 2 * 3 * cap; }
static double* peer_scal(void* base, size_t cap, size_t cap_atoms, int parity) {
    return reinterpret_cast<double*>(base) + 2 * 3 * cap + 3 * cap_atoms + (size_t)parity * 64;
}
static unsigned long long* peer_flags(void* base, size_t cap, size_t cap_atoms, int ch) {
    return reinterpret_cast<unsigned long long*>(reinterpret_cast<double*>(base) + peer_doubles(cap, cap_atoms)) + ch * PEER_MAX;
}
// Which Jacobi contraction runs.  The symmetric kernel's work item is half a 256 x 256 tile pair (~65 us of one CTA):
// for a small system its grid does not fill the GPU and an evaluation costs one item per iteration however small the
// system, while the row form with the j range fully split ((n/128)^2 CTAs of 128 x 128 pairs) finishes in 10-60 us
// despite doing every pair twice.  Crossover on one GPU (profiles/r02_sym_min.md, half-tile items): 3,840 padded
// polarizable atoms 0.93 (rows) against 0.96 ms (symmetric) per energy(), 4,864: 1.66 against 1.08, 7,296: 2.93 against
// 1.65; with the items cut down to single j-clusters for short launches (dipole_sym_hsplit) also 3,072: 0.99 against 0.91,
// 3,840: 1.31 against 1.09, while 2,560 stays with the rows (0.73 against 1.13 with half items) -- so 3,072 (8,192 while
// the item was a whole tile pair).  Sharded runs always use the symmetric form (its exchange protocol).
static bool sym_now(const Ctx* c) {
    const char* v = getenv("MPMC_B200_SYM_MIN");  // read per call: the tests switch it inside one process
    const int min_pad = v ? atoi(v) : 3072;
    return c->use_sym && (c->world > 1 || c->pol.n_pad >= min_pad);
}
static bool peer_active(const Ctx* c) {
    return c->peer_ready && c->world > 1 && c->use_sym && (size_t)c->pol.n_pad == c->peer_cap &&
           (size_t)c->n_pad == c->peer_cap_atoms;
}
static void peer_teardown(Ctx* c) {
    for (int r = 0; r < PEER_MAX; r++) {
        if (c->peer_base[r] && r != c->rank && c->peer_ipc) cudaIpcCloseMemHandle(c->peer_base[r]);
        c->peer_base[r] = nullptr;
    }
    if (c->peer_local) cudaFree(c->peer_local);
    c->peer_local = nullptr;
    c->peer_cap = c->peer_cap_atoms = 0;
    c->peer_ready = false;
    c->peer_ipc = false;
}

// consumer side of channel `ch` at its current sequence number
static PeerWait peer_wait_args(Ctx* c, int ch = CH_IND) {
    PeerWait pw;
    memset(&pw, 0, sizeof(pw));
    pw.world = c->world;
    pw.seq = c->peer_seq[ch];
    pw.my_flags = peer_flags(c->peer_base[c->rank], c->peer_cap, c->peer_cap_atoms, ch);
    const int parity = (int)(pw.seq & 1ull);
    for (int r = 0; r < c->world; r++)
        pw.src[r] = ch == CH_IND ? peer_buf(c->peer_base[r], c->peer_cap, parity)
                    : ch == CH_FIELD ? peer_field(c->peer_base[r], c->peer_cap)
                                     : peer_scal(c->peer_base[r], c->peer_cap, c->peer_cap_atoms, parity);
    pw.timeout_flag = c->scalars.p + S_PEER_TIMEOUT;
    return pw;
}
// producer side: next sequence number of channel `ch`, flag slot `rank` of every peer
static PeerSignal peer_signal_args(Ctx* c, int ch) {
    PeerSignal sg;
    memset(&sg, 0, sizeof(sg));
    sg.world = c->world;
    sg.me = c->rank;
    sg.seq = ++c->peer_seq[ch];
    for (int r = 0; r < c->world; r++) sg.flags[r] = peer_flags(c->peer_base[r], c->peer_cap, c->peer_cap_atoms, ch);
    return sg;
}

// ---- polar(): static field + solver + energy.  Enqueue only (precision mode syncs per iteration).
// row ranges of this shard (tile aligned) and per-evaluation counters
static void polar_prepare(Ctx* c) {
    const int np = c->n_pad, pp = c->pol.n_pad;
    c->pending_iterations = 0;
    c->pending_iter_success = 0;
    c->eic_np_valid = false;
    c->eind_np_valid = false;
    c->cnt.contractions = 0;
    c->cnt.n_polarizable = c->pol.n;
    c->cnt.pairs_dipole = (double)c->pol.n * std::max(c->pol.n - 1, 0);
    int tb, te;
    shard_range(np / TILE, c->rank, c->world, tb, te);
    c->fb = tb * TILE;
    c->fe = std::min(te * TILE, c->n);
    shard_range(pp / TILE, c->rank, c->world, tb, te);
    c->pb = tb * TILE;
    c->pe = te * TILE;
}

// thole_field() (thole_field.c:14-36) for the atom rows [fb, fe) of this shard
static int polar_field(Ctx* c) {
    const mpmc_b200_params& P = c->P;
    cudaStream_t st = c->stream;
    const Cell cell = make_cell(P);
    const bool ortho = is_ortho(P);
    const AtomsDev a = atoms_dev(c);
    const int np = c->n_pad;
    const int fb = c->fb, fe = c->fe;
    // ---------------- static field: thole_field() (thole_field.c:14-36) ----------------
    phase_begin(c, PH_FIELD);
    // sharded_eval(): this rank's rows go to its exported buffer (zero elsewhere); peer_sum_kernel of every rank then
    // adds the `world` buffers in rank order into ef_static -- the all-gather of the row-owned field over NVLink
    double* const ef_dst = c->field_via_peer ? peer_field(c->peer_base[c->rank], c->peer_cap) : c->ef_static.p;
    CU(cudaMemsetAsync(ef_dst, 0, sizeof(double) * 3 * np, st));
    FieldOpts fo;
    fo.table = nullptr;
    fo.table_max = 0;
    fo.cutoffterm = 0.0;
    fo.a = 0.0;
    bool accumulate = false;
    if (P.polar_ewald || P.polar_ewald_full) {  // ewald_full() computes its own Ewald static field (polar_ewald.c:379-382)
        fo.mode = FIELD_EWALD_REAL;
        fo.a = P.polar_ewald_alpha;
        // recip_term(): S(k) over ALL atoms (polar_ewald.c:104-108); zero-charge atoms add exact zeros
        if (run_structure_factor(c, c->charged, c->kweight_field, 0.0, nullptr, false)) return 10;
        KVecDev kv;
        kv.nk = c->nk;
        kv.kmax = P.ewald_kmax;
        kv.l = c->kl.p;
        kv.weight = c->kweight_field.p;
        launch_recip_field(cell, a, kv, c->kvec.p, c->sf.p, 8.0 * M_PI / P.volume, fb, fe, ef_dst, st);
        c->launches++;
        accumulate = true;
    } else if (P.polar_wolf || P.polar_wolf_full) {
        const double aa = P.polar_wolf_alpha, R = P.cutoff, rR = 1.0 / R;
        fo.a = aa;
        fo.cutoffterm = (erfc(aa * R) * rR * rR + 2.0 * aa * 0.56418958354 * exp(-aa * aa * R * R) * rR);
        if (aa == 0.0)
            fo.mode = FIELD_WOLF_A0;
        else if (P.polar_wolf_alpha_lookup) {
            fo.mode = FIELD_WOLF_LOOKUP;
            if (c->wolf_table_alpha != aa || c->wolf_table_cut != P.polar_wolf_alpha_lookup_cutoff) {
                // polar_wolf_alpha_lookup_init(): src/polarization/polar_wolf_lookup.c:8-26
                const int tmax = (int)(ceil(P.polar_wolf_alpha_lookup_cutoff) * 1000);
                std::vector<double> tab(std::max(tmax, 1), 0.0);
                for (int i = 1; i < tmax; i++) {
                    double r = (double)i / 1000., rr = 1.0 / r;
                    tab[i] = erfc(aa * r) * rr * rr + 2.0 * aa * 0.56418958354 * exp(-aa * aa * r * r) * rr;
                }
                if (tmax > 1) tab[0] = tab[1];
                if (upload(c->wolf_table, tab, st)) return 11;
                c->wolf_table_max = tmax;
                c->wolf_table_alpha = aa;
                c->wolf_table_cut = P.polar_wolf_alpha_lookup_cutoff;
            }
            fo.table = c->wolf_table.p;
            fo.table_max = c->wolf_table_max;
        } else
            fo.mode = FIELD_WOLF;
    } else
        fo.mode = FIELD_NOPBC;
    if (c->use_clusters) {
        if (c->charged.n > 0) {
            const SubsetDev sd = c->sorted.dev();  // cluster boxes were computed by the pair phase of this evaluation
            if (!c->boxes_fresh) {
                launch_cluster_boxes(cell, ortho, sd, sd.cbox, st);
                c->launches++;
            }
            int cb, ce;
            shard_range(c->sorted.n_pad / 32, c->rank, c->world, cb, ce);
            launch_static_field_clusters(cell, ortho, sd, fo, cb, ce, P.cutoff + 1e-6, accumulate, ef_dst, np,
                                         c->field_cnt.p, st);
            sum_reduce_kernel<<<1, 1024, 0, st>>>(c->field_cnt.p, std::max(ce - cb, 0), c->scalars.p + S_FIELDCNT);
            c->launches += 2;
        }
    } else
    if (c->charged.n > 0 && fe > fb) {
        const int gx = (fe - fb + TILE - 1) / TILE;
        launch_static_field(cell, ortho, a, c->charged.dev(), fo, fb, fe, c->field_jsplits, c->epart.p, c->field_cnt.p, st);
        launch_field_reduce(c->epart.p, c->field_jsplits, np, fb, fe, accumulate, ef_dst, st);
        // work counter: per-CTA counts of (row, charged source) contributions -> one scalar
        sum_reduce_kernel<<<1, 1024, 0, st>>>(c->field_cnt.p, gx * c->field_jsplits, c->scalars.p + S_FIELDCNT);
        c->launches += 3;
    }
    if (c->field_via_peer) {
        const PeerSignal sg = peer_signal_args(c, CH_FIELD);
        launch_peer_signal(sg, st);
        c->launches++;
    }
    phase_end(c, PH_FIELD);
    CU(cudaGetLastError());
    return 0;
}

static DipoleOpts dipole_opts(const mpmc_b200_params& P) {
    DipoleOpts dop;
    dop.damp_type = P.damp_type;
    dop.wolf_full = P.polar_wolf_full;
    dop.damp = P.polar_damp;
    return dop;
}
static UpdateOpts update_opts(const mpmc_b200_params& P, int iteration) {
    UpdateOpts uo;
    uo.sor = P.polar_sor;
    uo.esor = P.polar_esor;
    uo.gamma = P.polar_gamma;
    uo.save_old = (P.polar_rrms || P.polar_precision > 0 || P.polar_sor || P.polar_esor) ? 1 : 0;
    uo.do_rrms = (P.polar_rrms || P.polar_precision > 0) ? 1 : 0;
    uo.iteration = iteration;
    return uo;
}

// Warm start applies where it buys time and nothing else: the tolerance-mode iteration (polar_precision), one GPU, not
// the zodid / full-Ewald branches.  Fixed-iteration runs keep init_dipoles (their result depends on the start).
static bool warm_start_applies(const Ctx* c) {
    const mpmc_b200_params& P = c->P;
    return c->warm_enabled && c->world == 1 && P.polar_precision != 0.0 && !P.polar_zodid && !P.polar_ewald_full;
}
static bool warm_start_now(const Ctx* c) { return warm_start_applies(c) && c->warm_valid; }

// init_dipoles (thole_iterative.c:13-25) for ALL packed rows (needs the complete static field) and
// the per-evaluation clears
static int polar_init(Ctx* c) {
    const mpmc_b200_params& P = c->P;
    cudaStream_t st = c->stream;
    const int np = c->n_pad, pp = c->pol.n_pad;
    const double gscale = (!P.polar_sor && !P.polar_esor && !P.polar_ewald_full) ? P.polar_gamma : 1.0;  // polar_ewald.c:144
    launch_dipole_init(c->pol.dev(), c->ef_static.p, np, gscale, c->es.p, c->mu.p, 0, pp, st,
                       warm_start_now(c) ? c->mu_keep.p : nullptr);
    c->launches++;
    if ((c->use_sym || P.polar_ewald_full) && c->pol.n > 0) {
        launch_cluster_boxes(make_cell(P), is_ortho(P), c->pol.dev(), c->cbox.p, st);
        c->launches++;
    }
    const bool want_rrms = P.polar_rrms || P.polar_precision > 0;
    if (!want_rrms) CU(cudaMemsetAsync(c->rrms.p, 0, sizeof(double) * pp, st));
    CU(cudaMemsetAsync(c->eind.p, 0, sizeof(double) * 3 * pp, st));
    CU(cudaMemsetAsync(c->eic.p, 0, sizeof(double) * 3 * pp, st));
    CU(cudaMemsetAsync(c->mu_new.p, 0, sizeof(double) * 3 * pp, st));
    CU(cudaGetLastError());
    return 0;
}

// one Jacobi-family pass over this shard's rows: contract_dipoles + calc_dipole_rrms + the
// are_we_done_yet residual (thole_iterative.c:27-117); mu itself is not touched
// symmetric contraction of this shard's I tiles -> c->esum (contribution to E_ind of ALL rows)
static int polar_contract_sym(Ctx* c) {
    const mpmc_b200_params& P = c->P;
    kev_mark(c);
    if (peer_active(c)) {
        // this rank's contribution goes to its exported buffer (double buffered by the parity of the sequence
        // number: a peer may still be reading the previous one); peer_signal_kernel then raises the flags
        const PeerSignal sg = peer_signal_args(c, CH_IND);
        launch_dipole_contract_sym(make_cell(P), is_ortho(P), c->pol.dev(), dipole_opts(P), c->mu.p, c->cbox.p, c->rank, c->world,
                                   c->sym_ysplits, c->sym_partI.p, c->sym_partJ.p,
                                   peer_buf(c->peer_base[c->rank], c->peer_cap, (int)(sg.seq & 1ull)), c->stream, &sg);
    } else
    launch_dipole_contract_sym(make_cell(P), is_ortho(P), c->pol.dev(), dipole_opts(P), c->mu.p, c->cbox.p, c->rank, c->world,
                               c->sym_ysplits, c->sym_partI.p, c->sym_partJ.p, c->esum.p, c->stream);
    kev_mark(c);
    c->launches += peer_active(c) ? 3 : 2;
    c->cnt.contractions += 1;
    CU(cudaGetLastError());
    return 0;
}

// new_mu, residual, rrms from the reduced field (all rows: after the all-reduce every rank holds the complete E_ind)
static int polar_update_sym(Ctx* c, int it) {
    const mpmc_b200_params& P = c->P;
    unsigned long long* errbits = reinterpret_cast<unsigned long long*>(c->scalars.p + S_ERR);
    const bool precision_mode = P.polar_precision != 0.0;
    if (precision_mode) CU(cudaMemsetAsync(errbits, 0, sizeof(unsigned long long), c->stream));
    if (peer_active(c))  // all-reduce + update (+ commit unless palmo still needs the old dipoles) in one kernel
        launch_dipole_update_peer(c->pol.dev(), update_opts(P, it), peer_wait_args(c), c->es.p, c->mu.p, c->mu_old.p,
                                  c->mu_new.p, c->eind.p, c->rrms.p, precision_mode ? errbits : nullptr,
                                  P.polar_palmo ? 0 : 1, c->stream);
    else
    launch_dipole_update(c->pol.dev(), update_opts(P, it), c->esum.p, 1, c->es.p, c->mu.p, c->mu_old.p, c->mu_new.p,
                         c->eind.p, c->rrms.p, precision_mode ? errbits : nullptr, 0, c->pol.n_pad, c->stream,
                         P.polar_palmo ? nullptr : c->mu.p);
    c->launches++;
    CU(cudaGetLastError());
    return 0;
}

// one Jacobi-family pass over this shard's rows: contract_dipoles + calc_dipole_rrms + the
// are_we_done_yet residual (thole_iterative.c:27-117); mu itself is not touched
static int polar_jacobi_pass(Ctx* c, int it) {
    const mpmc_b200_params& P = c->P;
    cudaStream_t st = c->stream;
    if (sym_now(c)) {
        if (polar_contract_sym(c)) return 10;
        if (c->world > 1) return 0;  // the caller all-reduces esum, then runs MPMC_STAGE_UPDATE
        return polar_update_sym(c, it);
    }
    const SubsetDev pol = c->pol.dev();
    unsigned long long* errbits = reinterpret_cast<unsigned long long*>(c->scalars.p + S_ERR);
    const bool precision_mode = P.polar_precision != 0.0;
    if (precision_mode) CU(cudaMemsetAsync(errbits, 0, sizeof(unsigned long long), st));
    kev_mark(c);
    launch_dipole_contract(make_cell(P), is_ortho(P), pol, dipole_opts(P), c->mu.p, c->pb, c->pe, c->dipole_jsplits,
                           c->epart.p, st);
    kev_mark(c);
    launch_dipole_update(pol, update_opts(P, it), c->epart.p, c->dipole_jsplits, c->es.p, c->mu.p, c->mu_old.p,
                         c->mu_new.p, c->eind.p, c->rrms.p, precision_mode ? errbits : nullptr, c->pb, c->pe, st,
                         P.polar_palmo ? nullptr : c->mu.p);
    c->launches += 2;
    c->cnt.contractions += 1;
    CU(cudaGetLastError());
    return 0;
}

// palmo rows of the non-polarizable atoms (write-back only, see dipole_field_at_kernel); `src`/`mu`
// are the polarizable atoms and their current dipoles in any common order
static int palmo_nonpolar_rows(Ctx* c, const SubsetDev& src, const double* mu) {
    if (c->nonpol.n <= 0) return 0;
    int tb, te;
    shard_range(c->nonpol.n_pad / TILE, c->rank, c->world, tb, te);
    CU(cudaMemsetAsync(c->eic_np.p, 0, sizeof(double) * 3 * c->nonpol.n_pad, c->stream));
    if (te > tb) {
        const int js = dipole_field_at_splits(te - tb, src.n_pad / TILE);
        if (c->eic_np_part.ensure((size_t)js * 3 * c->nonpol.n_pad)) FAIL(11, "cudaMalloc failed");
        launch_dipole_field_at(make_cell(c->P), is_ortho(c->P), c->nonpol.dev(), src, dipole_opts(c->P), mu, tb * TILE,
                               te * TILE, c->eic_np.p, c->stream, c->eic_np_part.p);
        c->launches += 2;
    }
    c->eic_np_valid = true;
    CU(cudaGetLastError());
    return 0;
}

// palmo_contraction (thole_iterative.c:119-141) with the CURRENT mu, this shard's rows
static int polar_palmo_finish_sym(Ctx* c) {
    if (peer_active(c))
        launch_palmo_finish_peer(c->pol.dev(), peer_wait_args(c), c->eind.p, c->eic.p, c->stream);
    else
    launch_palmo_finish(c->pol.dev(), c->esum.p, 1, c->eind.p, c->eic.p, 0, c->pol.n_pad, c->stream);
    c->launches++;
    CU(cudaGetLastError());
    return palmo_nonpolar_rows(c, c->pol.dev(), c->mu.p);
}

static int polar_palmo_pass(Ctx* c) {
    const mpmc_b200_params& P = c->P;
    const SubsetDev pol = c->pol.dev();
    if (sym_now(c)) {
        if (polar_contract_sym(c)) return 10;
        if (c->world > 1) return 0;  // all-reduce of esum, then MPMC_STAGE_PALMO_FINISH
        return polar_palmo_finish_sym(c);
    }
    launch_dipole_contract(make_cell(P), is_ortho(P), pol, dipole_opts(P), c->mu.p, c->pb, c->pe, c->dipole_jsplits,
                           c->epart.p, c->stream);
    launch_palmo_finish(pol, c->epart.p, c->dipole_jsplits, c->eind.p, c->eic.p, c->pb, c->pe, c->stream);
    c->launches += 2;
    c->cnt.contractions += 1;
    CU(cudaGetLastError());
    return palmo_nonpolar_rows(c, pol, c->mu.p);
}

// mu = new_mu | SOR | ESOR mix (thole_iterative.c:242-252), this shard's rows
static int polar_commit(Ctx* c, int it) {
    if (!c->P.polar_palmo) return 0;  // committed by the update kernel itself (only a palmo pass needs the old dipoles)
    const bool all = sym_now(c);  // symmetric K4: every rank updates every row (no dipole all-gather)
    launch_dipole_commit(c->pol.dev(), update_opts(c->P, it), c->mu_old.p, c->mu_new.p, c->mu.p, all ? 0 : c->pb,
                         all ? c->pol.n_pad : c->pe, c->stream);
    c->launches++;
    CU(cudaGetLastError());
    return 0;
}

// U_pol partial (polar.c:108-116) and rrms partial (polar.c:13-28) over this shard's rows
static int polar_energy_partial(Ctx* c) {
    launch_polar_energy(c->pol.dev(), c->mu.p, c->es.p, c->P.polar_palmo ? c->eic.p : nullptr, c->rrms.p, c->pb,
                        std::min(c->pe, c->pol.n), c->scalars.p + S_POLAR, c->stream);
    c->launches++;
    CU(cudaGetLastError());
    return 0;
}


// ---- ewald_full(): src/polarization/polar_ewald.c:375-418 on ONE device.  c->es / c->mu hold E_s and
// mu0 = alpha E_s (polar_init).  The pass counter is 0-based and are_we_done_yet sees it as is, the rrms is
// taken before new_dipoles -- both as in the reference.
static int ewald_full_field(Ctx* c, const EwaldFullOpts& eo, const KVecDev& kv, int apc) {
    const mpmc_b200_params& P = c->P;
    cudaStream_t st = c->stream;
    const Cell cell = make_cell(P);
    const SubsetDev pol = c->pol.dev();
    kev_mark(c);
    int js = 1;
    if (c->k7_list_on)
        launch_induced_list_apply(pol, c->k7_offs.p, c->k7_list.p, c->mu.p, c->epart.p, st);
    else
        js = launch_induced_real_clusters(cell, is_ortho(P), pol, eo, c->cbox.p, c->mu.p, c->dipole_jsplits, c->epart.p, st);
    kev_mark(c);
    KGroupsDev kg;
    kg.ngroups = c->nkgroups;
    kg.max_count = c->kgroup_max;
    kg.groups = c->kgroups.p;
    const int chunks = launch_dipole_sf(cell, pol, kv, c->kvec.p, c->mu.p, apc, c->sf_partial.p, st, c->k7_grouped_sf ? &kg : nullptr);
    launch_recip_finish(kv, c->sf_partial.p, chunks, 0.0, c->sf.p, nullptr, 0, 0, st, pol.n, pol.n_pad, c->mu.p,
                        c->scalars.p + S_TOTMU);  // + sum(mu) for the surface term
    launch_induced_finish(cell, pol, kv, c->kvec.p, c->sf.p, eo, c->epart.p, js, c->scalars.p + S_TOTMU, c->mu.p, c->eind.p, st);
    c->launches += 4;
    c->cnt.contractions += 1;
    CU(cudaGetLastError());
    return 0;
}

static int enqueue_ewald_full(Ctx* c) {
    const mpmc_b200_params& P = c->P;
    cudaStream_t st = c->stream;
    const int pp = c->pol.n_pad;
    if (c->world > 1) FAIL(3, "polar_ewald_full is single-GPU in this round (replicas only)");
    if (c->pol.n <= 0) return 0;
    const SubsetDev pol = c->pol.dev();
    unsigned long long* errbits = reinterpret_cast<unsigned long long*>(c->scalars.p + S_ERR);
    EwaldFullOpts eo;
    eo.a = P.polar_ewald_alpha;
    eo.damp = P.polar_damp;
    KVecDev kv;
    kv.nk = c->nk;
    kv.kmax = P.ewald_kmax;
    kv.l = c->kl.p;
    kv.weight = c->kweight_field.p;
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device);
    // chunks of the dipole structure factor: the grouped kernel holds one 384..512-thread CTA per SM (register-bound), so
    // one chunk per SM; the per-k kernel two
    const bool grouped = c->k7_grouped_sf && c->nkgroups > 0 && c->kgroup_max <= 15 && ((c->nkgroups + 31) / 32) * 32 * 4 <= 512;
    const int per = grouped ? sms : 2 * sms;
    const int apc = std::max(64, round_up((c->pol.n + per - 1) / per, 64));
    const int chunks = (c->pol.n + apc - 1) / apc;
    if (c->sf_partial.ensure((size_t)std::max(chunks, 1) * 2 * std::max(c->nk, 1)) ||
        c->epart.ensure((size_t)c->dipole_jsplits * 3 * pp))
        FAIL(11, "cudaMalloc failed");
    // Pair list of the real-space term: count, size, fill (one host sync per energy(); the solver passes then stream it).
    // Above the memory cap (a fluid with cutoff = L/2 and 10^5 atoms) the tensor is re-evaluated on the fly every pass.
    c->k7_list_on = false;
    if (c->k7_list_cap_gb > 0.0 && (pp >> 5) <= 32 * CULL_MAX_BATCHES) {
        const Cell cell = make_cell(P);
        if (c->k7_cnt.ensure((size_t)4 * pp) || c->k7_offs.ensure((size_t)4 * pp + 1)) FAIL(11, "cudaMalloc failed");
        launch_induced_list_count(cell, is_ortho(P), pol, eo, c->cbox.p, c->k7_cnt.p, c->k7_offs.p, st);
        CU(cudaMemcpyAsync(c->h_scalars + 40, c->k7_offs.p + (size_t)4 * pp, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        unsigned long long total;
        memcpy(&total, c->h_scalars + 40, sizeof(total));
        const double bytes = (double)total * (double)induced_list_entry_bytes();
        c->launches += 2;
        c->cnt.pairs_induced_list = 0.0;
        if (bytes <= c->k7_list_cap_gb * 1073741824.0) {
            if (c->k7_list.ensure((size_t)std::max<unsigned long long>(total, 1) * induced_list_entry_bytes())) {
                cudaGetLastError();  // not enough memory for the list: on-the-fly kernel
            } else {
                launch_induced_list_fill(cell, is_ortho(P), pol, eo, c->cbox.p, c->k7_offs.p, c->k7_list.p, st);
                c->launches++;
                c->k7_list_on = true;
                c->cnt.pairs_induced_list = (double)total;
            }
        }
    }
    // init_dipoles_ewald (:135-148): old_mu = 0, new_mu = mu = alpha E_s
    CU(cudaMemsetAsync(c->mu_old.p, 0, sizeof(double) * 3 * pp, st));
    CU(cudaMemcpyAsync(c->mu_new.p, c->mu.p, sizeof(double) * 3 * pp, cudaMemcpyDeviceToDevice, st));
    const bool precision_mode = P.polar_precision != 0.0;
    const double allowed = P.polar_precision * P.polar_precision * kDEBYE2SKA * kDEBYE2SKA;
    UpdateOpts uo = update_opts(P, 0);
    int count = 0;
    bool keep = true;
    while (keep) {
        if (count >= kMAX_ITERATION_COUNT && precision_mode) {  // :391-394: dipoles are left as they are
            c->pending_iter_success = 1;
            break;
        }
        if (ewald_full_field(c, eo, kv, apc)) return 10;
        uo.iteration = count;
        if (precision_mode) CU(cudaMemsetAsync(errbits, 0, sizeof(unsigned long long), st));
        launch_ewald_update(pol, uo, c->es.p, c->eind.p, c->mu.p, c->mu_old.p, c->mu_new.p, c->rrms.p,
                            precision_mode ? errbits : nullptr, st);
        c->launches++;
        if (!precision_mode)
            keep = (count != P.polar_max_iter);
        else {
            CU(cudaMemcpyAsync(c->h_scalars + S_ERR, c->scalars.p + S_ERR, sizeof(double), cudaMemcpyDeviceToHost, st));
            CU(cudaStreamSynchronize(st));
            double worst;
            memcpy(&worst, c->h_scalars + S_ERR, sizeof(double));
            keep = worst > allowed;
        }
        if (P.polar_palmo && !keep) {  // ewald_palmo_contraction (:347-371)
            if (ewald_full_field(c, eo, kv, apc)) return 10;
            launch_ewald_palmo_finish(pol, c->es.p, c->eind.p, c->mu_new.p, c->eic.p, st);
            c->launches++;
        }
        count++;
    }
    c->pending_iterations = count;
    // ef_induced of the alpha == 0 atoms (write-back only): reciprocal + surface terms of the last field pass
    if (c->nonpol.n > 0 && count > 0) {
        launch_induced_finish(make_cell(P), c->nonpol.dev(), kv, c->kvec.p, c->sf.p, eo, nullptr, 0, c->scalars.p + S_TOTMU,
                              nullptr, c->eind_np.p, st);
        c->launches++;
        c->eind_np_valid = true;
    }
    CU(cudaGetLastError());
    return 0;
}

// ---- polar(): static field + solver + energy on ONE device.  Enqueue only (precision mode syncs
// per iteration).  thole_iterative(): thole_iterative.c:168-259.
static int enqueue_polar(Ctx* c) {
    const mpmc_b200_params& P = c->P;
    cudaStream_t st = c->stream;
    const Cell cell = make_cell(P);
    const bool ortho = is_ortho(P);
    const int pp = c->pol.n_pad;
    SubsetDev pol = c->pol.dev();
    unsigned long long* errbits = reinterpret_cast<unsigned long long*>(c->scalars.p + S_ERR);
    polar_prepare(c);
    if (polar_field(c)) return 10;
    phase_begin(c, PH_DIPOLE);
    if (polar_init(c)) return 10;
    const bool want_rrms = P.polar_rrms || P.polar_precision > 0;
    const DipoleOpts dop = dipole_opts(P);
    UpdateOpts uo = update_opts(P, 0);
    const bool gs = (P.polar_gs || P.polar_gs_ranked) && !P.polar_ewald_full;
    const double allowed = P.polar_precision * P.polar_precision * kDEBYE2SKA * kDEBYE2SKA;
    const int js = c->dipole_jsplits;

    if (P.polar_ewald_full) {  // polar.c:54-57 takes this branch before looking at polar_iterative / zodid
        if (enqueue_ewald_full(c)) return 10;
    } else if (!P.polar_zodid && c->pol.n > 0) {
        if (gs && c->world > 1) FAIL(3, "Gauss-Seidel polarization is single-GPU only (sequential dependency)");
        // ---- Gauss-Seidel working set (processing order = hperm) ----
        SubsetDev g = pol;
        bool permuted = false;
        std::vector<double> hrank;
        if (gs) {
            std::iota(c->hperm.begin(), c->hperm.end(), 0);
            if (P.polar_gs_ranked) {
                // rank metric (pairs.c:337-360) -> host, stable descending order (update_ranking,
                // thole_iterative.c:143-164: stable bubble sort over all atoms; non-polarizable atoms
                // are skipped by contract_dipoles so only the relative order of polarizable ones matters)
                launch_rank_metric(cell, ortho, pol, c->rmin_part.p, c->rank_packed.p, st);
                c->launches += 3;
                hrank.resize(pp);
                CU(cudaMemcpyAsync(hrank.data(), c->rank_packed.p, sizeof(double) * pp, cudaMemcpyDeviceToHost, st));
                CU(cudaStreamSynchronize(st));
            }
            if (c->gs.x.ensure(pp) || c->gs.y.ensure(pp) || c->gs.z.ensure(pp) || c->gs.q.ensure(pp) ||
                c->gs.alpha.ensure(pp) || c->gs.aux.ensure(pp) || c->gs.mol.ensure(pp) || c->gs.flags.ensure(pp) ||
                c->gs.didx.ensure(pp) || c->perm.ensure(pp) || c->g_es.ensure(3 * (size_t)pp) ||
                c->g_mu.ensure(3 * (size_t)pp) || c->g_mu_old.ensure(3 * (size_t)pp) || c->g_mu_new.ensure(3 * (size_t)pp) ||
                c->g_eind.ensure(3 * (size_t)pp) || c->g_eic.ensure(3 * (size_t)pp) || c->g_rrms.ensure(pp) ||
                c->g_delta.ensure(3 * (size_t)pp) ||
                (c->gs_delta && !c->gs_persistent && c->g_coef.ensure(gs_block_coeff_doubles(pp))) ||
                (c->gs_delta && c->gs_persistent && (c->g_tens.ensure(gs_block_tensor_doubles(pp)) || c->g_pushacc.ensure(3 * (size_t)pp) ||
                                                      c->g_ctab.ensure(gs_pair_coeff_doubles(pp)))))
                FAIL(11, "cudaMalloc failed");
            if (c->gs_delta && c->gs_persistent && getenv("MPMC_B200_GS_TRACE") && c->gs_trace.ensure((size_t)(pp / TILE) * 8 + 8))
                FAIL(11, "cudaMalloc failed");
            if (c->gs_delta && c->gs_persistent) {  // sweep flags carry an epoch and are never cleared: zero them once per allocation
                const size_t before = c->gs_flags.cap;
                if (c->gs_flags.ensure(2 * (size_t)(pp / TILE) + 2)) FAIL(11, "cudaMalloc failed");
                if (c->gs_flags.cap != before) {
                    CU(cudaMemsetAsync(c->gs_flags.p, 0, sizeof(unsigned) * c->gs_flags.cap, st));
                    c->gs_epoch = 0;
                }
            }
        }
        auto to_gs_space = [&](bool first) -> int {
            // (re)build the permuted working set from the packed-order arrays
            if (upload(c->perm, c->hperm, st)) return 11;
            const int nb = (c->pol.n + 255) / 256;
            const int n = c->pol.n;
            permute_kernel<<<nb, 256, 0, st>>>(c->pol.x.p, c->gs.x.p, c->perm.p, n, pp, 1, 0);
            permute_kernel<<<nb, 256, 0, st>>>(c->pol.y.p, c->gs.y.p, c->perm.p, n, pp, 1, 0);
            permute_kernel<<<nb, 256, 0, st>>>(c->pol.z.p, c->gs.z.p, c->perm.p, n, pp, 1, 0);
            permute_kernel<<<nb, 256, 0, st>>>(c->pol.q.p, c->gs.q.p, c->perm.p, n, pp, 1, 0);
            permute_kernel<<<nb, 256, 0, st>>>(c->pol.alpha.p, c->gs.alpha.p, c->perm.p, n, pp, 1, 0);
            permute_kernel<<<nb, 256, 0, st>>>(c->pol.aux.p, c->gs.aux.p, c->perm.p, n, pp, 1, 0);
            permute_int_kernel<<<nb, 256, 0, st>>>(c->pol.mol.p, c->gs.mol.p, c->perm.p, n);
            permute_kernel<<<nb, 256, 0, st>>>(c->es.p, c->g_es.p, c->perm.p, n, pp, 3, 0);
            permute_kernel<<<nb, 256, 0, st>>>(c->mu.p, c->g_mu.p, c->perm.p, n, pp, 3, 0);
            if (!first) {
                permute_kernel<<<nb, 256, 0, st>>>(c->mu_old.p, c->g_mu_old.p, c->perm.p, n, pp, 3, 0);
                permute_kernel<<<nb, 256, 0, st>>>(c->mu_new.p, c->g_mu_new.p, c->perm.p, n, pp, 3, 0);
            }
            c->launches += first ? 9 : 11;
            g = pol;
            g.x = c->gs.x.p;
            g.y = c->gs.y.p;
            g.z = c->gs.z.p;
            g.q = c->gs.q.p;
            g.alpha = c->gs.alpha.p;
            g.aux = c->gs.aux.p;
            g.mol = c->gs.mol.p;
            if (c->gs_delta && c->gsp.n > 0) {  // processing order <-> spatial order of the Jacobi part
                std::vector<int> inv(pp), s2g(pp), g2s(pp);
                for (int k = 0; k < pp; k++) inv[c->hperm[k]] = k;            // packed index -> processing index
                for (int k = 0; k < pp; k++) s2g[k] = inv[c->gsp_packed[k]];  // spatial index -> processing index
                for (int k = 0; k < pp; k++) g2s[s2g[k]] = k;
                if (upload(c->map_s2g, s2g, st) || upload(c->map_g2s, g2s, st)) return 11;
                CU(cudaStreamSynchronize(st));  // the uploads read from locals
                launch_cluster_boxes(cell, ortho, c->gsp.dev(), c->cbox_sp.p, st);
                c->launches++;
            }
            if (c->gs_delta) {  // cluster boxes of the atoms in processing order, for the symmetric Jacobi part of a sweep
                launch_cluster_boxes(cell, ortho, g, c->cbox.p, st);
                if (c->gs_persistent)
                    launch_gs_block_tensors(cell, ortho, g, dop, c->g_tens.p, c->g_ctab.p, st);  // tables of this order
                else
                    launch_gs_block_coeffs(cell, ortho, g, dop, c->g_coef.p, st);  // in-block tensor coefficients of this order
                c->launches += 2;
            }
            CU(cudaGetLastError());
            return 0;
        };
        auto from_gs_space = [&]() -> int {
            const int nb = (c->pol.n + 255) / 256, n = c->pol.n;
            permute_kernel<<<nb, 256, 0, st>>>(c->g_mu.p, c->mu.p, c->perm.p, n, pp, 3, 1);
            permute_kernel<<<nb, 256, 0, st>>>(c->g_mu_old.p, c->mu_old.p, c->perm.p, n, pp, 3, 1);
            permute_kernel<<<nb, 256, 0, st>>>(c->g_mu_new.p, c->mu_new.p, c->perm.p, n, pp, 3, 1);
            permute_kernel<<<nb, 256, 0, st>>>(c->g_eind.p, c->eind.p, c->perm.p, n, pp, 3, 1);
            permute_kernel<<<nb, 256, 0, st>>>(c->g_eic.p, c->eic.p, c->perm.p, n, pp, 3, 1);
            permute_kernel<<<nb, 256, 0, st>>>(c->g_rrms.p, c->rrms.p, c->perm.p, n, pp, 1, 1);
            c->launches += 6;
            CU(cudaGetLastError());
            return 0;
        };
        // Gauss-Seidel, Jacobi part of a sweep / Palmo contraction: esum (processing order) = - sum_{j != i} T_ij mu_j for the
        // dipoles in g_mu -- evaluated on the spatially ordered image when there is one, gathered back by index maps
        auto gs_jacobi = [&]() -> int {
            if (c->gsp.n > 0) {
                const int nb = (pp + 255) / 256;
                permute_kernel<<<nb, 256, 0, st>>>(c->g_mu.p, c->g_mu_sp.p, c->map_s2g.p, pp, pp, 3, 0);
                launch_dipole_contract_sym(cell, ortho, c->gsp.dev(), dop, c->g_mu_sp.p, c->cbox_sp.p, 0, 1, c->sym_ysplits,
                                           c->sym_partI.p, c->sym_partJ.p, c->g_esum_sp.p, st);
                permute_kernel<<<nb, 256, 0, st>>>(c->g_esum_sp.p, c->esum.p, c->map_g2s.p, pp, pp, 3, 0);
                c->launches += 4;
            } else {
                launch_dipole_contract_sym(cell, ortho, g, dop, c->g_mu.p, c->cbox.p, 0, 1, c->sym_ysplits, c->sym_partI.p,
                                           c->sym_partJ.p, c->esum.p, st);
                c->launches += 2;
            }
            CU(cudaGetLastError());
            return 0;
        };
        if (gs) {
            CU(cudaMemsetAsync(c->g_mu_old.p, 0, sizeof(double) * 3 * pp, st));
            CU(cudaMemsetAsync(c->g_mu_new.p, 0, sizeof(double) * 3 * pp, st));
            CU(cudaMemsetAsync(c->g_eind.p, 0, sizeof(double) * 3 * pp, st));
            CU(cudaMemsetAsync(c->g_eic.p, 0, sizeof(double) * 3 * pp, st));
            CU(cudaMemsetAsync(c->g_rrms.p, 0, sizeof(double) * pp, st));
            if (to_gs_space(true)) return 10;
            permuted = true;
        }

        // Small systems, Jacobi family: every iteration, the convergence test, the bail-out and the Palmo pass in ONE
        // cooperative launch (dipole_fused_solver_kernel); the iteration count comes back with the scalars.
        bool fused = false;
        {
            static const int fused_max = [] { const char* v = getenv("MPMC_B200_FUSED_MAX"); return v ? atoi(v) : 4096; }();
            if (!gs && c->world == 1 && !sym_now(c) && pp <= fused_max && c->epart.cap >= (size_t)((c->pol.n + TILE - 1) / TILE) * 3 * pp) {
                uo.iteration = 0;
                fused = launch_dipole_fused_solver(cell, ortho, pol, dop, uo, P.polar_max_iter, allowed, P.polar_precision != 0.0,
                                                   P.polar_palmo != 0, c->es.p, c->mu.p, c->mu_old.p, c->mu_new.p, c->eind.p,
                                                   c->eic.p, c->rrms.p, c->epart.p, P.polar_palmo ? c->esum.p : nullptr,
                                                   reinterpret_cast<unsigned long long*>(c->scalars.p + S_FUSED_ERR),
                                                   c->scalars.p + S_FUSED_OUT, st);
                if (fused) {
                    c->launches++;
                    c->iter_from_device = true;
                    if (P.polar_palmo && palmo_nonpolar_rows(c, pol, c->esum.p)) return 10;
                }
            }
        }
        int it = 0;
        bool keep = !fused;
        while (keep) {
            it++;
            if (it >= kMAX_ITERATION_COUNT && P.polar_precision != 0.0) {  // divergence bail-out (:199-210)
                if (gs)
                    reset_dipoles_kernel<<<(c->pol.n + 255) / 256, 256, 0, st>>>(c->pol.n, pp, g.alpha, c->g_es.p, c->g_mu.p,
                                                                                c->g_eic.p);
                else
                    reset_dipoles_kernel<<<(c->pol.n + 255) / 256, 256, 0, st>>>(c->pol.n, pp, pol.alpha, c->es.p, c->mu.p,
                                                                                c->eic.p);
                c->launches++;
                c->pending_iter_success = 1;
                break;
            }
            uo.iteration = it;
            const bool precision_mode = P.polar_precision != 0.0;
            if (precision_mode) CU(cudaMemsetAsync(errbits, 0, sizeof(unsigned long long), st));
            if (gs) {
                if (uo.save_old) CU(cudaMemcpyAsync(c->g_mu_old.p, c->g_mu.p, sizeof(double) * 3 * pp, cudaMemcpyDeviceToDevice, st));
                if (c->gs_delta) {
                    kev_mark(c);
                    bool one_launch = false;
                    if (c->gs_persistent) {
                        // Jacobi part with the dipoles as they are at the start of the sweep, then the whole
                        // lower-triangular part in one cooperative launch (resolver + pushers with look-ahead)
                        if (gs_jacobi()) return 10;
                        one_launch = launch_gs_sweep_persistent(cell, ortho, g, dop, c->esum.p, c->g_es.p, c->g_mu.p, c->g_mu_new.p,
                                                                c->g_eind.p, c->g_delta.p, c->g_tens.p, c->g_ctab.p, c->g_pushacc.p,
                                                                c->gs_flags.p, ++c->gs_epoch, c->scalars.p + S_GS_TIMEOUT, st,
                                                                c->gs_trace.p);
                        c->launches += 1;
                    }
                    if (!one_launch) {
                        launch_gs_sweep_delta(cell, ortho, g, dop, c->cbox.p, c->sym_ysplits, c->sym_partI.p, c->sym_partJ.p,
                                              c->esum.p, c->g_es.p, c->g_mu.p, c->g_mu_new.p, c->g_eind.p, c->g_delta.p,
                                              c->gs_persistent ? nullptr : c->g_coef.p, st);
                        c->launches += 2 + 2 * ((c->pol.n + TILE - 1) / TILE);
                    }
                    kev_mark(c);
                } else {
                    launch_gs_sweep(cell, ortho, g, dop, nullptr, c->g_es.p, c->g_mu.p, c->g_mu_new.p, c->g_eind.p, c->epart.p, st);
                    c->launches += 2 * (pp / TILE);
                }
                if (want_rrms || precision_mode) {
                    gs_check_kernel<<<(c->pol.n + 255) / 256, 256, 0, st>>>(c->pol.n, pp, c->g_mu_new.p, c->g_mu_old.p,
                                                                          uo.do_rrms, c->g_rrms.p, errbits);
                    c->launches++;
                }
            } else {
                if (polar_jacobi_pass(c, it)) return 10;
            }
            if (gs) c->cnt.contractions += 1;
            CU(cudaGetLastError());

            if (!precision_mode)
                keep = (it != P.polar_max_iter);
            else {
                CU(cudaMemcpyAsync(c->h_scalars + S_ERR, c->scalars.p + S_ERR, sizeof(double), cudaMemcpyDeviceToHost, st));
                CU(cudaStreamSynchronize(st));
                double worst;
                memcpy(&worst, c->h_scalars + S_ERR, sizeof(double));
                keep = worst > allowed;
            }

            if (P.polar_palmo && !keep) {  // palmo_contraction (:119-141) with the CURRENT mu
                if (gs && c->gs_delta) {
                    if (gs_jacobi()) return 10;
                    launch_palmo_finish(g, c->esum.p, 1, c->g_eind.p, c->g_eic.p, 0, pp, st);
                    c->launches += 1;
                    c->cnt.contractions += 1;
                    if (palmo_nonpolar_rows(c, g, c->g_mu.p)) return 10;
                } else if (gs) {
                    launch_dipole_contract(cell, ortho, g, dop, c->g_mu.p, 0, pp, js, c->epart.p, st);
                    launch_palmo_finish(g, c->epart.p, js, c->g_eind.p, c->g_eic.p, 0, pp, st);
                    c->launches += 2;
                    c->cnt.contractions += 1;
                    if (palmo_nonpolar_rows(c, g, c->g_mu.p)) return 10;
                } else if (polar_palmo_pass(c))
                    return 10;
            }

            if (gs && P.polar_gs_ranked && keep && it == 1) {
                // update_ranking (:143-164): from the second sweep on, atoms are processed in stable
                // descending rank order (the metric is constant within one energy() call)
                if (from_gs_space()) return 10;
                std::vector<int> order(c->pol.n);
                std::iota(order.begin(), order.end(), 0);
                std::stable_sort(order.begin(), order.end(), [&](int u, int v) { return hrank[u] > hrank[v]; });
                std::copy(order.begin(), order.end(), c->hperm.begin());
                if (to_gs_space(false)) return 10;
            }

            // save the dipoles for the next pass (:242-252)
            if (gs) {
                if (uo.sor || uo.esor) {
                    launch_dipole_commit(g, uo, c->g_mu_old.p, c->g_mu_new.p, c->g_mu.p, 0, pp, st);
                    c->launches++;
                }  // else mu already equals new_mu
            } else if (polar_commit(c, it))
                return 10;
        }
        c->pending_iterations = it;
        if (permuted)
            if (from_gs_space()) return 10;
    }
    if (warm_start_applies(c) && c->pol.n > 0 && !c->pending_iter_success) {  // keep the converged dipoles (atom space)
        if (c->mu_keep.ensure((size_t)3 * c->n_pad)) FAIL(11, "cudaMalloc failed");
        launch_dipole_keep(c->pol.dev(), c->mu.p, c->n_pad, c->mu_keep.p, st);
        c->launches++;
        c->warm_valid = true;
    } else
        c->warm_valid = false;
    phase_end(c, PH_DIPOLE);

    return polar_energy_partial(c);
}

static int enqueue_energy(Ctx* c, bool with_pair_part, bool with_polar_part) {
    if (!c->have_params) FAIL(2, "mpmc_b200_set_params has not been called");
    if (!c->have_atoms) FAIL(2, "mpmc_b200_set_atoms has not been called");
    if (check_supported(c->P)) return 3;
    const mpmc_b200_params& P = c->P;
    if (c->world > 1 && P.cavity_autoreject_absolute)  // a minimum over ALL pairs decides before anything else is summed
        FAIL(3, "cavity_autoreject_absolute is evaluated on one GPU only (sharded contexts refuse it)");
    cudaStream_t st = c->stream;
    CU(cudaSetDevice(c->device));
    c->scalars_complete = false;
    c->inc_valid = false;  // S(k) and the sums are rewritten: mpmc_b200_energy() re-arms the incremental path afterwards
    if (c->kdirty)
        if (build_kvectors(c)) return 10;
    if (c->analytic_dirty) compute_analytic(c);
    refresh_subsets(c);
    c->boxes_fresh = false;
    c->launches = 0;
    for (int p = 0; p < PH_COUNT; p++) c->ev_used[p] = false;
    c->kev_used = 0;
    memset(&c->cnt, 0, sizeof(c->cnt));
    c->cnt.n_atoms = c->n;
    c->pending_iterations = 0;
    c->pending_iter_success = 0;
    c->iter_from_device = false;
    CU(cudaMemsetAsync(c->scalars.p, 0, sizeof(double) * 32, st));
    phase_begin(c, PH_TOTAL);
    const Cell cell = make_cell(P);
    const bool ortho = is_ortho(P);

    if (with_pair_part) {
        // fork: with polarization on (and nothing that needs the pair result on the host first), K1 + K2 run on
        // the side stream while the static field and the dipole solve proceed on the engine's stream
        static const bool overlap_env = [] { const char* v = getenv("MPMC_B200_OVERLAP"); return !(v && v[0] == '0'); }();
        const bool will_polar = P.polarization && !P.rd_only && (with_polar_part || c->world > 1);
        const bool overlap = overlap_env && will_polar && !P.cavity_autoreject_absolute && !P.polar_ewald &&
                             !P.polar_ewald_full && c->use_clusters && c->side != nullptr;
        cudaStream_t ps = st;
        if (c->use_clusters) {
            // cluster boxes of the sorted image (this evaluation's positions): read by K1 and by K3
            const SubsetDev sd = c->sorted.dev();
            launch_cluster_boxes(cell, ortho, sd, sd.cbox, st);
            c->boxes_fresh = true;
            c->launches++;
        }
        if (overlap) {
            CU(cudaEventRecord(c->ev_fork, st));
            CU(cudaStreamWaitEvent(c->side, c->ev_fork, 0));
            ps = c->side;
        }
        phase_begin(c, PH_PAIR, ps);
        PairOpts po;
        po.ewald_alpha = P.ewald_alpha;
        po.temperature = P.temperature;
        po.fh_order = P.feynman_hibbs ? (P.feynman_hibbs_order == 2 ? 2 : 4) : 0;
        po.mixing = P.mixing;
        po.rd_only = P.rd_only;
        po.es_wolf = P.es_wolf;
        po.no_lj = P.rd_crystal ? 1 : 0;
        if (c->use_clusters) {
            // K1 over this shard's i-clusters
            const SubsetDev sd = c->sorted.dev();
            int cb, ce;
            shard_range(c->sorted.n_pad / 32, c->rank, c->world, cb, ce);
            const double cull = std::max(P.cutoff, P.cavity_autoreject_absolute ? P.cavity_autoreject_scale : 0.0) + 1e-6;
            launch_pair_energy_clusters(cell, ortho, sd, po, cb, ce, c->pair_kslices, cull, c->pair_part.p, c->pair_stride, ps);
            const int cols = pair_cluster_columns(c->sorted.n_pad, cb, ce, c->pair_kslices);
            launch_pair_reduce(c->pair_part.p, std::max(cols, 0), c->pair_stride, c->scalars.p + S_PAIR, ps);
            // intramolecular erf term (coulombic.c:176-183), list-order rows of this shard; overwrites slot 2
            int rb, re;
            shard_range(c->n_pad / TILE, c->rank, c->world, rb, re);
            rb *= TILE;
            re = std::min(re * TILE, c->n);
            if (!P.rd_only && !P.es_wolf) {  // coulombic_wolf() has no intramolecular term
                launch_intra_erf(atoms_dev(c), c->follow.p, P.ewald_alpha, rb, std::max(re, rb), c->intra_part.p,
                                 c->scalars.p + S_PAIR + 2, ps);
                c->launches += 2;
            }
            c->launches += 2;
        } else {
            int tb, te;
            shard_range(c->ntiles, c->rank, c->world, tb, te);
            if (te > tb) {
                launch_pair_energy(cell, ortho, atoms_dev(c), po, c->tiles.p + tb, te - tb, c->pair_part.p, c->pair_stride, ps);
                c->launches++;
            }
            launch_pair_reduce(c->pair_part.p, std::max(te - tb, 0), c->pair_stride, c->scalars.p + S_PAIR, ps);
            c->launches++;
        }
        if (P.rd_crystal) {  // LJ image sums over this shard's tile pairs (lj.c:196-216)
            if (c->crystal_img_dirty) {
                if (upload(c->crystal_img, c->h_crystal_img, ps)) return 11;
                c->crystal_img_dirty = false;
            }
            int tb, te;
            shard_range(c->crystal_ntiles, c->rank, c->world, tb, te);
            launch_crystal_lj(cell, ortho, atoms_dev(c), po, rd_cutoff(P), c->crystal_img.p, c->crystal_nimg,
                              c->crystal_tiles.p + tb, te - tb, c->crystal_part.p, c->crystal_stride,
                              c->scalars.p + S_CRYSTAL, ps);
            c->launches += (te > tb) ? 2 : 1;
        }
        phase_end(c, PH_PAIR, ps);
        CU(cudaGetLastError());

        if (P.cavity_autoreject_absolute) {  // energy.c:99-102: a bad contact skips every other term
            CU(cudaMemcpyAsync(c->h_scalars, c->scalars.p, sizeof(double) * 8, cudaMemcpyDeviceToHost, st));
            CU(cudaStreamSynchronize(st));
            if (c->h_scalars[S_PAIR + 3] < P.cavity_autoreject_scale && c->world == 1) {
                c->pending = true;
                c->last_had_polar = false;
                phase_end(c, PH_TOTAL);
                return 0;
            }
        }
        if (!P.rd_only && !P.es_wolf) {  // coulombic.c:27-28: wolf replaces real + reciprocal + self
            phase_begin(c, PH_RECIP, ps);
            c->cnt.kvectors = c->nk;
            c->cnt.recip_atoms = c->recip.n;
            if (run_structure_factor(c, c->recip, c->kweight, 4.0 * M_PI / P.volume, c->scalars.p + S_RECIP, true, ps)) return 10;
            phase_end(c, PH_RECIP, ps);
        }
        if (overlap) {
            CU(cudaEventRecord(c->ev_join, c->side));
            c->side_active = true;
        }
    }
    c->last_had_polar = false;
    if (with_polar_part && P.polarization && !P.rd_only) {
        if (enqueue_polar(c)) return 10;
        c->last_had_polar = true;
    }
    if (with_polar_part || !(P.polarization && !P.rd_only)) join_side(c);  // sharded runs join in STAGE_ENERGY
    phase_end(c, PH_TOTAL);
    c->pending = true;
    return 0;
}

// The 32 device scalars of one evaluation (+ the host-side analytic terms) -> what energy() reports.  Shared by
// finish_energy() and the resident MC steps, which keep one scalar set per step.
static int assemble_energies(Ctx* c, const double* h, mpmc_b200_energies* out) {
    const mpmc_b200_params& P = c->P;
    if (h[S_PEER_TIMEOUT] != 0.0) FAIL(12, "peer exchange timed out: a rank never published its induced-field partial sums");
    if (h[S_GS_TIMEOUT] != 0.0) FAIL(12, "Gauss-Seidel sweep timed out: a CTA of the persistent sweep waited 2 s for another");
    memset(out, 0, sizeof(*out));
    // host-side analytic terms are added once: by rank 0 when the caller sums the ranks' partial energies (staged
    // protocol), by every rank when the device scalars were already combined over the ranks (sharded_eval)
    const bool lead = (c->rank == 0) || c->scalars_complete;
    out->min_intermolecular_r = h[S_PAIR + 3];
    c->cnt.pairs_tested = c->use_clusters ? (lead ? c->pairs_tested_analytic : 0.0) : h[S_PAIR + 4];
    c->cnt.pairs_lj = h[S_PAIR + 5] + (P.rd_crystal ? h[S_CRYSTAL + 1] : 0.0);
    c->cnt.pairs_es = h[S_PAIR + 6];
    c->cnt.kernel_launches = c->launches;
    c->cnt.pairs_field = 0.5 * h[S_FIELDCNT];
    if (P.cavity_autoreject_absolute && c->world == 1 && h[S_PAIR + 3] < P.cavity_autoreject_scale) {
        out->autorejected = 1;
        out->total = kMAXVALUE;
        return 0;
    }
    out->rd_pair = h[S_PAIR + 0];
    if (P.rd_crystal) out->rd_pair += h[S_CRYSTAL] + (lead ? c->crystal_self : 0.0);  // lj.c:196-216, 264-267
    out->rd_lrc = lead ? (c->lrc_pair + c->lrc_self) : 0.0;
    out->rd = (out->rd_pair + (lead ? c->lrc_pair : 0.0)) + (lead ? c->lrc_self : 0.0);
    if (!P.rd_only) {
        out->es_real = h[S_PAIR + 1] - h[S_PAIR + 2];
        out->es_recip = P.es_wolf ? 0.0 : h[S_RECIP];
        out->es_self = (lead && !P.es_wolf) ? c->es_self : 0.0;
        out->coulombic = out->es_real + out->es_recip + out->es_self;
    }
    if (c->last_had_polar && c->iter_from_device) {
        c->pending_iterations = (int)h[S_FUSED_OUT];
        c->pending_iter_success = (int)h[S_FUSED_OUT + 1];
        if (c->pending_iter_success) c->warm_valid = false;  // the bail-out left alpha E_s, not converged dipoles
        c->cnt.contractions = c->pending_iterations - (c->pending_iter_success ? 1 : 0) + (P.polar_palmo && !c->pending_iter_success ? 1 : 0);
    }
    if (c->last_had_polar) {
        double pot = h[S_POLAR + 0];
        if (P.polar_palmo) pot += h[S_POLAR + 1];
        out->polar = -0.5 * pot;
        out->dipole_rrms = c->n ? h[S_POLAR + 2] / c->n : 0.0;
        out->polar_iterations = c->pending_iterations;
        out->iter_success = c->pending_iter_success;
        if (P.polar_ewald_full) {  // polar.c:54-57: this branch reports neither the pass count nor the rrms
            out->polar_iterations = 0;
            out->dipole_rrms = 0.0;
        }
    }
    out->total = out->rd + out->coulombic + out->polar;  // energy.c:194
    return 0;
}

static int finish_energy(Ctx* c, mpmc_b200_energies* out) {
    if (!c->pending) FAIL(2, "no evaluation pending");
    const mpmc_b200_params& P = c->P;
    cudaStream_t st = c->stream;
    join_side(c);
    CU(cudaMemcpyAsync(c->h_scalars, c->scalars.p, sizeof(double) * 32, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    c->pending = false;
    for (int p = 0; p < PH_COUNT; p++) {
        c->phase_ms[p] = 0.f;
        if (c->ev_used[p]) cudaEventElapsedTime(&c->phase_ms[p], c->ev[p][0], c->ev[p][1]);
    }
    c->contract_ms_sum = 0.f;
    c->contract_launches = c->kev_used / 2;
    for (int k = 0; k + 1 < c->kev_used; k += 2) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, c->kev[k], c->kev[k + 1]);
        c->contract_ms_sum += ms;
    }
    const double* h = c->h_scalars;
    if (c->gs_trace.p && c->last_had_polar && (c->P.polar_gs || c->P.polar_gs_ranked) && c->pol.n > 0) {
        // profiling aid: where the resolver of the last persistent sweep spent its time, per 128-atom block
        const int nb = (c->pol.n + TILE - 1) / TILE;
        std::vector<unsigned long long> t((size_t)nb * 8);
        if (cudaMemcpy(t.data(), c->gs_trace.p, sizeof(unsigned long long) * t.size(), cudaMemcpyDeviceToHost) == cudaSuccess && nb > 4) {
            double acc[5] = {0, 0, 0, 0, 0};
            for (int b = 2; b < nb - 1; b++)
                for (int k = 0; k < 5; k++) acc[k] += (double)(t[(size_t)b * 8 + k + 1] - t[(size_t)b * 8 + k]);
            fprintf(stderr, "MPMC_B200_GS_TRACE: %d blocks, sweep %.3f ms; per block (us): wait for pushers %.2f, load %.2f, "
                            "chains + cross pushes %.2f, publish %.2f, push to next block %.2f\n",
                    nb, (double)(t[(size_t)(nb - 1) * 8 + 5] - t[0]) * 1e-6, acc[0] / (nb - 3) * 1e-3, acc[1] / (nb - 3) * 1e-3,
                    acc[2] / (nb - 3) * 1e-3, acc[3] / (nb - 3) * 1e-3, acc[4] / (nb - 3) * 1e-3);
        }
    }
    return assemble_energies(c, h, out);
}

// ================================================================================================
// Sharded evaluation inside the library: `nl` contexts of THIS process, each one rank of the same
// world (nl == world: one process drives every GPU of the box, mpmc_b200_create_group; nl == 1: one process per GPU,
// the ranks' exported buffers mapped through CUDA IPC).  thole_iterative() (thole_iterative.c:168-259) cut where partial
// results have to be combined, with every exchange done by the engine's own kernels over NVLink peer memory:
//   static field   K3 leaves this rank's rows in its exported buffer      -> peer_sum_kernel        (all-gather)
//   every pass     K4 leaves its contribution to E_ind of all rows there  -> dipole_update_peer     (all-reduce + update)
//   the end        the 32 device scalars                                  -> peer_scalars_kernel    (all-reduce)
// No NCCL call, no host round trip between stages (tolerance mode reads one scalar per pass, as on one GPU).  Stages are
// issued rank by rank from one host thread; a kernel that waits for a peer's flag is only ever enqueued after every
// allocation of the evaluation has been made (cudaMalloc may synchronise devices that map each other's memory).
// ================================================================================================
static int sharded_enqueue(Ctx** cs, int nl, bool with_pair) {
    Ctx* c0 = cs[0];
    const mpmc_b200_params& P = c0->P;
    for (int i = 0; i < nl; i++) {
        Ctx* c = cs[i];
        if (!c->have_params || !c->have_atoms) FAIL(2, "no system");
        if (!peer_active(c)) FAIL(2, "sharded evaluation needs the peer exchange (mpmc_b200_peer_export / _connect after set_atoms)");
        if (c->world != c0->world || c->n != c0->n || c->pol.n_pad != c0->pol.n_pad) FAIL(2, "ranks hold different systems");
    }
    const bool polar = P.polarization && !P.rd_only;
    if (polar && (P.polar_gs || P.polar_gs_ranked) && !P.polar_ewald_full)
        FAIL(3, "Gauss-Seidel polarization is single-GPU only (sequential dependency)");
    if (polar && P.polar_ewald_full) FAIL(3, "polar_ewald_full is single-GPU in this round (replicas only)");
    // run `fn` for every local rank with its device current; the first non-zero return code ends the evaluation
    auto each = [&](auto&& fn) -> int {
        for (int i = 0; i < nl; i++) {
            Ctx* c = cs[i];
            if (cudaSetDevice(c->device) != cudaSuccess) {
                g_err = "cudaSetDevice failed";
                return 10;
            }
            if (const int rc = fn(c)) return rc;
        }
        return 0;
    };
#define RUN(...)                                             \
    do {                                                     \
        if (const int rc__ = each(__VA_ARGS__)) return rc__; \
    } while (0)
    RUN([&](Ctx* c) -> int { return enqueue_energy(c, with_pair, false); });
    int it = 0;
    if (polar) {
        RUN([&](Ctx* c) -> int {
            polar_prepare(c);
            c->field_via_peer = true;
            const int rc = polar_field(c);
            c->field_via_peer = false;
            c->last_had_polar = true;
            return rc ? 10 : 0;
        });
        RUN([&](Ctx* c) -> int {
            phase_begin(c, PH_DIPOLE);
            launch_peer_sum(peer_wait_args(c, CH_FIELD), (size_t)3 * c->n_pad, c->ef_static.p, c->stream);
            c->launches++;
            return polar_init(c) ? 10 : 0;
        });
        if (!P.polar_zodid && c0->pol.n > 0) {
            const bool precision_mode = P.polar_precision != 0.0;
            const double allowed = P.polar_precision * P.polar_precision * kDEBYE2SKA * kDEBYE2SKA;
            bool keep = true;
            while (keep) {
                it++;
                if (it >= kMAX_ITERATION_COUNT && precision_mode) {  // divergence bail-out (:199-210), every rank alike
                    RUN([&](Ctx* c) -> int {
                        reset_dipoles_kernel<<<(c->pol.n + 255) / 256, 256, 0, c->stream>>>(c->pol.n, c->pol.n_pad, c->pol.alpha.p,
                                                                                           c->es.p, c->mu.p, c->eic.p);
                        c->launches++;
                        c->pending_iter_success = 1;
                        CU(cudaGetLastError());
                        return 0;
                    });
                    break;
                }
                RUN([&](Ctx* c) -> int { return polar_jacobi_pass(c, it) ? 10 : 0; });  // contraction + reduce + signal
                RUN([&](Ctx* c) -> int { return polar_update_sym(c, it) ? 10 : 0; });   // wait + rank-ordered sum + update
                if (!precision_mode)
                    keep = (it != P.polar_max_iter);
                else {  // every rank computed the same residual from the same sums: no exchange, one scalar per rank
                    RUN([&](Ctx* c) -> int {
                        CU(cudaMemcpyAsync(c->h_scalars + S_ERR, c->scalars.p + S_ERR, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
                        return 0;
                    });
                    RUN([&](Ctx* c) -> int {
                        CU(cudaStreamSynchronize(c->stream));
                        return 0;
                    });
                    double worst;
                    memcpy(&worst, c0->h_scalars + S_ERR, sizeof(double));
                    for (int i = 1; i < nl; i++)
                        if (memcmp(cs[i]->h_scalars + S_ERR, c0->h_scalars + S_ERR, sizeof(double)) != 0)
                            FAIL(12, "ranks disagree on the dipole residual (replicated dipoles drifted apart)");
                    keep = worst > allowed;
                }
                if (P.polar_palmo && !keep) {
                    RUN([&](Ctx* c) -> int { return polar_palmo_pass(c) ? 10 : 0; });
                    RUN([&](Ctx* c) -> int { return polar_palmo_finish_sym(c) ? 10 : 0; });
                }
                RUN([&](Ctx* c) -> int { return polar_commit(c, it) ? 10 : 0; });
            }
        }
        RUN([&](Ctx* c) -> int {
            phase_end(c, PH_DIPOLE);
            c->pending_iterations = it;
            return polar_energy_partial(c) ? 10 : 0;
        });
    }
    // the 32 device scalars, combined over the ranks by every rank (rank order => identical bits everywhere)
    RUN([&](Ctx* c) -> int {
        join_side(c);
        const PeerSignal sg = peer_signal_args(c, CH_SCAL);
        CU(cudaMemcpyAsync(peer_scal(c->peer_base[c->rank], c->peer_cap, c->peer_cap_atoms, (int)(sg.seq & 1ull)), c->scalars.p,
                           sizeof(double) * 32, cudaMemcpyDeviceToDevice, c->stream));
        launch_peer_signal(sg, c->stream);
        return 0;
    });
    RUN([&](Ctx* c) -> int {
        launch_peer_scalars(peer_wait_args(c, CH_SCAL), 32, S_PAIR + 3, c->scalars.p, c->stream);
        c->launches += 2;
        phase_end(c, PH_TOTAL);
        c->pending = true;
        c->scalars_complete = true;
        CU(cudaGetLastError());
        return 0;
    });
#undef RUN
    return 0;
}

static int sharded_eval(Ctx** cs, int nl, bool with_pair, mpmc_b200_energies* out) {
    Ctx* c0 = cs[0];
    // The ranks of a device group are issued by one host thread each -- except for the first evaluation after the
    // layout or the options changed: that one may still allocate (cudaMalloc can synchronise devices that map each
    // other's memory, and a rank blocked there while a peer's kernel waits for its flag would dead-lock), so it goes
    // stage by stage from this thread, where every allocation precedes every waiting kernel.
    bool parallel = nl > 1 && (int)c0->workers.size() == nl - 1;
    for (int i = 0; i < nl; i++) parallel = parallel && !cs[i]->layout_dirty;
    int rc_all = 0;
    if (parallel) {
        std::vector<int> rcs(nl, 0);
        std::vector<std::string> msgs(nl);
        for (int i = 1; i < nl; i++)
            c0->workers[i - 1]->submit([&, i] {
                Ctx* one = cs[i];
                rcs[i] = sharded_enqueue(&one, 1, with_pair);
                if (rcs[i]) msgs[i] = g_err;  // g_err is per thread
            });
        rcs[0] = sharded_enqueue(cs, 1, with_pair);
        if (rcs[0]) msgs[0] = g_err;
        for (int i = 1; i < nl; i++) c0->workers[i - 1]->wait();
        for (int i = 0; i < nl; i++)
            if (rcs[i] && !rc_all) {
                rc_all = rcs[i];
                g_err = msgs[i];
            }
    } else
        rc_all = sharded_enqueue(cs, nl, with_pair);
    if (rc_all) {  // leave no evaluation half enqueued: wait for what was launched (peer waits time out on their own)
        const std::string keep = g_err;
        for (int i = 0; i < nl; i++) {
            cudaSetDevice(cs[i]->device);
            cudaStreamSynchronize(cs[i]->stream);
            if (cs[i]->side) cudaStreamSynchronize(cs[i]->side);
            cs[i]->side_active = false;
            cs[i]->pending = false;
        }
        cudaGetLastError();
        cudaSetDevice(c0->device);
        g_err = keep;
        return rc_all;
    }
    for (int i = 0; i < nl; i++) {
        Ctx* c = cs[i];
        CU(cudaSetDevice(c->device));
        mpmc_b200_energies e;
        const int rc = finish_energy(c, &e);
        c->scalars_complete = false;
        if (rc && !rc_all) rc_all = rc;
        if (i == 0 && !rc) *out = e;
        if (!rc) c->layout_dirty = false;
    }
    CU(cudaSetDevice(c0->device));
    return rc_all;
}

// every rank's context of a group led by `c` (or `c` alone)
static int local_ranks(Ctx* c, Ctx** cs) {
    if (c->group.empty()) {
        cs[0] = c;
        return 1;
    }
    for (size_t i = 0; i < c->group.size(); i++) cs[i] = c->group[i];
    return (int)c->group.size();
}

// One process, several GPUs: every context allocates its exchange buffer, peer access is enabled pairwise and the
// buffers are wired directly (no IPC).  Called after set_atoms / set_params of the whole group when a size changed.
static int group_connect(Ctx* lead) {
    const int w = (int)lead->group.size();
    bool ok = w > 1;
    for (Ctx* c : lead->group) ok = ok && peer_active(c);
    if (ok || w <= 1) return 0;
    for (Ctx* c : lead->group) {
        CU(cudaSetDevice(c->device));
        CU(cudaStreamSynchronize(c->stream));
        if (c->side) CU(cudaStreamSynchronize(c->side));
    }
    for (Ctx* c : lead->group) {
        CU(cudaSetDevice(c->device));
        peer_teardown(c);
        if (!c->use_sym) continue;
        c->peer_cap = (size_t)c->pol.n_pad;
        c->peer_cap_atoms = (size_t)c->n_pad;
        const size_t bytes = peer_bytes(c->peer_cap, c->peer_cap_atoms);
        CU(cudaMalloc(&c->peer_local, bytes));
        CU(cudaMemset(c->peer_local, 0, bytes));
        for (Ctx* o : lead->group) {
            if (o == c) continue;
            cudaError_t e = cudaDeviceEnablePeerAccess(o->device, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled)
                cudaGetLastError();
            else if (e != cudaSuccess) {
                cudaGetLastError();
                FAIL(13, "cudaDeviceEnablePeerAccess failed: the GPUs of a group must reach each other over NVLink / PCIe P2P");
            }
        }
    }
    for (Ctx* c : lead->group) {
        if (!c->use_sym) continue;
        for (Ctx* o : lead->group) c->peer_base[o->rank] = o->peer_local;
        memset(c->peer_seq, 0, sizeof(c->peer_seq));
        c->peer_ipc = false;
        c->peer_ready = true;
    }
    CU(cudaSetDevice(lead->device));
    return 0;
}

// ================================================================================================
// C ABI
// ================================================================================================
extern "C" {

int mpmc_b200_abi_version(void) { return MPMC_B200_ABI_VERSION; }
const char* mpmc_b200_last_error(void) { return g_err.c_str(); }

int mpmc_b200_create(mpmc_b200_ctx** out, int device) {
    if (!out) FAIL(1, "null ctx pointer");
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev <= 0) FAIL(20, "no CUDA device available (this library has no CPU fallback)");
    if (device < 0 || device >= ndev) FAIL(20, "CUDA device ordinal out of range");
    CU(cudaSetDevice(device));
    Ctx* c = new Ctx();
    c->device = device;
    CU(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    c->own_stream = true;
    {   // A/B switch for profiling: MPMC_B200_K4=rows selects the row-by-row contraction of dipole_kernel.cu
        const char* v = getenv("MPMC_B200_K4");
        c->use_sym = !(v && strcmp(v, "rows") == 0);
        // MPMC_B200_CUTOFF=tiles selects the all-pairs tile form of K1 / K3 (no cluster culling)
        v = getenv("MPMC_B200_CUTOFF");
        c->use_clusters = !(v && strcmp(v, "tiles") == 0);
        v = getenv("MPMC_B200_WARM_START");
        c->warm_enabled = v && v[0] == '1';
        v = getenv("MPMC_B200_INCREMENTAL");
        c->inc_enabled = v && v[0] == '1';
        v = getenv("MPMC_B200_K7_SF");
        c->k7_grouped_sf = !(v && strcmp(v, "plain") == 0);
        v = getenv("MPMC_B200_K7_LIST_GB");
        if (v) c->k7_list_cap_gb = atof(v);
        v = getenv("MPMC_B200_GS");
        c->gs_delta = c->use_sym && !(v && strcmp(v, "panels") == 0);
        c->gs_persistent = !(v && strcmp(v, "blocks") == 0);
        v = getenv("MPMC_B200_GS_JACOBI");
        c->gs_spatial = !(v && strcmp(v, "order") == 0);
    }
    for (int p = 0; p < PH_COUNT; p++) {
        CU(cudaEventCreate(&c->ev[p][0]));
        CU(cudaEventCreate(&c->ev[p][1]));
    }
    {   // lowest priority: the static field + dipole solve on the engine's stream are the critical path, K1 / K2 fill in
        int lo = 0, hi = 0;
        cudaDeviceGetStreamPriorityRange(&lo, &hi);
        CU(cudaStreamCreateWithPriority(&c->side, cudaStreamNonBlocking, lo));
    }
    CU(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming));
    if (c->scalars.ensure(64)) FAIL(11, "cudaMalloc failed");
    CU(cudaMallocHost(&c->h_scalars, sizeof(double) * 64));
    *out = c;
    return 0;
}

// One process driving `ndev` GPUs of one NVLink / NVSwitch box: rank i of the sharded evaluation lives
// on devices[i]; the returned context is rank 0 and leads the group -- set_params / set_atoms / set_positions / energy /
// polar / get_atom_arrays on it address the whole group, so a caller (the drop-in shim) needs no other change.
int mpmc_b200_create_group(mpmc_b200_ctx** out, int ndev, const int* devices) {
    if (!out) FAIL(1, "null ctx pointer");
    *out = nullptr;
    if (ndev < 1 || ndev > PEER_MAX) FAIL(1, "a group has 1..8 devices");
    if (ndev == 1) return mpmc_b200_create(out, devices ? devices[0] : 0);
    for (int i = 0; i < ndev; i++)
        for (int j = 0; j < i; j++)
            if ((devices ? devices[i] : i) == (devices ? devices[j] : j)) FAIL(1, "a device appears twice in the group");
    std::vector<mpmc_b200_ctx*> g;
    for (int i = 0; i < ndev; i++) {
        mpmc_b200_ctx* m = nullptr;
        const int rc = mpmc_b200_create(&m, devices ? devices[i] : i);
        if (rc) {
            const std::string keep = g_err;
            for (mpmc_b200_ctx* x : g) mpmc_b200_destroy(x);
            g_err = keep;
            return rc;
        }
        m->rank = i;
        m->world = ndev;
        m->group_member = i > 0;
        g.push_back(m);
    }
    for (int i = 0; i < ndev; i++)
        for (int j = 0; j < ndev; j++) {
            int can = 1;
            if (i != j) cudaDeviceCanAccessPeer(&can, g[i]->device, g[j]->device);
            if (!can) {
                for (mpmc_b200_ctx* x : g) {
                    x->group_member = false;
                    mpmc_b200_destroy(x);
                }
                FAIL(13, "the GPUs of a group must have peer access to each other (NVLink / NVSwitch)");
            }
        }
    g[0]->group = g;
    {
        const char* v = getenv("MPMC_B200_GROUP_THREADS");  // 0: every rank issued from the caller's thread, stage by stage
        if (!(v && v[0] == '0'))
            for (int i = 1; i < ndev; i++) {
                RankWorker* w = new RankWorker();
                w->th = std::thread([w] { w->loop(); });
                g[0]->workers.push_back(w);
            }
    }
    cudaSetDevice(g[0]->device);
    *out = g[0];
    return 0;
}

int mpmc_b200_group_size(mpmc_b200_ctx* c) { return !c ? 0 : c->group.empty() ? 1 : (int)c->group.size(); }

void mpmc_b200_destroy(mpmc_b200_ctx* c) {
    if (!c) return;
    for (RankWorker* w : c->workers) {
        w->stop();
        delete w;
    }
    c->workers.clear();
    if (!c->group.empty()) {  // a group leader owns the other ranks' contexts
        std::vector<mpmc_b200_ctx*> members(c->group.begin() + 1, c->group.end());
        c->group.clear();
        for (mpmc_b200_ctx* m : members) {
            m->group_member = false;
            mpmc_b200_destroy(m);
        }
    }
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    if (c->side) {
        cudaStreamSynchronize(c->side);
        cudaStreamDestroy(c->side);
    }
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    if (c->ev_join) cudaEventDestroy(c->ev_join);
    for (int p = 0; p < PH_COUNT; p++) {
        cudaEventDestroy(c->ev[p][0]);
        cudaEventDestroy(c->ev[p][1]);
    }
    for (cudaEvent_t e : c->kev) cudaEventDestroy(e);
    for (DevBuf<double>* b : {&c->x, &c->y, &c->z, &c->q, &c->eps, &c->sig, &c->molmass, &c->alpha, &c->pair_part, &c->kweight,
                              &c->kweight_field, &c->kvec, &c->sf_partial, &c->sf, &c->ef_static, &c->epart, &c->es, &c->mu,
                              &c->mu_old, &c->mu_new, &c->eind, &c->eic, &c->rrms, &c->rank_packed, &c->rmin_part,
                              &c->field_cnt, &c->g_es, &c->g_mu, &c->g_mu_old, &c->g_mu_new, &c->g_eind, &c->g_eic, &c->g_rrms, &c->g_delta, &c->g_coef,
                              &c->g_tens, &c->g_pushacc, &c->g_ctab,
                              &c->wolf_table, &c->scalars, &c->cbox, &c->sym_partI, &c->sym_partJ, &c->esum})
        b->release();
    for (DevBuf<int>* b : {&c->mol, &c->flags, &c->kl, &c->perm}) b->release();
    c->kgroups.release();
    c->gs_flags.release();
    c->gs_trace.release();
    c->k7_cnt.release();
    c->k7_offs.release();
    c->k7_list.release();
    c->tiles.release();
    c->charged.release();
    c->recip.release();
    c->pol.release();
    c->nonpol.release();
    c->eic_np.release();
    c->eic_np_part.release();
    c->eind_np.release();
    c->wb.release();
    c->gs.release();
    c->gsp.release();
    c->map_s2g.release();
    c->map_g2s.release();
    c->g_mu_sp.release();
    c->g_esum_sp.release();
    c->cbox_sp.release();
    c->sorted.release();
    c->follow.release();
    c->intra_part.release();
    if (c->h_scalars) cudaFreeHost(c->h_scalars);
    if (c->h_stage) cudaFreeHost(c->h_stage);
    if (c->h_wb) cudaFreeHost(c->h_wb);
    if (c->h_mc_res) cudaFreeHost(c->h_mc_res);
    if (c->h_mc_ring) cudaFreeHost(c->h_mc_ring);
    peer_teardown(c);
    if (c->own_stream) cudaStreamDestroy(c->stream);
    delete c;
}

int mpmc_b200_set_stream(mpmc_b200_ctx* c, void* cuda_stream) {
    if (!c) FAIL(1, "null ctx");
    CU(cudaStreamSynchronize(c->stream));
    if (c->side) CU(cudaStreamSynchronize(c->side));
    c->side_active = false;
    if (c->own_stream) cudaStreamDestroy(c->stream);
    if (cuda_stream) {
        c->stream = (cudaStream_t)cuda_stream;
        c->own_stream = false;
    } else {
        CU(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
        c->own_stream = true;
    }
    return 0;
}

int mpmc_b200_set_params(mpmc_b200_ctx* c, const mpmc_b200_params* p) {
    if (!c || !p) FAIL(1, "null argument");
    if (c->group.size() > 1) {  // the same options on every rank of the group
        if (check_supported(*p)) return 3;
        std::vector<mpmc_b200_ctx*> g;
        g.swap(c->group);
        int rc = 0;
        for (mpmc_b200_ctx* m : g)
            if (!rc) rc = mpmc_b200_set_params(m, p);
        c->group.swap(g);
        if (!rc && c->have_atoms) rc = group_connect(c);
        return rc;
    }
    const bool repack_needed = c->have_atoms && (p->damp_type != c->P.damp_type || p->polar_gs != c->P.polar_gs ||
                                                  p->polar_gs_ranked != c->P.polar_gs_ranked ||
                                                  p->polar_ewald_full != c->P.polar_ewald_full ||
                                                  (p->rd_crystal != 0) != (c->P.rd_crystal != 0));
    if (check_supported(*p)) return 3;  // refused parameters leave the context as it was
    c->inc_valid = false;
    c->warm_valid = false;
    c->layout_dirty = true;
    if (repack_needed || !c->have_params || memcmp(p->basis, c->P.basis, sizeof(p->basis)) != 0) c->orders_valid = false;
    c->P = *p;
    c->have_params = true;
    c->kdirty = true;
    c->analytic_dirty = true;
    if (repack_needed) {
        CU(cudaSetDevice(c->device));
        if (repack(c)) return 10;
    }
    return 0;
}

int mpmc_b200_set_atoms(mpmc_b200_ctx* c, int n, const double* pos, const double* charge, const double* polarizability,
                        const double* epsilon, const double* sigma, const double* mol_mass, const int* mol_id,
                        const int* frozen) {
    if (!c) FAIL(1, "null ctx");
    if (!c->have_params) FAIL(2, "call mpmc_b200_set_params before mpmc_b200_set_atoms");
    if (n < 0 || (n > 0 && (!pos || !charge || !polarizability || !epsilon || !sigma || !mol_mass || !mol_id || !frozen)))
        FAIL(1, "bad atom arrays");
    if (n > CULL_MAX_BATCHES * 32 * 32 - 4096) FAIL(3, "more than ~2M atoms: raise CULL_MAX_BATCHES (common.cuh)");
    if (c->group.size() > 1) {  // every rank holds the whole system (the work is sharded, not the atoms)
        std::vector<mpmc_b200_ctx*> g;
        g.swap(c->group);
        int rc = 0;
        for (mpmc_b200_ctx* m : g)
            if (!rc) rc = mpmc_b200_set_atoms(m, n, pos, charge, polarizability, epsilon, sigma, mol_mass, mol_id, frozen);
        c->group.swap(g);
        if (!rc) rc = group_connect(c);
        return rc;
    }
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    // a small edit of the atom set (uVT insert / remove / restore): common prefix and suffix with what is held
    c->edit_map.clear();
    c->drifted.clear();
    c->drift_tracked = false;
    {
        static const bool reuse = [] { const char* v = getenv("MPMC_B200_REUSE_ORDER"); return !(v && v[0] == '0'); }();
        const int no = c->n;
        if (reuse && c->have_atoms && c->orders_valid && !c->resort_pending && no >= 1024 && n >= 1024) {
            auto same = [&](int io, int in) {
                return c->hx[io] == pos[3 * in] && c->hy[io] == pos[3 * in + 1] && c->hz[io] == pos[3 * in + 2] &&
                       c->hq[io] == charge[in] && c->halpha[io] == polarizability[in] && c->heps[io] == epsilon[in] &&
                       c->hsig[io] == sigma[in] && c->hfrozen[io] == (frozen[in] ? 1 : 0);
            };
            // Two-pointer walk over the held and the incoming arrays.  At a mismatch the shortest block (<= 64 atoms: a
            // molecule) whose removal from the held side, insertion on the incoming side, or replacement makes the next atoms
            // agree again is taken as one edit.  A uVT step is usually TWO edits at different places of the list -- the
            // driver's restore of a rejected insertion / removal never reaches the engine on its own, it arrives together
            // with the next trial -- which a common prefix + suffix would read as one long changed span (every second
            // step of the PCN-61 sample then sorted afresh).
            auto agree = [&](int io, int in) {  // the next (up to) 8 atoms agree; both lists ending together counts
                const int len = std::min(8, std::min(no - io, n - in));
                if (len == 0) return io == no && in == n;
                for (int q = 0; q < len; q++)
                    if (!same(io + q, in + q)) return false;
                return true;
            };
            std::vector<int> map(no, -1), fresh_list;
            int io = 0, in = 0, gone = 0;
            bool ok = true;
            while (ok && (io < no || in < n)) {
                if (io < no && in < n && same(io, in)) {
                    map[io++] = in++;
                    continue;
                }
                bool found = false;
                for (int k = 1; k <= 64 && !found; k++) {
                    if (io + k <= no && agree(io + k, in)) {  // k held atoms are gone
                        io += k, gone += k, found = true;
                    } else if (in + k <= n && agree(io, in + k)) {  // k new atoms
                        for (int q = 0; q < k; q++) fresh_list.push_back(in + q);
                        in += k, found = true;
                    } else if (io + k <= no && in + k <= n && agree(io + k, in + k)) {  // k atoms replaced
                        for (int q = 0; q < k; q++) fresh_list.push_back(in + q);
                        io += k, in += k, gone += k, found = true;
                    }
                }
                if (!found || gone + (int)fresh_list.size() > 256) ok = false;
            }
            const int fresh = (int)fresh_list.size();
            // every edit shifts the clusters behind it off the bricks the sort made (a few % of K4 / K1 / K3 once many have
            // piled up): a full sort again after n/128 edited atoms -- its cost, spread over them, is ~0.15 ms per edit
            // (small systems: a sort costs ~2 ms at 8k atoms against an evaluation of ~1.3 ms, so up to 256 edited atoms)
            if (ok && gone + fresh <= 256 && c->order_edits + gone + fresh <= std::max(std::min(256, no / 16), no / 128)) {
                c->edit_map.swap(map);
                c->edit_fresh.swap(fresh_list);
                c->order_edits += gone + fresh;
            }
        }
    }
    c->n = n;
    c->hx.resize(n), c->hy.resize(n), c->hz.resize(n), c->hq.resize(n), c->halpha.resize(n), c->heps.resize(n);
    c->hsig.resize(n), c->hmm.resize(n), c->hmol.resize(n), c->hfrozen.resize(n);
    for (int i = 0; i < n; i++) {
        c->hx[i] = pos[3 * i];
        c->hy[i] = pos[3 * i + 1];
        c->hz[i] = pos[3 * i + 2];
        c->hq[i] = charge[i];
        c->halpha[i] = polarizability[i];
        c->heps[i] = epsilon[i];
        c->hsig[i] = sigma[i];
        c->hmm[i] = mol_mass[i];
        c->hmol[i] = mol_id[i];
        c->hfrozen[i] = frozen[i] ? 1 : 0;
    }
    if (repack(c)) return 10;
    CU(cudaStreamSynchronize(c->stream));  // uploads read from temporaries
    return 0;
}

int mpmc_b200_set_positions(mpmc_b200_ctx* c, int first, int count, const double* pos) {
    if (!c || !pos) FAIL(1, "null argument");
    if (!c->have_atoms) FAIL(2, "mpmc_b200_set_atoms has not been called");
    if (first < 0 || count < 0 || first + count > c->n) FAIL(1, "position range out of bounds");
    if (count == 0) return 0;
    if (c->group.size() > 1) {
        std::vector<mpmc_b200_ctx*> g;
        g.swap(c->group);
        int rc = 0;
        for (mpmc_b200_ctx* m : g)
            if (!rc) rc = mpmc_b200_set_positions(m, first, count, pos);
        c->group.swap(g);
        CU(cudaSetDevice(c->device));
        return rc;
    }
    CU(cudaSetDevice(c->device));
    if (c->h_stage_cap < (size_t)3 * count) {
        if (c->h_stage) {
            CU(cudaStreamSynchronize(c->stream));
            cudaFreeHost(c->h_stage);
        }
        c->h_stage_cap = (size_t)3 * count + 1024;
        CU(cudaMallocHost(&c->h_stage, sizeof(double) * c->h_stage_cap));
    } else {
        CU(cudaStreamSynchronize(c->stream));  // staging buffer reuse
    }
    double *sx = c->h_stage, *sy = sx + count, *sz = sy + count;
    if (c->inc_valid) {  // remember which atoms really move, and where they were at the last evaluation
        for (int k = 0; k < count && c->inc_valid; k++) {
            const int g = first + k;
            if (pos[3 * k] == c->hx[g] && pos[3 * k + 1] == c->hy[g] && pos[3 * k + 2] == c->hz[g]) continue;
            if (!c->inc_slot[g]) {
                if ((int)c->inc_moved.size() >= INC_MAX_MOVED) {
                    c->inc_valid = false;  // too many: the next evaluation is a full one
                    break;
                }
                MovedRec r;
                r.xo = c->hx[g], r.yo = c->hy[g], r.zo = c->hz[g];
                r.q = c->hq[g], r.eps = c->heps[g], r.sig = c->hsig[g], r.mm = c->hmm[g];
                r.mol = c->hmol[g], r.fl = F_VALID | (c->hfrozen[g] ? F_FROZEN : 0), r.idx = g, r.pad = 0;
                c->inc_moved.push_back(r);
                c->inc_slot[g] = (int)c->inc_moved.size();
            }
            MovedRec& r = c->inc_moved[c->inc_slot[g] - 1];
            r.xn = pos[3 * k], r.yn = pos[3 * k + 1], r.zn = pos[3 * k + 2];
        }
        if (!c->inc_valid) {
            for (const MovedRec& r : c->inc_moved) c->inc_slot[r.idx] = 0;
            c->inc_moved.clear();
        }
    }
    for (int k = 0; k < count; k++) {
        sx[k] = c->hx[first + k] = pos[3 * k];
        sy[k] = c->hy[first + k] = pos[3 * k + 1];
        sz[k] = c->hz[first + k] = pos[3 * k + 2];
    }
    if ((c->pol_sorted || c->use_clusters) && c->hx_sort.size() == c->hx.size() && (!c->resort_pending || c->drift_tracked)) {
        // clusters stay compact as long as no atom has drifted far from where it was when the orders were made; the ones
        // that have are named, so that refresh_subsets() can relocate them instead of sorting everything again
        if (!c->resort_pending) {
            c->drifted.clear();
            c->drift_tracked = true;
        }
        const double lim2 = 1.5 * 1.5;
        for (int k = 0; k < count; k++) {
            const int g = first + k;
            const double dx = c->hx[g] - c->hx_sort[g], dy = c->hy[g] - c->hy_sort[g], dz = c->hz[g] - c->hz_sort[g];
            if (dx * dx + dy * dy + dz * dz > lim2) {
                c->resort_pending = true;
                if (c->drifted.size() <= 512) c->drifted.push_back(g);
            }
        }
    }
    CU(cudaMemcpyAsync(c->x.p + first, sx, sizeof(double) * count, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(c->y.p + first, sy, sizeof(double) * count, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(c->z.p + first, sz, sizeof(double) * count, cudaMemcpyHostToDevice, c->stream));
    c->subsets_stale = true;
    c->boxes_fresh = false;
    return 0;
}

int mpmc_b200_set_shard(mpmc_b200_ctx* c, int rank, int world) {
    if (!c) FAIL(1, "null ctx");
    if (world < 1 || rank < 0 || rank >= world) FAIL(1, "bad shard");
    if (!c->group.empty() || c->group_member) FAIL(2, "the ranks of a group are fixed by mpmc_b200_create_group");
    const bool changed = (world != c->world);
    if (changed || rank != c->rank) peer_teardown(c);  // the exchange buffers belong to the old shard layout
    c->rank = rank;
    c->world = world;
    if (changed && c->have_atoms) {
        CU(cudaSetDevice(c->device));
        if (repack(c)) return 10;
        CU(cudaStreamSynchronize(c->stream));
    }
    return 0;
}

// ---- incremental evaluation -----------------------------------------------------------------------------
// Applies to what changes with a few moved atoms and nothing else: the K1 pair sums and S(k).  Polarization (every dipole
// changes), the cavity minimum distance (not decomposable) and the rd_crystal image sums keep the full evaluation.
static bool inc_supported(const Ctx* c) {
    const mpmc_b200_params& P = c->P;
    return c->inc_enabled && c->world == 1 && !(P.polarization && !P.rd_only) && !P.cavity_autoreject_absolute &&
           !P.rd_crystal;
}

static void inc_clear_moved(Ctx* c) {
    for (const MovedRec& r : c->inc_moved) c->inc_slot[r.idx] = 0;
    c->inc_moved.clear();
}

static int incremental_energy(Ctx* c, mpmc_b200_energies* out) {
    const mpmc_b200_params& P = c->P;
    cudaStream_t st = c->stream;
    CU(cudaSetDevice(c->device));
    // an atom that moved and came back (a rejected move restored bit for bit) has nothing to add
    std::vector<MovedRec> mv;
    for (const MovedRec& r : c->inc_moved)
        if (!(r.xn == r.xo && r.yn == r.yo && r.zn == r.zo)) mv.push_back(r);
    c->launches = 0;
    *out = c->inc_base;
    c->cnt = c->inc_cnt;
    if (!mv.empty()) {
        const int nb = c->n_pad / TILE;
        const int stride = round_up(nb, 32);
        if (c->inc_moved_dev.ensure(INC_MAX_MOVED) || c->inc_part.ensure((size_t)INC_NPART * stride) || c->inc_out.ensure(8))
            FAIL(11, "cudaMalloc failed");
        CU(cudaMemcpyAsync(c->inc_moved_dev.p, mv.data(), sizeof(MovedRec) * mv.size(), cudaMemcpyHostToDevice, st));
        PairOpts po;
        po.ewald_alpha = P.ewald_alpha;
        po.temperature = P.temperature;
        po.fh_order = P.feynman_hibbs ? (P.feynman_hibbs_order == 2 ? 2 : 4) : 0;
        po.mixing = P.mixing;
        po.rd_only = P.rd_only;
        po.es_wolf = P.es_wolf;
        const Cell cell = make_cell(P);
        launch_pair_delta(cell, is_ortho(P), atoms_dev(c), po, c->inc_moved_dev.p, (int)mv.size(), c->inc_part.p, stride,
                          c->inc_out.p, st);
        c->launches += 2;
        const bool recip = !P.rd_only && !P.es_wolf;
        if (recip) {
            KVecDev kv;
            kv.nk = c->nk;
            kv.kmax = P.ewald_kmax;
            kv.l = c->kl.p;
            kv.weight = c->kweight.p;
            launch_recip_delta(kv, c->kvec.p, c->inc_moved_dev.p, (int)mv.size(), c->sf.p, 4.0 * M_PI / P.volume,
                               c->inc_out.p + INC_NPART, st);
            c->launches += 2;
        }
        CU(cudaGetLastError());
        CU(cudaMemcpyAsync(c->h_scalars, c->inc_out.p, sizeof(double) * 8, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        const double* h = c->h_scalars;
        out->rd_pair += h[0];
        out->rd = (out->rd_pair + c->lrc_pair) + c->lrc_self;
        if (!P.rd_only) {
            out->es_real += h[1] - h[2];
            if (recip) out->es_recip = h[INC_NPART];
            out->coulombic = out->es_real + out->es_recip + out->es_self;
        }
        out->total = out->rd + out->coulombic + out->polar;
        c->cnt.pairs_lj += h[3];
        c->cnt.pairs_es += h[4];
    }
    out->min_intermolecular_r = 0.0;  // not maintained by the incremental path (cavity_autoreject disables it)
    c->cnt.kernel_launches = c->launches;
    c->cnt.incremental = 1;
    c->inc_base = *out;
    c->inc_cnt = c->cnt;
    inc_clear_moved(c);
    c->inc_since_full++;
    return 0;
}

int mpmc_b200_energy(mpmc_b200_ctx* c, mpmc_b200_energies* out) {
    if (!c || !out) FAIL(1, "null argument");
    if (c->world > 1) {  // sharded: the whole staged evaluation inside the library, exchanges over peer memory
        Ctx* cs[PEER_MAX];
        const int nl = local_ranks(c, cs);
        if (nl > 1 || peer_active(c)) return sharded_eval(cs, nl, true, out);
        FAIL(2, "sharded context without the peer exchange: call mpmc_b200_peer_export / _connect, or drive the stages "
                "with mpmc_b200_energy_enqueue / mpmc_b200_polar_stage and your own collectives");
    }
    static const int refresh = [] { const char* v = getenv("MPMC_B200_INCREMENTAL_REFRESH"); return v ? atoi(v) : 256; }();
    if (c->inc_valid && inc_supported(c) && !c->kdirty && !c->analytic_dirty && c->inc_since_full < refresh) {
        const int rc = incremental_energy(c, out);
        if (rc) c->inc_valid = false;
        return rc;
    }
    c->inc_valid = false;
    inc_clear_moved(c);
    int rc = enqueue_energy(c, true, true);
    if (rc) return rc;
    rc = finish_energy(c, out);
    if (!rc && inc_supported(c) && !out->autorejected) {  // this evaluation is the base of the next incremental ones
        c->inc_base = *out;
        c->inc_cnt = c->cnt;
        c->inc_since_full = 0;
        if ((int)c->inc_slot.size() != c->n) c->inc_slot.assign(c->n, 0);
        c->inc_valid = true;
    }
    return rc;
}

int mpmc_b200_set_warm_start(mpmc_b200_ctx* c, int on) {
    if (!c) FAIL(1, "null ctx");
    c->warm_enabled = on != 0;
    c->warm_valid = false;
    return 0;
}

int mpmc_b200_set_incremental(mpmc_b200_ctx* c, int on) {
    if (!c) FAIL(1, "null ctx");
    c->inc_enabled = on != 0;
    c->inc_valid = false;
    inc_clear_moved(c);
    return 0;
}

int mpmc_b200_polar(mpmc_b200_ctx* c, mpmc_b200_energies* out) {
    if (!c || !out) FAIL(1, "null argument");
    int rc;
    if (c->world > 1) {
        Ctx* cs[PEER_MAX];
        const int nl = local_ranks(c, cs);
        if (!(nl > 1 || peer_active(c))) FAIL(2, "sharded context without the peer exchange (see mpmc_b200_energy)");
        rc = sharded_eval(cs, nl, false, out);
    } else {
        rc = enqueue_energy(c, false, true);
        if (rc) return rc;
        rc = finish_energy(c, out);
    }
    if (!rc) {
        out->rd = out->rd_pair = out->rd_lrc = out->coulombic = out->es_real = out->es_recip = out->es_self = 0.0;
        out->total = out->polar;
    }
    return rc;
}

int mpmc_b200_energy_enqueue(mpmc_b200_ctx* c) {
    if (!c) FAIL(1, "null ctx");
    if (c->world > 1) return enqueue_energy(c, true, false);  // polar part is staged by the caller
    return enqueue_energy(c, true, true);
}

int mpmc_b200_energy_finish(mpmc_b200_ctx* c, mpmc_b200_energies* out) {
    if (!c || !out) FAIL(1, "null argument");
    return finish_energy(c, out);
}

// ================================================================================================
// Resident MC steps: a batch of displacement moves evaluated and decided on the device.  Per step, on the
// engine's stream and without any host synchronisation: mc_move_kernel (translate + rotate of one molecule in the
// list-order coordinates, old coordinates kept), the whole evaluation (enqueue_energy: the subsets and cluster boxes are
// refreshed from the list-order coordinates), the write-back image of that evaluation into the step's ring slot, and
// mc_decide_kernel (total, Metropolis test, coordinates put back on a reject).  One D2H of the per-step records and the
// ring at the end.  The spatial order is left alone during a batch (it is a matter of speed, not of results) and
// re-examined afterwards from the final coordinates.
// ================================================================================================
static int enqueue_writeback(mpmc_b200_ctx* c, double* dst, int kb, int ke);
int mpmc_b200_mc_steps(mpmc_b200_ctx* c, int nsteps, const mpmc_b200_mc_move* moves, double temperature,
                       double current_energy, mpmc_b200_mc_result* results) {
    if (!c || !moves || !results) FAIL(1, "null argument");
    if (!c->have_params || !c->have_atoms) FAIL(2, "no system");
    if (nsteps < 1 || nsteps > MPMC_B200_MC_MAX_STEPS) FAIL(1, "mpmc_b200_mc_steps: nsteps outside 1..MPMC_B200_MC_MAX_STEPS");
    if (c->world > 1 || c->group.size() > 1) FAIL(3, "mpmc_b200_mc_steps: single GPU only");
    const mpmc_b200_params& P = c->P;
    const bool polar = P.polarization && !P.rd_only;
    if (P.cavity_autoreject_absolute) FAIL(3, "mpmc_b200_mc_steps: cavity_autoreject_absolute needs the host between the pair sums and the rest");
    if (polar && P.polar_precision != 0.0 && !P.polar_zodid) FAIL(3, "mpmc_b200_mc_steps: fixed-iteration solvers only (polar_precision reads a scalar per pass)");
    if (!(temperature > 0.0)) FAIL(1, "mpmc_b200_mc_steps: temperature must be positive");
    for (int j = 0; j < nsteps; j++)
        if (moves[j].count < 1 || moves[j].count > MPMC_B200_MC_MAX_ATOMS || moves[j].first < 0 || moves[j].first + moves[j].count > c->n)
            FAIL(1, "mpmc_b200_mc_steps: a move's atom range is empty, longer than MPMC_B200_MC_MAX_ATOMS or out of bounds");
    CU(cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    const size_t ring_per_step = polar ? (size_t)14 * c->n : 0;
    if (c->mc_moves.ensure(MPMC_B200_MC_MAX_STEPS) || c->mc_res.ensure(MPMC_B200_MC_MAX_STEPS) || c->mc_ecur.ensure(1) ||
        (ring_per_step && c->mc_ring.ensure(ring_per_step * nsteps)))
        FAIL(11, "cudaMalloc failed");
    if (c->h_mc_res_cap < (size_t)MPMC_B200_MC_MAX_STEPS) {
        c->h_mc_res_cap = MPMC_B200_MC_MAX_STEPS;
        CU(cudaMallocHost(&c->h_mc_res, sizeof(McStepDev) * c->h_mc_res_cap));
    }
    if (c->h_mc_ring_cap < ring_per_step * nsteps) {
        if (c->h_mc_ring) cudaFreeHost(c->h_mc_ring);
        c->h_mc_ring = nullptr;
        c->h_mc_ring_cap = ring_per_step * nsteps + ring_per_step;
        CU(cudaMallocHost(&c->h_mc_ring, sizeof(double) * c->h_mc_ring_cap));
    }
    c->mc_ring_valid = false;
    c->inc_valid = false;
    inc_clear_moved(c);
    c->warm_valid = false;
    CU(cudaMemcpyAsync(c->mc_moves.p, moves, sizeof(mpmc_b200_mc_move) * nsteps, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(c->mc_ecur.p, &current_energy, sizeof(double), cudaMemcpyHostToDevice, st));
    if (c->kdirty)
        if (build_kvectors(c)) return 10;
    if (c->analytic_dirty) compute_analytic(c);
    McDecide k;
    k.lrc_pair = c->lrc_pair, k.lrc_self = c->lrc_self, k.es_self = c->es_self, k.crystal_self = c->crystal_self;
    k.temperature = temperature;
    k.rd_only = P.rd_only, k.es_wolf = P.es_wolf, k.rd_crystal = P.rd_crystal, k.polar = polar ? 1 : 0, k.palmo = P.polar_palmo;
    k.s_pair = S_PAIR, k.s_recip = S_RECIP, k.s_polar = S_POLAR, k.s_crystal = S_CRYSTAL, k.s_fused_out = S_FUSED_OUT;
    int launches = 0;
    if (c->mc_ctr.ensure(1) || (ring_per_step && c->wb.ensure(ring_per_step))) FAIL(11, "cudaMalloc failed");
    CU(cudaMemsetAsync(c->mc_ctr.p, 0, sizeof(int), st));
    // one step: j >= 0 names it, j < 0 takes the step the device counter points at
    auto step = [&](int j) -> int {
        launch_mc_move(c->mc_moves.p, j, c->mc_ctr.p, c->x.p, c->y.p, c->z.p, c->mc_res.p, st);
        c->subsets_stale = true;
        c->boxes_fresh = false;
        const bool resort = c->resort_pending;
        if (j != 0) c->resort_pending = false;  // the host mirror of the coordinates is stale inside a batch
        const int rc = enqueue_energy(c, true, true);
        c->resort_pending = c->resort_pending || (j != 0 && resort);
        if (rc) return rc;
        if (ring_per_step) {
            if (enqueue_writeback(c, c->wb.p, 0, c->pol.n)) return 10;
            launch_mc_ring_copy(c->wb.p, c->mc_ring.p, ring_per_step, j, c->mc_ctr.p, st);
        }
        k.iter_from_device = c->iter_from_device ? 1 : 0;
        k.iter_success = c->pending_iter_success;
        launch_mc_decide(c->mc_moves.p, j, c->mc_ctr.p, k, c->scalars.p, c->mc_ecur.p, c->x.p, c->y.p, c->z.p, c->mc_res.p, st);
        launches += c->launches + 2 + (ring_per_step ? 2 : 0);
        return 0;
    };
    // Steps 0 and 1 are enqueued launch by launch (step 0 may still re-sort and allocate; step 1 runs every launcher once in
    // its steady state); from step 2 on a CUDA graph captured from one more plain enqueue of a step is replayed: the ~40
    // launches of a small system's step become one graph launch, and the GPU no longer idles between tiny kernels.
    // Gauss-Seidel is left out (its persistent sweep is handed a fresh epoch by the host at every launch).
    if (c->mc_graph < 0) {
        const char* v = getenv("MPMC_B200_MC_GRAPH");
        c->mc_graph = (v && v[0] == '0') ? 0 : 1;
    }
    const bool want_graph = c->mc_graph == 1 && nsteps >= 4 && !(polar && (P.polar_gs || P.polar_gs_ranked || P.polar_ewald_full));
    int j = 0;
    for (; j < nsteps && (!want_graph || j < 2); j++)
        if (const int rc = step(j)) return rc;
    if (j < nsteps) {
        cudaGraph_t graph = nullptr;
        cudaGraphExec_t exec = nullptr;
        bool ok = cudaStreamBeginCapture(st, cudaStreamCaptureModeRelaxed) == cudaSuccess;
        int rc_step = 0;
        if (ok) {
            const int before = launches;
            rc_step = step(-1);
            const int per_graph = launches - before;
            ok = cudaStreamEndCapture(st, &graph) == cudaSuccess && graph && rc_step == 0;
            if (ok) ok = cudaGraphInstantiate(&exec, graph, 0) == cudaSuccess;
            launches = before;
            for (; ok && j < nsteps; j++) {
                ok = cudaGraphLaunch(exec, st) == cudaSuccess;
                launches += per_graph;
            }
        }
        if (exec) cudaGraphExecDestroy(exec);
        if (graph) cudaGraphDestroy(graph);
        if (!ok) {  // something in the step does not capture on this driver: this context stays on plain launches
            cudaGetLastError();
            c->mc_graph = 0;
            // the device counter says how far the batch got; finish it launch by launch
            CU(cudaStreamSynchronize(st));
            int done = 0;
            CU(cudaMemcpy(&done, c->mc_ctr.p, sizeof(int), cudaMemcpyDeviceToHost));
            for (j = done; j < nsteps; j++)
                if (const int rc = step(j)) return rc;
        }
    }
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(c->h_mc_res, c->mc_res.p, sizeof(McStepDev) * nsteps, cudaMemcpyDeviceToHost, st));
    if (ring_per_step)
        CU(cudaMemcpyAsync(c->h_mc_ring, c->mc_ring.p, sizeof(double) * ring_per_step * nsteps, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    c->pending = false;
    c->subsets_stale = true;  // the last decision may have put coordinates back after the subsets were gathered
    c->boxes_fresh = false;
    const int pend_it = c->pending_iterations, pend_ok = c->pending_iter_success;
    for (int j = 0; j < nsteps; j++) {
        const McStepDev& r = c->h_mc_res[j];
        c->pending_iterations = pend_it;
        c->pending_iter_success = pend_ok;
        if (const int rc = assemble_energies(c, r.scalars, &results[j].e)) return rc;
        results[j].accepted = r.accepted != 0.0 ? 1 : 0;
        results[j].reserved = 0;
        memcpy(results[j].new_pos, r.new_pos, sizeof(r.new_pos));
        memcpy(results[j].old_pos, r.old_pos, sizeof(r.old_pos));
        // the host mirror follows the decisions
        const double* fin = results[j].accepted ? r.new_pos : r.old_pos;
        for (int i = 0; i < moves[j].count; i++) {
            const int g = moves[j].first + i;
            c->hx[g] = fin[3 * i], c->hy[g] = fin[3 * i + 1], c->hz[g] = fin[3 * i + 2];
        }
    }
    c->cnt.kernel_launches = launches;
    if ((c->pol_sorted || c->use_clusters) && !c->resort_pending && c->hx_sort.size() == c->hx.size()) {
        const double lim2 = 1.5 * 1.5;  // as mpmc_b200_set_positions: clusters stay compact while no atom drifted far
        c->drifted.clear();
        c->drift_tracked = true;
        for (int j = 0; j < nsteps; j++)
            for (int i = 0; i < moves[j].count; i++) {
                const int g = moves[j].first + i;
                const double dx = c->hx[g] - c->hx_sort[g], dy = c->hy[g] - c->hy_sort[g], dz = c->hz[g] - c->hz_sort[g];
                if (dx * dx + dy * dy + dz * dz > lim2) {
                    c->resort_pending = true;
                    if (c->drifted.size() <= 512) c->drifted.push_back(g);
                }
            }
    }
    c->mc_last_steps = nsteps;
    c->mc_ring_valid = ring_per_step != 0;
    return 0;
}

int mpmc_b200_mc_step_arrays_view(mpmc_b200_ctx* c, int step, const double** mu, const double** ef_static,
                                  const double** ef_induced, const double** ef_induced_change, const double** dipole_rrms,
                                  const double** rank_metric) {
    if (!c) FAIL(1, "null ctx");
    if (!c->mc_ring_valid || step < 0 || step >= c->mc_last_steps) FAIL(2, "no per-atom results of that step (mpmc_b200_mc_steps with polarization on)");
    const size_t n = (size_t)c->n;
    const double* h = c->h_mc_ring + (size_t)14 * n * step;
    if (mu) *mu = h;
    if (ef_static) *ef_static = h + 3 * n;
    if (ef_induced) *ef_induced = h + 6 * n;
    if (ef_induced_change) *ef_induced_change = h + 9 * n;
    if (dipole_rrms) *dipole_rrms = h + 12 * n;
    if (rank_metric) *rank_metric = h + 13 * n;
    return 0;
}

int mpmc_b200_polar_stage(mpmc_b200_ctx* c, int stage, int iteration) {
    if (!c) FAIL(1, "null ctx");
    if (!c->have_params || !c->have_atoms) FAIL(2, "no system");
    const mpmc_b200_params& P = c->P;
    if (!P.polarization || P.rd_only) FAIL(2, "polarization is off");
    if (P.polar_gs || P.polar_gs_ranked) FAIL(3, "Gauss-Seidel polarization is single-GPU only (sequential dependency)");
    if (check_supported(P)) return 3;
    CU(cudaSetDevice(c->device));
    switch (stage) {
        case MPMC_STAGE_FIELD:
            polar_prepare(c);
            if (polar_field(c)) return 10;
            c->last_had_polar = true;
            return 0;
        case MPMC_STAGE_INIT:
            phase_begin(c, PH_DIPOLE);
            return polar_init(c) ? 10 : 0;
        case MPMC_STAGE_PASS:
            if (P.polar_zodid || c->pol.n <= 0) return 0;
            return polar_jacobi_pass(c, iteration) ? 10 : 0;
        case MPMC_STAGE_UPDATE:
            if (P.polar_zodid || c->pol.n <= 0 || !c->use_sym || c->world == 1) return 0;  // done inside PASS
            return polar_update_sym(c, iteration) ? 10 : 0;
        case MPMC_STAGE_PALMO:
            if (P.polar_zodid || c->pol.n <= 0) return 0;
            return polar_palmo_pass(c) ? 10 : 0;
        case MPMC_STAGE_PALMO_FINISH:
            if (P.polar_zodid || c->pol.n <= 0 || !c->use_sym || c->world == 1) return 0;  // done inside PALMO
            return polar_palmo_finish_sym(c) ? 10 : 0;
        case MPMC_STAGE_COMMIT:
            if (P.polar_zodid || c->pol.n <= 0) return 0;
            return polar_commit(c, iteration) ? 10 : 0;
        case MPMC_STAGE_FAIL:  // divergence bail-out (thole_iterative.c:199-210): mu = alpha E_s on every rank
            reset_dipoles_kernel<<<(c->pol.n + 255) / 256, 256, 0, c->stream>>>(c->pol.n, c->pol.n_pad, c->pol.alpha.p,
                                                                               c->es.p, c->mu.p, c->eic.p);
            c->launches++;
            c->pending_iter_success = 1;
            CU(cudaGetLastError());
            return 0;
        case MPMC_STAGE_ENERGY:
            phase_end(c, PH_DIPOLE);
            c->pending_iterations = iteration;
            if (polar_energy_partial(c)) return 10;
            join_side(c);
            phase_end(c, PH_TOTAL);
            c->pending = true;
            return 0;
        default: FAIL(1, "unknown polar stage");
    }
}

int mpmc_b200_uses_symmetric_contraction(mpmc_b200_ctx* c) { return (c && c->use_sym) ? 1 : 0; }

// ---- peer-memory exchange set-up (CUDA IPC; one process per GPU on one node)
int mpmc_b200_peer_export(mpmc_b200_ctx* c, unsigned char* handle64) {
    if (!c || !handle64) FAIL(1, "null argument");
    if (!c->have_atoms) FAIL(2, "mpmc_b200_set_atoms has not been called");
    if (sizeof(cudaIpcMemHandle_t) > 64) FAIL(3, "unexpected cudaIpcMemHandle_t size");
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    peer_teardown(c);
    if (!c->group.empty() || c->group_member) FAIL(2, "the ranks of a group are wired by the library itself");
    c->peer_cap = (size_t)c->pol.n_pad;
    c->peer_cap_atoms = (size_t)c->n_pad;
    const size_t bytes = peer_bytes(c->peer_cap, c->peer_cap_atoms);
    CU(cudaMalloc(&c->peer_local, bytes));
    CU(cudaMemset(c->peer_local, 0, bytes));
    cudaIpcMemHandle_t h;
    CU(cudaIpcGetMemHandle(&h, c->peer_local));
    memset(handle64, 0, 64);
    memcpy(handle64, &h, sizeof(h));
    return 0;
}

int mpmc_b200_peer_connect(mpmc_b200_ctx* c, int world, const unsigned char* handles) {
    if (!c || !handles) FAIL(1, "null argument");
    if (world != c->world || world > PEER_MAX) FAIL(1, "peer_connect: world does not match set_shard (or exceeds 8)");
    if (!c->peer_local) FAIL(2, "mpmc_b200_peer_export has not been called");
    CU(cudaSetDevice(c->device));
    for (int r = 0; r < world; r++) {
        if (r == c->rank) {
            c->peer_base[r] = c->peer_local;
            continue;
        }
        cudaIpcMemHandle_t h;
        memcpy(&h, handles + 64 * (size_t)r, sizeof(h));
        void* p = nullptr;
        if (cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
            cudaGetLastError();
            c->peer_ipc = true;  // close the handles opened so far
            peer_teardown(c);
            FAIL(13, "cudaIpcOpenMemHandle failed (no peer access between the ranks' GPUs?)");
        }
        c->peer_base[r] = p;
    }
    memset(c->peer_seq, 0, sizeof(c->peer_seq));
    c->peer_ipc = true;
    c->peer_ready = true;
    return 0;
}

int mpmc_b200_peer_active(mpmc_b200_ctx* c) { return (c && peer_active(c)) ? 1 : 0; }
int mpmc_b200_uses_cluster_field(mpmc_b200_ctx* c) { return (c && c->use_clusters) ? 1 : 0; }

int mpmc_b200_shard_field_rows(mpmc_b200_ctx* c, int* begin, int* end, int* n_atoms_padded) {
    if (!c) FAIL(1, "null ctx");
    if (!c->have_atoms) FAIL(2, "no atoms");
    int tb, te;
    shard_range(c->n_pad / TILE, c->rank, c->world, tb, te);
    if (begin) *begin = tb * TILE;
    if (end) *end = te * TILE;
    if (n_atoms_padded) *n_atoms_padded = c->n_pad;
    return 0;
}

int mpmc_b200_device_buffer(mpmc_b200_ctx* c, int which, void** dev_ptr, long long* n_doubles) {
    if (!c || !dev_ptr || !n_doubles) FAIL(1, "null argument");
    if (!c->have_atoms) FAIL(2, "no atoms");
    switch (which) {
        case 0: *dev_ptr = c->mu.p; *n_doubles = 3LL * c->pol.n_pad; break;
        case 1: *dev_ptr = c->mu_new.p; *n_doubles = 3LL * c->pol.n_pad; break;
        case 2: *dev_ptr = c->scalars.p; *n_doubles = 32; break;
        case 3: *dev_ptr = c->ef_static.p; *n_doubles = 3LL * c->n_pad; break;
        case 4: *dev_ptr = c->esum.p; *n_doubles = 3LL * c->pol.n_pad; break;
        default: FAIL(1, "unknown buffer");
    }
    return 0;
}

int mpmc_b200_shard_rows(mpmc_b200_ctx* c, int* begin, int* end, int* n_pol, int* n_pol_padded) {
    if (!c) FAIL(1, "null ctx");
    if (!c->have_atoms) FAIL(2, "no atoms");
    int tb, te;
    shard_range(c->pol.n_pad / TILE, c->rank, c->world, tb, te);
    if (begin) *begin = tb * TILE;
    if (end) *end = te * TILE;
    if (n_pol) *n_pol = c->pol.n;
    if (n_pol_padded) *n_pol_padded = c->pol.n_pad;
    return 0;
}

// what polar() leaves in atom_t, assembled on the device in the caller's layout -> the pinned host image c->h_wb
// the write-back image of the evaluation that was just enqueued (writeback_kernel), into `dst` (14 n doubles)
static int enqueue_writeback(mpmc_b200_ctx* c, double* dst, int kb, int ke) {
    const int n = c->n, np = c->n_pad, pp = c->pol.n_pad;
    cudaStream_t st = c->stream;
    CU(cudaMemsetAsync(dst, 0, sizeof(double) * (size_t)14 * n, st));
    const int m = std::max(n, std::max(c->pol.n, c->nonpol.n));
    writeback_kernel<<<(m + 255) / 256, 256, 0, st>>>(
        n, np, c->ef_static.p, c->pol.n, pp, c->pol.didx.p, c->mu.p, c->eind.p, c->eic.p, c->rrms.p, c->rank_packed.p, kb, ke,
        c->nonpol.n, c->nonpol.n_pad, c->nonpol.didx.p, c->eind_np_valid ? c->eind_np.p : nullptr,
        c->eic_np_valid ? c->eic_np.p : nullptr, dst);
    CU(cudaGetLastError());
    return 0;
}

static int fetch_atom_arrays(mpmc_b200_ctx* c) {
    if (!c) FAIL(1, "null ctx");
    if (!c->have_atoms) FAIL(2, "no atoms");
    CU(cudaSetDevice(c->device));
    const int n = c->n, np = c->n_pad, pp = c->pol.n_pad;
    if (n <= 0) return 0;
    // Sharded runs: mu and ef_static are complete on every rank.  With the symmetric contraction every rank
    // also holds the complete E_ind / rrms (all-reduced field, every rank updates every row); in the row form
    // the row-owned arrays are reported for this rank's rows only (zero elsewhere) and the caller sums over ranks.
    int kb = 0, ke = c->pol.n;
    if (c->world > 1 && !mpmc_b200_atom_arrays_complete(c)) {
        int tb, te;
        shard_range(pp / TILE, c->rank, c->world, tb, te);
        kb = tb * TILE;
        ke = std::min(te * TILE, c->pol.n);
    }
    const size_t total = (size_t)14 * n;
    if (c->wb.ensure(total)) FAIL(11, "cudaMalloc failed");
    if (c->h_wb_cap < total) {
        if (c->h_wb) cudaFreeHost(c->h_wb);
        c->h_wb = nullptr;
        c->h_wb_cap = total + total / 4;
        CU(cudaMallocHost(&c->h_wb, sizeof(double) * c->h_wb_cap));
    }
    cudaStream_t st = c->stream;
    if (c->group.size() > 1 && c->eic_np_valid && c->nonpol.n > 0) {
        // palmo rows of the non-polarizable atoms (write-back only) are computed tile range by tile range across the
        // ranks: fetch the other ranks' ranges over the peer link (their streams were synchronised by finish_energy)
        const int qp = c->nonpol.n_pad;
        for (size_t r = 1; r < c->group.size(); r++) {
            mpmc_b200_ctx* o = c->group[r];
            int tb, te;
            shard_range(qp / TILE, o->rank, o->world, tb, te);
            if (te <= tb) continue;
            for (int d = 0; d < 3; d++)
                CU(cudaMemcpyPeerAsync(c->eic_np.p + (size_t)d * qp + (size_t)tb * TILE, c->device,
                                       o->eic_np.p + (size_t)d * qp + (size_t)tb * TILE, o->device,
                                       sizeof(double) * (size_t)(te - tb) * TILE, st));
        }
    }
    if (enqueue_writeback(c, c->wb.p, kb, ke)) return 10;
    CU(cudaMemcpyAsync(c->h_wb, c->wb.p, sizeof(double) * total, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return 0;
}

int mpmc_b200_get_atom_arrays(mpmc_b200_ctx* c, double* mu, double* ef_static, double* ef_induced,
                              double* ef_induced_change, double* dipole_rrms, double* rank_metric) {
    if (const int rc = fetch_atom_arrays(c)) return rc;
    const int n = c->n;
    if (n <= 0) return 0;
    const double* h = c->h_wb;
    if (mu) memcpy(mu, h, sizeof(double) * 3 * n);
    if (ef_static) memcpy(ef_static, h + 3 * (size_t)n, sizeof(double) * 3 * n);
    if (ef_induced) memcpy(ef_induced, h + 6 * (size_t)n, sizeof(double) * 3 * n);
    if (ef_induced_change) memcpy(ef_induced_change, h + 9 * (size_t)n, sizeof(double) * 3 * n);
    if (dipole_rrms) memcpy(dipole_rrms, h + 12 * (size_t)n, sizeof(double) * n);
    if (rank_metric) memcpy(rank_metric, h + 13 * (size_t)n, sizeof(double) * n);
    return 0;
}

int mpmc_b200_atom_arrays_view(mpmc_b200_ctx* c, const double** mu, const double** ef_static, const double** ef_induced,
                               const double** ef_induced_change, const double** dipole_rrms, const double** rank_metric) {
    if (const int rc = fetch_atom_arrays(c)) return rc;
    const size_t n = (size_t)c->n;
    const double* h = c->h_wb;
    if (mu) *mu = h;
    if (ef_static) *ef_static = h + 3 * n;
    if (ef_induced) *ef_induced = h + 6 * n;
    if (ef_induced_change) *ef_induced_change = h + 9 * n;
    if (dipole_rrms) *dipole_rrms = h + 12 * n;
    if (rank_metric) *rank_metric = h + 13 * n;
    return 0;
}

int mpmc_b200_atom_arrays_complete(mpmc_b200_ctx* c) {
    if (!c) return 0;
    if (c->world <= 1) return 1;
    const bool gs = c->P.polar_gs || c->P.polar_gs_ranked;
    return (c->use_sym && !gs) ? 1 : 0;
}

int mpmc_b200_get_counters(mpmc_b200_ctx* c, mpmc_b200_counters* out) {
    if (!c || !out) FAIL(1, "null argument");
    *out = c->cnt;
    return 0;
}

int mpmc_b200_phase_ms(mpmc_b200_ctx* c, const char* phase, double* ms) {
    if (!c || !phase || !ms) FAIL(1, "null argument");
    if (!strcmp(phase, "contract")) {  // mean device time of one K4 launch in the last evaluation
        *ms = c->contract_launches ? c->contract_ms_sum / c->contract_launches : 0.0;
        return 0;
    }
    static const char* names[PH_COUNT] = {"pair", "recip", "field", "dipole", "total"};
    for (int p = 0; p < PH_COUNT; p++)
        if (!strcmp(phase, names[p])) {
            *ms = c->phase_ms[p];
            return 0;
        }
    FAIL(1, "unknown phase");
}

int mpmc_b200_measure_fp64_peak(mpmc_b200_ctx* c, double* tflops) {
    if (!c || !tflops) FAIL(1, "null argument");
    CU(cudaSetDevice(c->device));
    *tflops = measure_fp64_peak_tflops(c->stream);
    CU(cudaGetLastError());
    return 0;
}

}  // extern "C"
