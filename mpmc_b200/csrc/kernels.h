// kernels.h -- host-callable launchers of the sm_100a kernels (one per reference loop).
#pragma once
#include <cuda_runtime.h>

#include "common.cuh"

namespace mpmc {

enum : int { F_FROZEN = 1, F_VALID = 2 };

// message behind mpmc_b200_last_error() (engine.cu), for entry points defined in other translation units
void set_last_error(const char* msg);

// All atoms, reference list order, padded to a multiple of TILE (pad atoms have flags == 0).
struct AtomsDev {
    int n, n_pad;
    double *x, *y, *z, *q, *eps, *sig, *molmass, *alpha;
    int *mol, *flags;
};

// A compacted subset of atoms (charged atoms for the static field, charged mobile atoms for the
// structure factor, polarizable atoms for the dipole solve), padded to a multiple of TILE.
struct SubsetDev {
    int n, n_pad;
    double *x, *y, *z, *q, *alpha;
    double *aux;  // dipole kernels: alpha^(1/6) for linear damping, the charge for damping off
    int *mol, *flags, *idx;  // idx = index in AtomsDev
    // only in the spatially sorted image of ALL atoms used by the cutoff kernels (else null):
    double *eps, *sig, *molmass;
    double* cbox;  // [7][n_pad/32] cluster boxes of this evaluation (centre, half extent, one-image flag)
};

// ---- K1: fused LJ(+FH) + Ewald real(+FH) + intramolecular erf + min distance ------------------
// replaces pair_exclusions/minimum_image (src/energy/pairs.c:56-290), the pair loops of lj()
// (src/energy/lj.c:181-260) and coulombic_real() (src/energy/coulombic.c:159-191), and
// cavity_absolute_check() (src/energy/energy.c:49-64).
struct PairOpts {
    double ewald_alpha, temperature;
    int fh_order;  // 0 = off, 2, 4
    int mixing;
    int rd_only;
    int es_wolf = 0;  // coulombic_wolf() instead of coulombic_real() (needs fh_order == 0)
    int no_lj = 0;    // K1 without the LJ term (rd_crystal: launch_crystal_lj computes it)
};
// rd_crystal (lj.c:175-216): LJ image sums over all tile pairs of the list-ordered atoms.  img [nimg][3] = B^T l in the
// reference's loop order; partial [2][stride] (energy, pair count); out[0..1] = their fixed-order sums.
int crystal_max_images();
void launch_crystal_lj(const Cell& c, bool ortho, const AtomsDev& a, const PairOpts& o, double rd_cutoff, const double* img,
                       int nimg, const int2* tiles, int ntiles, double* partial, int stride, double* out, cudaStream_t s);
// ---- incremental evaluation after a small move (opt-in) ---------------------------------------------
// The atoms whose positions changed since the last evaluation, with their positions then (o) and now (n).
struct MovedRec {
    double xn, yn, zn, xo, yo, zo, q, eps, sig, mm;
    int mol, fl, idx, pad;
};
constexpr int INC_MAX_MOVED = 128;
constexpr int INC_NPART = 5;  // d(rd), d(es), d(intra), d(n_lj), d(n_es)
// Change of the K1 sums: for every pair with at least one moved atom, the pair terms at the new positions minus the
// pair terms at the old ones (each such pair once).  One thread per atom j of the list-ordered arrays (which hold the
// NEW positions), the moved atoms in shared memory.  partial [INC_NPART][stride], one column per CTA; out[INC_NPART].
void launch_pair_delta(const Cell& c, bool ortho, const AtomsDev& a, const PairOpts& o, const MovedRec* moved, int nmoved,
                       double* partial, int stride, double* out, cudaStream_t s);
constexpr int PAIR_NPART = 8;  // rd, es, intra, min_r, n_tested, n_lj, n_es, (unused)
void launch_pair_energy(const Cell& c, bool ortho, const AtomsDev& a, const PairOpts& o, const int2* tiles, int ntiles,
                        double* partial /*[PAIR_NPART][ntiles_stride]*/, int stride, cudaStream_t s);
// Cluster form of K1 on the spatially sorted image: one warp per 32-atom i-cluster, candidate j-clusters
// (ci + k) mod nclus, k = 0..nclus/2 (every unordered cluster pair once), culled by the bounding-box gap against
// `cull` (>= cutoff; <= 0: no culling).  partial [PAIR_NPART][stride], one column per (slice, own cluster).
void launch_pair_energy_clusters(const Cell& c, bool ortho, const SubsetDev& sorted, const PairOpts& o, int clus_begin,
                                 int clus_end, int kslices, double cull, double* partial, int stride, cudaStream_t s);
int pair_cluster_columns(int n_pad_sorted, int clus_begin, int clus_end, int kslices);  // columns of `partial` used
// intramolecular erf term of coulombic_real() (src/energy/coulombic.c:176-183) over list-order atoms;
// follow[i] = number of later atoms of the same (not entirely frozen) molecule; part[blocks] -> sum_out
void launch_intra_erf(const AtomsDev& a, const int* follow, double ewald_alpha, int row_begin, int row_end, double* part,
                      double* sum_out, cudaStream_t s);
void launch_pair_reduce(const double* partial, int ntiles, int stride, double* out /*[PAIR_NPART]*/, cudaStream_t s);

// ---- K2: reciprocal-space structure factor -----------------------------------------------------
// replaces coulombic_reciprocal() (src/energy/coulombic.c:42-95)
struct KVecDev {
    int nk, kmax;
    int* l;          // [3][nk]
    double* weight;  // exp(-k^2/4a^2)/k^2
};
void launch_recip_sf(const Cell& c, const SubsetDev& atoms, const KVecDev& kv, int chunk_begin, int chunk_end,
                     int atoms_per_chunk, double* sf_partial /*[nchunks][2][nk]*/, cudaStream_t s);
// sf = sum over chunks (always written); energy_out (optional) = scale * sum_{k in [k_begin,k_end)} w_k |S(k)|^2 : the
// k-vector range is this shard's share.  mu (optional, [3][mu_np]): mu_total[3] = sum of the first mu_n dipoles, in the same
// launch.  Returns the number of kernels launched.
int launch_recip_finish(const KVecDev& kv, const double* sf_partial, int nchunks, double scale,
                        double* sf /*[2][nk]*/, double* energy_out, int k_begin, int k_end, cudaStream_t s, int mu_n = 0,
                        int mu_np = 0, const double* mu = nullptr, double* mu_total = nullptr);
// S(k) += sum over the moved, charged, mobile atoms of q (e^{ik.r_new} - e^{ik.r_old}); then energy_out as
// launch_recip_finish computes it.  Two launches.
void launch_recip_delta(const KVecDev& kv, const double* kvec /*[3][nk]*/, const MovedRec* moved, int nmoved, double* sf,
                        double scale, double* energy_out, cudaStream_t s);
// recip_term() of the Ewald static field (src/polarization/polar_ewald.c:78-132)
void launch_recip_field(const Cell& c, const AtomsDev& a, const KVecDev& kv, const double* kvec /*[3][nk]*/,
                        const double* sf, double scale, int row_begin, int row_end, double* ef /*[3][n_pad]*/,
                        cudaStream_t s);

// ---- K3: static field --------------------------------------------------------------------------
// replaces thole_field_nopbc / thole_field_wolf (src/polarization/thole_field.c:39-124) and
// real_term() (src/polarization/polar_ewald.c:38-74)
enum : int { FIELD_NOPBC = 0, FIELD_WOLF_A0 = 1, FIELD_WOLF = 2, FIELD_WOLF_LOOKUP = 3, FIELD_EWALD_REAL = 4 };
struct FieldOpts {
    int mode;
    double a;           // polar_wolf_alpha or polar_ewald_alpha
    double cutoffterm;  // wolf: value of the kernel at R_c
    const double* table;
    int table_max;
    double in2 = 0.0, out2 = 0.0;  // (cutoff -+ 1e-9)^2, filled by the launchers: the range test on r^2 away from the boundary
};
void launch_static_field(const Cell& c, bool ortho, const AtomsDev& a, const SubsetDev& charged, const FieldOpts& o,
                         int row_begin, int row_end, int jsplits, double* epart /*[jsplits][3][n_pad]*/,
                         double* count_part, cudaStream_t s);
// ef[c][i] (+)= sum_js epart[js][c][i]
// Cluster form of K3 on the spatially sorted image: one warp per i-cluster, all j-clusters culled by the gap;
// writes ef[c][idx[i]] (list order), accumulating when `accumulate`; count_part[cluster] = contributions.
void launch_static_field_clusters(const Cell& c, bool ortho, const SubsetDev& sorted, const FieldOpts& o, int clus_begin,
                                  int clus_end, double cull, bool accumulate, double* ef, int ef_stride, double* count_part,
                                  cudaStream_t s);
void launch_field_reduce(const double* epart, int jsplits, int n_pad, int row_begin, int row_end, bool accumulate,
                         double* ef, cudaStream_t s);

// ---- K4/K5: matrix-free dipole field tensor contraction and dipole update -----------------------
// replaces thole_amatrix() (src/polarization/thole_matrix.c:38-146) and contract_dipoles(),
// palmo_contraction() (src/polarization/thole_iterative.c:27-59,119-141)
struct DipoleOpts {
    int damp_type;
    int wolf_full;
    double damp;  // polar_damp
};
void launch_dipole_contract(const Cell& c, bool ortho, const SubsetDev& pol, const DipoleOpts& o, const double* mu /*[3][n_pad]*/,
                            int row_begin, int row_end, int jsplits, double* epart /*[jsplits][3][n_pad]*/,
                            cudaStream_t s);
// palmo: field of the dipoles at the atoms of `rows` (the non-polarizable atoms), out[3][rows.n_pad]
// part (optional): [dipole_field_at_splits(row tiles, source tiles)][3][rows.n_pad] scratch -- the source tiles are then
// split over gridDim.y and the slices' partial rows added in slice order (few row tiles would leave the GPU empty)
int dipole_field_at_splits(int row_tiles, int src_tiles);
void launch_dipole_field_at(const Cell& c, bool ortho, const SubsetDev& rows, const SubsetDev& pol, const DipoleOpts& o,
                            const double* mu, int row_begin, int row_end, double* out, cudaStream_t s, double* part = nullptr);
// Symmetric form of the contraction (dipole_sym_kernel.cu): T_ij computed once per unordered pair and
// applied in both directions.  Work items = (I super tile of 256 atoms) x (slice of the offsets 0..S/2, `ysplits` slices); rank r of
// `world` takes the r-th contiguous share of the flattened items;
// partial rows: epartI [dipole_sym_hsplit() * ysplits][3][n_pad], epartJ [dipole_sym_partial_rows(n_pad, ysplits)][3][n_pad];
// esum [3][n_pad] receives this launch's contribution to E_ind for ALL rows (summed over ranks by the
// caller in sharded runs).  cbox [7][n_pad/32] from launch_cluster_boxes().
// Peer-memory exchange of the induced-field partial sums (multi-GPU, one process per GPU): every rank's reduce
// kernel leaves its contribution to E_ind of ALL rows in a buffer the other ranks can read over NVLink
// (CUDA IPC), then raises a flag in every peer's memory; the update kernel of each rank waits for all flags and
// adds the `world` contributions in rank order -- the all-reduce, the field sum and the dipole update are one
// kernel, and every rank gets bit-identical sums.
constexpr int PEER_MAX = 8;
struct PeerSignal {             // producer side (peer_signal_kernel, launched after dipole_sym_reduce_kernel)
    int world, me;
    unsigned long long seq;     // flag value to publish
    unsigned long long* flags[PEER_MAX];  // flags[r] = rank r's flag array (peer pointer); slot `me` is written
};
struct PeerWait {               // consumer side (dipole_update_peer_kernel, palmo_finish_peer_kernel)
    int world;
    unsigned long long seq;
    const unsigned long long* my_flags;  // local flag array, slot r written by rank r
    const double* src[PEER_MAX];         // src[r] = rank r's partial field [3][n_pad] (peer pointer)
    double* timeout_flag;                // set to 1.0 if a peer never signalled (host turns it into an error)
};
// The same flag protocol for the other two exchanges of a sharded evaluation (engine.cu: sharded_eval):
// out[i] = sum over ranks of src[r][i], i < n (static field rows: one owner per row, zero elsewhere => an all-gather)
void launch_peer_signal(const PeerSignal& sig, cudaStream_t s);
void launch_peer_sum(const PeerWait& pw, size_t n, double* out, cudaStream_t s);
// the evaluation's device scalars: sums over ranks, slot min_slot the minimum; n <= 64
void launch_peer_scalars(const PeerWait& pw, int n, int min_slot, double* out, cudaStream_t s);
int dipole_sym_tile();  // atoms per super tile (256): n_pad of the polarizable subset must be a multiple
int dipole_sym_hsplit(int n_pad, int ysplits, int world);  // gridDim.z of a contraction: I-side partial rows = hsplit * ysplits
int dipole_sym_partial_rows(int n_pad, int ysplits);  // J-side partial rows one launch writes (capped: MPMC_B200_SYM_PART_MB)
void launch_cluster_boxes(const Cell& c, bool ortho, const SubsetDev& pol, double* cbox, cudaStream_t s);
void launch_dipole_contract_sym(const Cell& c, bool ortho, const SubsetDev& pol, const DipoleOpts& o, const double* mu,
                                const double* cbox, int rank, int world, int ysplits, double* epartI,
                                double* epartJ, double* esum, cudaStream_t s, const PeerSignal* sig = nullptr);
struct UpdateOpts {
    int sor, esor, save_old, do_rrms;
    double gamma;
    int iteration;
};
// E_ind = sum_js epart; new_mu = alpha (E_s + E_ind); max (new-old)^2; rrms.  Touches mu only when mu_commit is given.
void launch_dipole_update(const SubsetDev& pol, const UpdateOpts& o, const double* epart, int jsplits, const double* es,
                          const double* mu, double* mu_old, double* mu_new, double* eind, double* rrms,
                          unsigned long long* max_err2_bits, int row_begin, int row_end, cudaStream_t s,
                          double* mu_commit = nullptr /* == mu: commit the rows in the same kernel */);
// Peer form: waits for every rank's flag, E_ind = sum_r src[r] (rank order), then the same update; with
// `commit` the dipoles are committed in the same kernel (Jacobi without palmo).
void launch_dipole_update_peer(const SubsetDev& pol, const UpdateOpts& o, const PeerWait& pw, const double* es, double* mu,
                               double* mu_old, double* mu_new, double* eind, double* rrms,
                               unsigned long long* max_err2_bits, int commit, cudaStream_t s);
void launch_palmo_finish_peer(const SubsetDev& pol, const PeerWait& pw, const double* eind, double* eic, cudaStream_t s);
// mu = new_mu | SOR | ESOR mix (src/polarization/thole_iterative.c:242-252)
void launch_dipole_commit(const SubsetDev& pol, const UpdateOpts& o, const double* mu_old, const double* mu_new,
                          double* mu, int row_begin, int row_end, cudaStream_t s);
// mu0 = gamma alpha E_s (init_dipoles, thole_iterative.c:13-25); also gathers E_s into packed order
void launch_dipole_init(const SubsetDev& pol, const double* ef_static /*[3][atoms n_pad]*/, int atoms_n_pad,
                        double gamma_scale, double* es /*[3][pol n_pad]*/, double* mu, int row_begin, int row_end,
                        cudaStream_t s, const double* mu_start /*[3][atoms n_pad], optional warm start*/ = nullptr);
// mu (packed) -> mu_keep [3][atoms n_pad]: the starting point of the next solve when the warm start is on
void launch_dipole_keep(const SubsetDev& pol, const double* mu, int atoms_n_pad, double* mu_keep, cudaStream_t s);
// ef_induced_change = -ef_induced + sum_js epart (palmo)
void launch_palmo_finish(const SubsetDev& pol, const double* epart, int jsplits, const double* eind, double* eic,
                         int row_begin, int row_end, cudaStream_t s);
// out[0] = sum mu.E_s, out[1] = sum mu.eic, out[2] = sum finite rrms   (fixed-order, one CTA)
void launch_polar_energy(const SubsetDev& pol, const double* mu, const double* es, const double* eic, const double* rrms,
                         int row_begin, int row_end, double* out, cudaStream_t s);

// Small systems: the whole Jacobi-family solve (every iteration, convergence test, bail-out, Palmo pass) in one
// cooperative launch (dipole_kernel.cu: dipole_fused_solver_kernel).  part [ceil(n/128)][3][n_pad]; err2 two device words;
// out2[0] = iterations done, out2[1] = iter_success.  mu_palmo (optional) receives the dipoles the Palmo pass saw.
// false = not launchable here (the caller takes the launch-per-iteration path).
bool launch_dipole_fused_solver(const Cell& c, bool ortho, const SubsetDev& pol, const DipoleOpts& o, const UpdateOpts& uo,
                                int max_iter, double allowed, bool precision_mode, bool palmo, const double* es, double* mu,
                                double* mu_old, double* mu_new, double* eind, double* eic, double* rrms, double* part,
                                double* mu_palmo, unsigned long long* err2, double* out2, cudaStream_t s);
// Gauss-Seidel sweep (polar_gs / polar_gs_ranked): blocked forward substitution in `order`.
void launch_gs_sweep(const Cell& c, bool ortho, const SubsetDev& pol, const DipoleOpts& o, const int* order,
                     const double* es, double* mu, double* mu_new, double* eind, double* scratch, cudaStream_t s);
// Delta form of the sweep (dipole_kernel.cu): one symmetric Jacobi contraction with the old dipoles + right-looking
// block pushes of delta = mu_new - mu_old.  cbox = cluster boxes of `pol` in ITS order; delta [3][n_pad] scratch.
// coef (optional, gs_block_coeff_doubles(n_pad) doubles from launch_gs_block_coeffs for the SAME order): tensor
// coefficients of the in-block pairs, read by the sequential block kernel instead of being recomputed.
size_t gs_block_coeff_doubles(int n_pad);
void launch_gs_block_coeffs(const Cell& c, bool ortho, const SubsetDev& pol, const DipoleOpts& o, double* coef, cudaStream_t s);
void launch_gs_sweep_delta(const Cell& c, bool ortho, const SubsetDev& pol, const DipoleOpts& o, const double* cbox, int ysplits,
                           double* epartI, double* epartJ, double* esum, const double* es, double* mu, double* mu_new,
                           double* eind, double* delta, const double* coef, cudaStream_t s);
// Persistent form of the same sweep (dipole_kernel.cu: gs_sweep_persistent_kernel): one cooperative launch per sweep, a
// resolver CTA walking the blocks and one pusher CTA per other SM spreading each finished delta with look-ahead.
// tens: gs_block_tensor_doubles(n_pad) doubles from launch_gs_block_tensors for the SAME order; pushacc [3][n_pad];
// flags [2 * n_pad / 128] unsigned zeroed at allocation; epoch > 0 and increasing.  false = not launchable here.
size_t gs_block_tensor_doubles(int n_pad);
size_t gs_pair_coeff_doubles(int n_pad);  // (c1, c2) of the pairs the resolver applies itself: [blocks][2][128][128] double2
void launch_gs_block_tensors(const Cell& c, bool ortho, const SubsetDev& pol, const DipoleOpts& o, double* tens, double* ctab,
                             cudaStream_t s);
bool launch_gs_sweep_persistent(const Cell& c, bool ortho, const SubsetDev& pol, const DipoleOpts& o, const double* esum,
                                const double* es, double* mu, double* mu_new, double* eind, double* delta, const double* tens,
                                const double* ctab, double* pushacc, unsigned* flags, unsigned epoch, double* timeout_flag,
                                cudaStream_t s, unsigned long long* trace = nullptr /* [blocks][8] resolver time stamps */);
// rank metric of pairs() (src/energy/pairs.c:337-360)
void launch_rank_metric(const Cell& c, bool ortho, const SubsetDev& pol, double* rmin_part, double* rank /*[pol n_pad]*/,
                        cudaStream_t s);

// ---- K7: full-Ewald polarization (ewald_full_kernel.cu) -----------------------------------------
// replaces induced_real_term / induced_recip_term / induced_corr_term / new_dipoles /
// ewald_palmo_contraction (src/polarization/polar_ewald.c:176-371)
struct EwaldFullOpts {
    double a;     // polar_ewald_alpha
    double damp;  // polar_damp
};
// row-owned real-space term over the spatially ordered polarizable subset; returns the number of j slices
// written to part [slices][3][n_pad]
int launch_induced_real_clusters(const Cell& c, bool ortho, const SubsetDev& pol, const EwaldFullOpts& o, const double* cbox,
                                 const double* mu, int jsplits, double* part, cudaStream_t s);
// Pair-list form of the same term: erfc / exp / rsqrt once per energy(), every pass streams 48-byte entries.
// count -> (host reads offs[4 n_pad] = total, sizes `list`) -> fill -> apply per pass (part [3][n_pad], one slice).
size_t induced_list_entry_bytes();
void launch_induced_list_count(const Cell& c, bool ortho, const SubsetDev& pol, const EwaldFullOpts& o, const double* cbox,
                               unsigned int* cnt, unsigned long long* offs, cudaStream_t s);
void launch_induced_list_fill(const Cell& c, bool ortho, const SubsetDev& pol, const EwaldFullOpts& o, const double* cbox,
                              const unsigned long long* offs, void* list, cudaStream_t s);
void launch_induced_list_apply(const SubsetDev& pol, const unsigned long long* offs, const void* list, const double* mu,
                               double* part, cudaStream_t s);
// P(k) = sum_j (k.mu_j) e^{ik.r_j}: per-chunk partials [chunks][2][nk] (sum with launch_recip_finish); returns chunks
// kg (optional): the (l0, l1) groups of the lexicographic k list, sorted by count (descending):
// {first k index, l0, l1, (first l2 + 64) | count << 8}; with it and max_count <= 15 the grouped kernel runs.
struct KGroupsDev {
    int ngroups, max_count;
    const int4* groups;
};
int launch_dipole_sf(const Cell& c, const SubsetDev& pol, const KVecDev& kv, const double* kvec, const double* mu,
                     int atoms_per_chunk, double* sf_partial, cudaStream_t s, const KGroupsDev* kg = nullptr);
void launch_total_mu(const SubsetDev& pol, const double* mu, double* out3, cudaStream_t s);
// out[3][rows.n_pad] = sum_js part + reciprocal term + surface term (+ self term when mu != null)
void launch_induced_finish(const Cell& c, const SubsetDev& rows, const KVecDev& kv, const double* kvec, const double* sf,
                           const EwaldFullOpts& o, const double* part, int jsplits, const double* total_mu, const double* mu,
                           double* out, cudaStream_t s);
void launch_ewald_update(const SubsetDev& pol, const UpdateOpts& o, const double* es, const double* eind, double* mu,
                         double* mu_old, double* mu_new, double* rrms, unsigned long long* max_err2_bits, cudaStream_t s);
void launch_ewald_palmo_finish(const SubsetDev& pol, const double* es, const double* eind, const double* mu_new, double* eic,
                               cudaStream_t s);

// fp64 FMA throughput probe
double measure_fp64_peak_tflops(cudaStream_t s);

// ---- resident MC steps (mc_step_kernel.cu) -----------------------------------------
// replaces, per step of a batch: make_move() DISPLACE (src/mc/mc_moves.c:376-464,713-731), the Metropolis test
// (src/mc/mc.c:106-114,320) and restore() of the coordinates (mc_moves.c:744-804)
struct McStepDev {  // what a step leaves on the device
    double scalars[32];
    double total, accepted;
    double new_pos[3 * MPMC_B200_MC_MAX_ATOMS], old_pos[3 * MPMC_B200_MC_MAX_ATOMS];
};
struct McDecide {  // what assemble_energies() (engine.cu) adds on the host, and which slots it reads
    double lrc_pair, lrc_self, es_self, crystal_self, temperature;
    int rd_only, es_wolf, rd_crystal, polar, palmo, iter_from_device, iter_success;
    int s_pair, s_recip, s_polar, s_crystal, s_fused_out;
};
// j >= 0: that step; j < 0: the step the device counter points at (a captured graph of one step replayed for the rest)
void launch_mc_move(const mpmc_b200_mc_move* moves, int j, const int* step_ctr, double* x, double* y, double* z, McStepDev* res,
                    cudaStream_t s);
void launch_mc_decide(const mpmc_b200_mc_move* moves, int j, int* step_ctr, const McDecide& k, const double* scalars,
                      double* e_current, double* x, double* y, double* z, McStepDev* res, cudaStream_t s);
void launch_mc_ring_copy(const double* wb, double* ring, size_t per_step, int j, const int* step_ctr, cudaStream_t s);

}  // namespace mpmc
