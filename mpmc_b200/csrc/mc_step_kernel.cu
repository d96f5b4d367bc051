// mc_step_kernel.cu -- resident MC steps: the two small kernels that bracket an evaluation when a batch
// of displacement moves is run without the host in between (engine.cu: mpmc_b200_mc_steps).
//
// Reference code restated (relative to the reference tree):
//   translate()                  src/mc/mc_moves.c:376-397   pos += trans, com += trans
//   rotate()                     src/mc/mc_moves.c:401-464   pos = q (pos - com) q* + com
//   quaternion_multiplication()  src/main/quaternion.c:76-86
//   update_com()                 src/energy/pairs.c:364-385  (what molecule->com holds when make_move() starts)
//   boltzmann_factor() NVT       src/mc/mc.c:106-114 and the Metropolis test mc.c:320; restore() mc_moves.c:744-804
// The host compiler of the reference (gcc, x86-64, no -mfma) rounds every product and sum separately; the device code
// does the same with the _rn intrinsics, in the reference's operation order, so the trial coordinates are the bits
// make_move() produces -- the drop-in shim checks exactly that for every step it serves from a batch.
#include "kernels.h"

namespace mpmc {

struct Quat {
    double x, y, z, w;
};

// w = a.w b.w - a.x b.x - a.y b.y - a.z b.z ... left to right, one rounding per operation (quaternion.c:77-80)
__device__ __forceinline__ Quat qmul(const Quat& a, const Quat& b) {
    Quat r;
    r.w = __dsub_rn(__dsub_rn(__dsub_rn(__dmul_rn(a.w, b.w), __dmul_rn(a.x, b.x)), __dmul_rn(a.y, b.y)), __dmul_rn(a.z, b.z));
    r.x = __dsub_rn(__dadd_rn(__dadd_rn(__dmul_rn(a.w, b.x), __dmul_rn(a.x, b.w)), __dmul_rn(a.y, b.z)), __dmul_rn(a.z, b.y));
    r.y = __dadd_rn(__dadd_rn(__dsub_rn(__dmul_rn(a.w, b.y), __dmul_rn(a.x, b.z)), __dmul_rn(a.y, b.w)), __dmul_rn(a.z, b.x));
    r.z = __dadd_rn(__dsub_rn(__dadd_rn(__dmul_rn(a.w, b.z), __dmul_rn(a.x, b.y)), __dmul_rn(a.y, b.x)), __dmul_rn(a.z, b.w));
    return r;
}

// The step index is either given (j >= 0) or read from the device counter the decide kernel advances: the second form
// lets one captured CUDA graph of a step be replayed for every later step of the batch.
__global__ void mc_move_kernel(const mpmc_b200_mc_move* __restrict__ moves, int j, const int* __restrict__ step_ctr,
                               double* __restrict__ x, double* __restrict__ y, double* __restrict__ z,
                               McStepDev* __restrict__ res) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    if (j < 0) j = *step_ctr;
    const mpmc_b200_mc_move& m = moves[j];
    McStepDev& r = res[j];
    const int n = m.count, f = m.first;
    if (m.explicit_positions) {  // a move the caller made on its lists: the coordinates are given
        for (int i = 0; i < n; i++) {
            for (int p = 0; p < 3; p++) {
                r.old_pos[3 * i + p] = m.old_pos[3 * i + p];
                r.new_pos[3 * i + p] = m.new_pos[3 * i + p];
            }
            x[f + i] = m.new_pos[3 * i];
            y[f + i] = m.new_pos[3 * i + 1];
            z[f + i] = m.new_pos[3 * i + 2];
        }
        return;
    }
    double pos[MPMC_B200_MC_MAX_ATOMS][3];
    double mass = 0.0, com[3] = {0.0, 0.0, 0.0};
    for (int i = 0; i < n; i++) {
        pos[i][0] = x[f + i];
        pos[i][1] = y[f + i];
        pos[i][2] = z[f + i];
        for (int p = 0; p < 3; p++) r.old_pos[3 * i + p] = pos[i][p];
        mass = __dadd_rn(mass, m.mass[i]);  // pairs.c:374-378
        for (int p = 0; p < 3; p++) com[p] = __dadd_rn(com[p], __dmul_rn(m.mass[i], pos[i][p]));
    }
    for (int p = 0; p < 3; p++) com[p] = __ddiv_rn(com[p], mass);
    // translate(): mc_moves.c:387-396
    for (int p = 0; p < 3; p++) com[p] = __dadd_rn(com[p], m.trans[p]);
    for (int i = 0; i < n; i++)
        for (int p = 0; p < 3; p++) pos[i][p] = __dadd_rn(pos[i][p], m.trans[p]);
    // rotate(): mc_moves.c:432-460
    const Quat q = {m.quat[0], m.quat[1], m.quat[2], m.quat[3]};
    const Quat qc = {-q.x, -q.y, -q.z, q.w};
    for (int i = 0; i < n; i++) {
        Quat pv = {__dsub_rn(pos[i][0], com[0]), __dsub_rn(pos[i][1], com[1]), __dsub_rn(pos[i][2], com[2]), 0.0};
        Quat a = qmul(pv, qc);
        a = qmul(q, a);
        pos[i][0] = __dadd_rn(a.x, com[0]);
        pos[i][1] = __dadd_rn(a.y, com[1]);
        pos[i][2] = __dadd_rn(a.z, com[2]);
        for (int p = 0; p < 3; p++) r.new_pos[3 * i + p] = pos[i][p];
        x[f + i] = pos[i][0];
        y[f + i] = pos[i][1];
        z[f + i] = pos[i][2];
    }
}

// After the evaluation of step j: keep its scalars, assemble the total exactly as engine.cu:assemble_energies() does on
// the host (same operations, same order), decide, and put the old coordinates back on a reject.
__global__ void mc_decide_kernel(const mpmc_b200_mc_move* __restrict__ moves, int j, int* __restrict__ step_ctr, McDecide k,
                                 const double* __restrict__ scalars, double* __restrict__ e_current, double* __restrict__ x,
                                 double* __restrict__ y, double* __restrict__ z, McStepDev* __restrict__ res) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    if (j < 0) j = *step_ctr;
    *step_ctr = j + 1;
    const mpmc_b200_mc_move& m = moves[j];
    McStepDev& r = res[j];
    for (int s = 0; s < 32; s++) r.scalars[s] = scalars[s];
    const double* h = scalars;
    double rd_pair = h[k.s_pair + 0];
    if (k.rd_crystal) rd_pair = __dadd_rn(rd_pair, __dadd_rn(h[k.s_crystal], k.crystal_self));
    const double rd = __dadd_rn(__dadd_rn(rd_pair, k.lrc_pair), k.lrc_self);
    double coulombic = 0.0;
    if (!k.rd_only) {
        const double es_real = __dsub_rn(h[k.s_pair + 1], h[k.s_pair + 2]);
        const double es_recip = k.es_wolf ? 0.0 : h[k.s_recip];
        const double es_self = k.es_wolf ? 0.0 : k.es_self;
        coulombic = __dadd_rn(__dadd_rn(es_real, es_recip), es_self);
    }
    double polar = 0.0;
    int iter_success = k.iter_success;
    if (k.polar) {
        double pot = h[k.s_polar + 0];
        if (k.palmo) pot = __dadd_rn(pot, h[k.s_polar + 1]);
        polar = __dmul_rn(-0.5, pot);
        if (k.iter_from_device) iter_success = (int)h[k.s_fused_out + 1];
    }
    const double total = __dadd_rn(__dadd_rn(rd, coulombic), polar);  // energy.c:194
    // mc.c:312-320: a non-finite energy is a reject; else u < exp(-(E_new - E_old) / T) and a converged solve
    double bf = 0.0;
    if (isfinite(total)) bf = exp(__ddiv_rn(-__dsub_rn(total, *e_current), k.temperature));
    const bool accept = (m.u_accept < bf) && iter_success == 0;
    r.total = total;
    r.accepted = accept ? 1.0 : 0.0;
    if (accept) {
        *e_current = total;
    } else {  // restore(): the backup's coordinates (mc_moves.c:790-799)
        for (int i = 0; i < m.count; i++) {
            x[m.first + i] = r.old_pos[3 * i];
            y[m.first + i] = r.old_pos[3 * i + 1];
            z[m.first + i] = r.old_pos[3 * i + 2];
        }
    }
}

// the write-back image of the step (14 n doubles) into the step's slot of the ring
__global__ void mc_ring_copy_kernel(const double* __restrict__ wb, double* __restrict__ ring, size_t per_step, int j,
                                    const int* __restrict__ step_ctr) {
    if (j < 0) j = *step_ctr;
    double* dst = ring + per_step * (size_t)j;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < per_step; i += (size_t)gridDim.x * blockDim.x) dst[i] = wb[i];
}

void launch_mc_move(const mpmc_b200_mc_move* moves, int j, const int* step_ctr, double* x, double* y, double* z, McStepDev* res,
                    cudaStream_t s) {
    mc_move_kernel<<<1, 32, 0, s>>>(moves, j, step_ctr, x, y, z, res);
}
void launch_mc_decide(const mpmc_b200_mc_move* moves, int j, int* step_ctr, const McDecide& k, const double* scalars,
                      double* e_current, double* x, double* y, double* z, McStepDev* res, cudaStream_t s) {
    mc_decide_kernel<<<1, 32, 0, s>>>(moves, j, step_ctr, k, scalars, e_current, x, y, z, res);
}
void launch_mc_ring_copy(const double* wb, double* ring, size_t per_step, int j, const int* step_ctr, cudaStream_t s) {
    const int blocks = (int)((per_step + 255) / 256 < 592 ? (per_step + 255) / 256 : 592);
    mc_ring_copy_kernel<<<blocks > 0 ? blocks : 1, 256, 0, s>>>(wb, ring, per_step, j, step_ctr);
}

}  // namespace mpmc
