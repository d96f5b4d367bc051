// issue_probe.cu -- does a non-fp64 instruction cost issue time next to DFMA on B200?
// Each thread runs 8 independent DFMA chains; variants add 2 extra instructions of another pipe per 8 DFMA.
// 4 CTAs x 128 threads per SM (the occupancy of K4).  Prints ms and the slowdown against the pure-DFMA loop.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o profiles/issue_probe profiles/issue_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(128, 4) probe(double* out, int iters, double x, int k) {
    __shared__ double sm[256];
    sm[threadIdx.x] = x + threadIdx.x;
    sm[threadIdx.x + 128] = x;
    __syncthreads();
    double a0 = x, a1 = x + 1, a2 = x + 2, a3 = x + 3, a4 = x + 4, a5 = x + 5, a6 = x + 6, a7 = x + 7;
    const double m = 0.999999, c = 1e-9;
    int u = threadIdx.x, v = k;
    double e0 = x * 0.3, e1 = x * 0.7;
    for (int i = 0; i < iters; i++) {
        a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
        a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
        if (MODE == 1) { u = __shfl_sync(0xffffffffu, u, (threadIdx.x + 1) & 31); v = __shfl_sync(0xffffffffu, v, (threadIdx.x + 3) & 31); }
        if (MODE == 2) { u = (u ^ v) + i; v = (v & u) | k; }
        if (MODE == 3) { e0 = rint(e0 * 1.0000001); e1 = rint(e1 + 0.5); }  // FRND + 1 fp64 each
        if (MODE == 4) { e0 += sm[(u + i) & 255]; u += 1; }                 // LDS (+1 DADD, + int)
        if (MODE == 5) { u = __shfl_sync(0xffffffffu, u, (threadIdx.x + 1) & 31); v = __shfl_sync(0xffffffffu, v, (threadIdx.x + 3) & 31);
                         u = (u ^ v) + i; v = (v & u) | k; }
        if (MODE == 6) { asm volatile("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(e0) : "d"(e0 + 1.5)); }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7 + u + v + e0 + e1;
}

template <int MODE>
float run(double* d, int iters) {
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    probe<MODE><<<148 * 4, 128>>>(d, 64, 1.0, 3);
    cudaEventRecord(a);
    probe<MODE><<<148 * 4, 128>>>(d, iters, 1.0, 3);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    return ms;
}

int main() {
    double* d;
    cudaMalloc(&d, sizeof(double) * 148 * 4 * 128);
    const int iters = 200000;
    const char* names[] = {"8 DFMA", "8 DFMA + 2 SHFL", "8 DFMA + 2x2 int ALU", "8 DFMA + 2 (FRND + fp64)", "8 DFMA + LDS + DADD + 2 int",
                           "8 DFMA + 2 SHFL + 4 int", "8 DFMA + MUFU.RSQ64H + DADD"};
    float t[7];
    t[0] = run<0>(d, iters); t[1] = run<1>(d, iters); t[2] = run<2>(d, iters); t[3] = run<3>(d, iters);
    t[4] = run<4>(d, iters); t[5] = run<5>(d, iters); t[6] = run<6>(d, iters);
    const double flops = 148.0 * 4 * 128 * 8.0 * 2.0 * iters;
    for (int i = 0; i < 7; i++)
        printf("%-32s %8.3f ms  x%.3f  (DFMA rate %.1f TFLOP/s)\n", names[i], t[i], t[i] / t[0], flops / (t[i] * 1e-3) / 1e12);
    return 0;
}
