// FP64 ceilings of this part, for the FP64 Gram kernel (k_gram_f64, the parity anchor of K7): the tensor-core DMMA rate
// (mma.sync m8n8k4 and m16n8k16, f64 x f64 -> f64) and the plain DFMA rate, all operands in registers, eight independent
// accumulator chains per warp / thread.  SURVEY.md 8(d) asks for the FP64 tensor peak next to the TF32 one
// (experiments/umma_peak.cu).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o build/dmma_peak dmma_peak.cu
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1);} } while (0)

constexpr int CHAINS = 8;

__global__ void __launch_bounds__(256) k_dmma884(int iters, double seed, double *sink) {
    double c0[CHAINS], c1[CHAINS];
    const double a = seed + threadIdx.x * 1e-9, b = 1.0 - seed;
#pragma unroll
    for (int k = 0; k < CHAINS; k++) { c0[k] = k; c1[k] = -k; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < CHAINS; k++)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0[k]), "+d"(c1[k]) : "d"(a), "d"(b));
    }
    double s = 0;
#pragma unroll
    for (int k = 0; k < CHAINS; k++) s += c0[k] + c1[k];
    if (s == 12345.678) sink[0] = s;
}

__global__ void __launch_bounds__(256) k_dmma16816(int iters, double seed, double *sink) {
    double c[CHAINS][4];
    double a[8], b[4];
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = seed + (threadIdx.x + i) * 1e-9;
#pragma unroll
    for (int i = 0; i < 4; i++) b[i] = 1.0 - seed + i * 1e-9;
#pragma unroll
    for (int k = 0; k < CHAINS; k++) { c[k][0] = k; c[k][1] = -k; c[k][2] = 0.5 * k; c[k][3] = 1; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < CHAINS; k++)
            asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
                         : "+d"(c[k][0]), "+d"(c[k][1]), "+d"(c[k][2]), "+d"(c[k][3])
                         : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]), "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
    }
    double s = 0;
#pragma unroll
    for (int k = 0; k < CHAINS; k++) s += c[k][0] + c[k][1] + c[k][2] + c[k][3];
    if (s == 12345.678) sink[0] = s;
}

__global__ void __launch_bounds__(256) k_dfma(int iters, double seed, double *sink) {
    double c[CHAINS];
    const double a = 1.0 + seed * 1e-9, b = seed * 1e-12 + threadIdx.x * 1e-15;
#pragma unroll
    for (int k = 0; k < CHAINS; k++) c[k] = k;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < CHAINS; k++) c[k] = fma(c[k], a, b);
    }
    double s = 0;
#pragma unroll
    for (int k = 0; k < CHAINS; k++) s += c[k];
    if (s == 12345.678) sink[0] = s;
}

template <typename F>
static void timed(const char *name, F launch, double flop_per_launch) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    double best = 1e30, sum = 0;
    const int reps = 5;
    for (int r = 0; r < reps + 1; r++) {
        CK(cudaEventRecord(e0));
        launch();
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        CK(cudaGetLastError());
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (r == 0) continue;
        sum += ms;
        if (ms < best) best = ms;
    }
    printf("%-44s best %.3f ms, mean %.3f ms | %.2f TFLOP/s best, %.2f mean\n", name, best, sum / reps, flop_per_launch / best * 1e-9, flop_per_launch / (sum / reps) * 1e-9);
}

int main(int argc, char **argv) {
    const int iters = argc > 1 ? atoi(argv[1]) : 4096;
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int ctas_per_sm = 4, grid = prop.multiProcessorCount * ctas_per_sm, warps = grid * 8;
    double *sink;
    CK(cudaMalloc(&sink, 8));
    printf("%s, %d SMs, %d CTAs x 256 threads, %d iterations x %d chains\n", prop.name, prop.multiProcessorCount, grid, iters, CHAINS);
    timed("DMMA mma.sync m8n8k4 f64", [&] { k_dmma884<<<grid, 256>>>(iters, 0.25, sink); }, (double)warps * iters * CHAINS * 2.0 * 8 * 8 * 4);
    timed("DMMA mma.sync m16n8k16 f64", [&] { k_dmma16816<<<grid, 256>>>(iters, 0.25, sink); }, (double)warps * iters * CHAINS * 2.0 * 16 * 8 * 16);
    timed("DFMA fma.rn.f64 (CUDA cores)", [&] { k_dfma<<<grid, 256>>>(iters * 8, 0.25, sink); }, (double)warps * 32 * (iters * 8.0) * CHAINS * 2.0);
    return 0;
}
