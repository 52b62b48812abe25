// Microbenchmark for the fused probe + aggregate of config 4: 125 M fact rows (uniform int32 keys over a
// 1 M-row packed dimension table), 1000 groups.  Variants isolate the probe and the aggregation styles.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1);} } while (0)
__device__ __host__ __forceinline__ uint32_t mix32(uint32_t h) { h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16; return h; }
__device__ __forceinline__ uint64_t mixq(uint64_t seed, uint64_t idx) {
    uint64_t z = seed + (idx + 1) * 0x9E3779B97F4A7C15ULL; z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL; z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL; return z ^ (z >> 31);
}
__global__ void gen(int *k, double *v, int64_t n, int G, int skew) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        uint64_t r = mixq(2, i);
        int key = (int)(((r >> 32) * (uint64_t)G) >> 32);
        if (skew && (r & 7) == 0) key = 12345;          // 12.5 % of the rows share one key
        k[i] = key; v[i] = (double)(mixq(3, i) >> 40);
    }
}
__global__ void build(unsigned long long *slots, unsigned long long mask, int G) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < G; i += gridDim.x * blockDim.x) {
        int g = (int)(mixq(5, i) % 1000);
        unsigned long long packed = ((unsigned long long)(uint32_t)g << 32) | (uint32_t)i, s = mix32(i) & mask;
        while (atomicCAS(&slots[s], ~0ULL, packed) != ~0ULL) s = (s + 1) & mask;
    }
}
__device__ __forceinline__ int probe(const unsigned long long *slots, unsigned long long mask, int k, unsigned long long first, unsigned long long s) {
    unsigned long long cur = first;
    while (cur != ~0ULL) { if ((uint32_t)cur == (uint32_t)k) return (int)(cur >> 32); s = (s + 1) & mask; cur = __ldg(&slots[s]); }
    return -1;
}
// V: 0 probe only (sum of groups), 1 shared atomics on one CTA table, 2 warp-private + match_any, 3 warp-private + smem atomics(u32 count + f64)
template <int V, int U>
__global__ void k(const int *__restrict__ fk, const double *__restrict__ fv, int64_t n, const unsigned long long *__restrict__ slots,
                  unsigned long long mask, int range, double *gsum, unsigned long long *gcnt) {
    extern __shared__ unsigned char smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    const bool priv = V >= 2;
    double *w_sum = (double *)(smem + (priv ? (size_t)warp * range * 12 : 0));
    unsigned *w_cnt = (unsigned *)(w_sum + range);
    if (V >= 1) { for (int i = priv ? lane : threadIdx.x; i < range; i += priv ? 32 : blockDim.x) { w_sum[i] = 0; w_cnt[i] = 0; } __syncthreads(); }
    long long acc = 0;
    const int64_t stride = (int64_t)gridDim.x * nwarps * 32 * U;
    for (int64_t b = ((int64_t)blockIdx.x * nwarps + warp) * 32 * U; b < n; b += stride) {
        int kk[U]; double v[U]; unsigned long long s[U], e[U]; int g[U];
#pragma unroll
        for (int u = 0; u < U; u++) { int64_t row = b + u * 32 + lane; bool ok = row < n; kk[u] = ok ? fk[row] : 0; v[u] = ok ? fv[row] : 0; s[u] = mix32(kk[u]) & mask; }
#pragma unroll
        for (int u = 0; u < U; u++) e[u] = __ldg(&slots[s[u]]);
#pragma unroll
        for (int u = 0; u < U; u++) g[u] = (b + u * 32 + lane < n) ? probe(slots, mask, kk[u], e[u], s[u]) : -1;
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (V == 0) acc += g[u] + (long long)v[u];
            if (V == 1 || V == 3) { if (g[u] >= 0) { atomicAdd(&w_cnt[g[u]], 1u); atomicAdd(&w_sum[g[u]], v[u]); } }
            if (V == 2) {
                const unsigned peers = __match_any_sync(0xffffffffu, g[u]);
                unsigned m = peers;
                const int iters = __reduce_max_sync(0xffffffffu, __popc(m));
                double sum = 0.0;
                for (int it = 0; it < iters; it++) { int src = m ? __ffs(m) - 1 : lane; double x = __shfl_sync(0xffffffffu, v[u], src); if (m) { sum += x; m &= m - 1; } }
                if (g[u] >= 0 && lane == __ffs(peers) - 1) { w_cnt[g[u]] += __popc(peers); w_sum[g[u]] += sum; }
                __syncwarp();
            }
        }
    }
    if (V == 0) { if (acc == 123456789) gcnt[0] = acc; return; }
    __syncthreads();
    for (int i = threadIdx.x; i < range; i += blockDim.x) {
        double sum = 0; unsigned long long c = 0;
        for (int w = 0; w < (priv ? nwarps : 1); w++) { const double *ws = (const double *)(smem + (size_t)w * range * 12); sum += ws[i]; c += ((const unsigned *)(ws + range))[i]; }
        if (c) { atomicAdd(&gcnt[i], c); atomicAdd(&gsum[i], sum); }
    }
}
// Second family: direct-index dimension (dense primary keys: group = dir[k]) and a fixed-point aggregator whose
// shared-memory updates are native 32-bit ATOMS.ADD (three limbs with carries) instead of 64-bit CAS loops.
// V: 0 probe only, 1 f64 CAS + u32 count, 4 fixed-point limbs + u32 count
template <int V, int U, bool DIRECT>
__global__ void k2(const int *__restrict__ fk, const double *__restrict__ fv, int64_t n, const unsigned long long *__restrict__ slots,
                   unsigned long long mask, const int *__restrict__ dir, int range, double scale, double *gsum, unsigned long long *gcnt,
                   unsigned long long *glo, unsigned long long *ghi) {
    extern __shared__ unsigned char smem[];
    double *s_sum = (double *)smem;                      // V1
    unsigned *l0 = (unsigned *)smem, *l1 = l0 + range, *l2 = l1 + range;   // V4
    unsigned *s_cnt = (unsigned *)(smem + (size_t)range * 12);
    if (V >= 1) { for (int i = threadIdx.x; i < range * 4; i += blockDim.x) ((unsigned *)smem)[i] = 0; __syncthreads(); }
    long long acc = 0;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x * U;
    for (int64_t b = (int64_t)blockIdx.x * blockDim.x * U; b < n; b += stride) {
        int kk[U]; double v[U]; int g[U]; unsigned long long s[U], e[U];
#pragma unroll
        for (int u = 0; u < U; u++) { int64_t row = b + (int64_t)u * blockDim.x + threadIdx.x; bool ok = row < n; kk[u] = ok ? fk[row] : -1; v[u] = ok ? fv[row] : 0; s[u] = mix32(kk[u]) & mask; }
        if (DIRECT) {
#pragma unroll
            for (int u = 0; u < U; u++) g[u] = kk[u] >= 0 ? __ldg(&dir[kk[u]]) : -1;
        } else {
#pragma unroll
            for (int u = 0; u < U; u++) e[u] = __ldg(&slots[s[u]]);
#pragma unroll
            for (int u = 0; u < U; u++) g[u] = kk[u] >= 0 ? probe(slots, mask, kk[u], e[u], s[u]) : -1;
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (V == 0) acc += g[u] + (long long)v[u];
            if (g[u] < 0) continue;
            if (V == 1) { atomicAdd(&s_cnt[g[u]], 1u); atomicAdd(&s_sum[g[u]], v[u]); }
            if (V == 4) {
                const long long q = __double2ll_rn(v[u] * scale);
                const unsigned lo = (unsigned)q, mid = (unsigned)((unsigned long long)q >> 32), hi = q < 0 ? 0xFFFFFFFFu : 0u;
                atomicAdd(&s_cnt[g[u]], 1u);
                unsigned old = atomicAdd(&l0[g[u]], lo);
                const unsigned c0 = (old + lo) < old;
                const unsigned a1 = mid + c0;
                unsigned c1 = a1 < c0;
                if (a1) { old = atomicAdd(&l1[g[u]], a1); c1 |= (old + a1) < old; }
                const unsigned a2 = hi + c1;
                if (a2) atomicAdd(&l2[g[u]], a2);
            }
        }
    }
    if (V == 0) { if (acc == 123456789) gcnt[0] = acc; return; }
    __syncthreads();
    for (int i = threadIdx.x; i < range; i += blockDim.x) {
        const unsigned c = s_cnt[i];
        if (!c) continue;
        atomicAdd(&gcnt[i], (unsigned long long)c);
        if (V == 1) atomicAdd(&gsum[i], s_sum[i]);
        if (V == 4) {   // 96-bit signed partial -> 128-bit global accumulator (exact, order independent)
            const unsigned long long lo = ((unsigned long long)l1[i] << 32) | l0[i];
            const unsigned long long hi = (unsigned long long)(long long)(int)l2[i];
            const unsigned long long old = atomicAdd(&glo[i], lo);
            atomicAdd(&ghi[i], hi + ((old + lo) < old ? 1ULL : 0ULL));
        }
    }
}
__global__ void build_dir(int *dir, int G) { for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < G; i += gridDim.x * blockDim.x) dir[i] = (int)(mixq(5, i) % 1000); }

int main() {
    const int64_t n = 125000000; const int G = 1000000, range = 1000;
    int *fk; double *fv; unsigned long long *slots, *gcnt; double *gsum;
    const unsigned long long cap = 1 << 21;
    CK(cudaMalloc(&fk, n * 4)); CK(cudaMalloc(&fv, n * 8)); CK(cudaMalloc(&slots, cap * 8)); CK(cudaMalloc(&gcnt, range * 8)); CK(cudaMalloc(&gsum, range * 8));
    CK(cudaMemset(slots, 0xFF, cap * 8));
    build<<<148 * 4, 256>>>(slots, cap - 1, G);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int skew = 0; skew < 2; skew++) {
        gen<<<148 * 8, 256>>>(fk, fv, n, G, skew);
        printf("--- %s keys\n", skew ? "skewed (12.5%% one key)" : "uniform");
#define RUN(name, V, U, threads, ctas_per_sm, smem) do { \
    CK(cudaFuncSetAttribute(k<V, U>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(smem))); float best = 1e9; \
    for (int it = 0; it < 3; it++) { cudaMemset(gcnt, 0, range * 8); cudaMemset(gsum, 0, range * 8); cudaEventRecord(e0); \
        k<V, U><<<148 * (ctas_per_sm), threads, smem>>>(fk, fv, n, slots, cap - 1, range, gsum, gcnt); cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); \
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms; } \
    unsigned long long c0; cudaMemcpy(&c0, gcnt, 8, cudaMemcpyDeviceToHost); \
    printf("%-58s %7.3f ms %6.1f Grows/s (cnt[0]=%llu)\n", name, best, n / best / 1e6, c0); } while (0)
        {
            static int *dir = nullptr; static unsigned long long *glo, *ghi;
            if (!dir) { CK(cudaMalloc(&dir, G * 4)); build_dir<<<148 * 4, 256>>>(dir, G); CK(cudaMalloc(&glo, range * 8)); CK(cudaMalloc(&ghi, range * 8)); }
            const double scale = 4096.0;   // values are integers < 2^24 here
#define RUN2(name, V, U, D, threads, ctas_per_sm) do { \
    CK(cudaFuncSetAttribute(k2<V, U, D>, cudaFuncAttributeMaxDynamicSharedMemorySize, range * 16)); float best = 1e9; \
    for (int it = 0; it < 3; it++) { cudaMemset(gcnt, 0, range * 8); cudaMemset(gsum, 0, range * 8); cudaMemset(glo, 0, range * 8); cudaMemset(ghi, 0, range * 8); cudaEventRecord(e0); \
        k2<V, U, D><<<148 * (ctas_per_sm), threads, range * 16>>>(fk, fv, n, slots, cap - 1, dir, range, scale, gsum, gcnt, glo, ghi); cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); \
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms; } \
    unsigned long long c0, lo0, hi0; double s0; cudaMemcpy(&c0, gcnt + 7, 8, cudaMemcpyDeviceToHost); cudaMemcpy(&s0, gsum + 7, 8, cudaMemcpyDeviceToHost); \
    cudaMemcpy(&lo0, glo + 7, 8, cudaMemcpyDeviceToHost); cudaMemcpy(&hi0, ghi + 7, 8, cudaMemcpyDeviceToHost); \
    printf("%-58s %7.3f ms %6.1f Grows/s (cnt[7]=%llu sum[7]=%.1f fx[7]=%.1f)\n", name, best, n / best / 1e6, c0, s0, ((double)(long long)hi0 * 18446744073709551616.0 + (double)lo0) / scale); } while (0)
            RUN2("K2 hash  probe only U=4 256x8", 0, 4, false, 256, 8);
            RUN2("K2 direct probe only U=4 256x8", 0, 4, true, 256, 8);
            RUN2("K2 direct probe only U=8 256x8", 0, 8, true, 256, 8);
            RUN2("K2 hash  + f64 CAS U=4 256x8", 1, 4, false, 256, 8);
            RUN2("K2 direct + f64 CAS U=4 256x8", 1, 4, true, 256, 8);
            RUN2("K2 hash  + limbs U=4 256x8", 4, 4, false, 256, 8);
            RUN2("K2 direct + limbs U=4 256x8", 4, 4, true, 256, 8);
            RUN2("K2 direct + limbs U=8 256x8", 4, 8, true, 256, 8);
            RUN2("K2 direct + limbs U=4 512x4", 4, 4, true, 512, 4);
            RUN2("K2 direct + limbs U=4 256x6", 4, 4, true, 256, 6);
        }
        RUN("V0 probe only, U=4, 256 thr x 8 CTA/SM", 0, 4, 256, 8, 0);
        RUN("V1 shared atomics, one table/CTA, U=4, 256 thr x 8 CTA/SM", 1, 4, 256, 8, range * 12);
        RUN("V1 shared atomics, one table/CTA, U=4, 1024 thr x 2 CTA/SM", 1, 4, 1024, 2, range * 12);
        RUN("V2 warp-private + match_any, U=4, 256 thr x 2", 2, 4, 256, 2, 8 * range * 12);
    }
    return 0;
}
