// scan.cuh -- device-wide exclusive prefix sum over u64 (three small kernels: per-tile sums,
// scan of the tile sums in one block, per-tile rescan with the carried offset).  Used by the
// join (match offsets) and the partitioner.  Deterministic.
#pragma once
#include "common.cuh"

namespace {

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

__device__ __forceinline__ unsigned long long block_excl_scan_256(unsigned long long v,
                                                                  unsigned long long *total) {
    __shared__ unsigned long long wsum[SCAN_THREADS / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned long long x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        unsigned long long y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
    }
    if (lane == 31) wsum[warp] = x;
    __syncthreads();
    unsigned long long before = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < SCAN_THREADS / 32; w++) {
        unsigned long long t = wsum[w];
        before += w < warp ? t : 0;
        tot += t;
    }
    __syncthreads();
    *total = tot;
    return before + x - v;
}

// in: u64 or u32 counts; tile_sums[tile]
template <typename CntT>
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_tile_sums(const CntT *__restrict__ in,
                                                                 unsigned long long n,
                                                                 unsigned long long *__restrict__ tile_sums) {
    unsigned long long base = (unsigned long long)blockIdx.x * SCAN_TILE + (unsigned long long)threadIdx.x * SCAN_ITEMS;
    unsigned long long s = 0;
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; j++) s += (base + j < n) ? (unsigned long long)in[base + j] : 0ULL;
    unsigned long long tot;
    (void)block_excl_scan_256(s, &tot);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(1024) k_scan_single_block(unsigned long long *data, unsigned long long n,
                                                            unsigned long long *total_out) {
    __shared__ unsigned long long warp_tot[32];
    __shared__ unsigned long long carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (unsigned long long base = 0; base < n; base += 1024) {
        unsigned long long i = base + threadIdx.x;
        unsigned long long v = i < n ? data[i] : 0, x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            unsigned long long y = __shfl_up_sync(0xffffffffu, x, o);
            if ((threadIdx.x & 31) >= o) x += y;
        }
        if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = x;
        __syncthreads();
        if (threadIdx.x < 32) {
            unsigned long long w = warp_tot[threadIdx.x], wx = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                unsigned long long y = __shfl_up_sync(0xffffffffu, wx, o);
                if (threadIdx.x >= o) wx += y;
            }
            warp_tot[threadIdx.x] = wx - w;
        }
        __syncthreads();
        unsigned long long excl = carry + warp_tot[threadIdx.x >> 5] + x - v;
        if (i < n) data[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total_out = carry;
}

template <typename CntT>
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_apply(const CntT *__restrict__ in, unsigned long long n,
                                                             const unsigned long long *__restrict__ tile_offsets,
                                                             unsigned long long *__restrict__ out) {
    unsigned long long base = (unsigned long long)blockIdx.x * SCAN_TILE + (unsigned long long)threadIdx.x * SCAN_ITEMS;
    unsigned long long v[SCAN_ITEMS], s = 0;
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; j++) { v[j] = (base + j < n) ? (unsigned long long)in[base + j] : 0ULL; s += v[j]; }
    unsigned long long tot;
    unsigned long long off = block_excl_scan_256(s, &tot) + tile_offsets[blockIdx.x];
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; j++) {
        if (base + j < n) out[base + j] = off;
        off += v[j];
    }
}

// out[i] = sum_{k<i} in[k]; total -> ctx->d_scalars[slot].  in and out may not alias unless
// CntT is u64 and the caller passes the same pointer (each thread reads its items before writing).
template <typename CntT>
int32_t exclusive_scan(fsx_ctx *ctx, const CntT *in, unsigned long long n, unsigned long long *out,
                       int scalar_slot) {
    if (n == 0) {
        FSX_CUDA(ctx, cudaMemsetAsync(ctx->d_scalars + scalar_slot, 0, 8, ctx->stream));
        return FSX_OK;
    }
    const unsigned long long ntiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    unsigned long long *tile_sums = nullptr;
    FSX_CHECK(fsx_dev_alloc(ctx, ntiles * 8, (void **)&tile_sums));
    k_scan_tile_sums<CntT><<<(unsigned)ntiles, SCAN_THREADS, 0, ctx->stream>>>(in, n, tile_sums);
    k_scan_single_block<<<1, 1024, 0, ctx->stream>>>(tile_sums, ntiles, ctx->d_scalars + scalar_slot);
    k_scan_apply<CntT><<<(unsigned)ntiles, SCAN_THREADS, 0, ctx->stream>>>(in, n, tile_sums, out);
    ctx->launches += 3;
    cudaError_t e = cudaGetLastError();
    fsx_dev_free(ctx, tile_sums);
    FSX_CUDA(ctx, e);
    return FSX_OK;
}

} // namespace
