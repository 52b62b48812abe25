// partition.cu -- K6 hash partitioner for the multi-rank exchange.
//
// K6 replaces Spark's `Exchange hashpartitioning(team_id, 200)` that sits between the partial
// and the final HashAggregate and in front of the join (implicit in ComputeFeatures.scala:
// 148-150).  Rows are routed to rank = mix(key) % nparts.  Two passes over
// the key column and no host round trip in between:
//   (1) k_part_count    per-destination row counts: lanes of a warp that share a destination are found with
//                       MATCH.ANY and one of them bumps the CTA's shared-memory counter once for all of them
//                       (256 threads hammering <= 8 shared counters was the cost of the first version);
//       k_part_offsets  exclusive scan of the counts -> segment start of every destination (one small block)
//   (2) k_part_scatter  every CTA takes 2048-row tiles: the same MATCH.ANY ranking gives each row its position
//                       inside the tile's run for its destination, the tile's runs are reserved with one global
//                       atomic per destination, and each column is staged through shared memory (read in row
//                       order, parked at the row's position, written in position order) so that both the reads and
//                       the writes of every column are full 32-byte sectors.
// The result is laid out exactly as an all-to-all wants it: nparts contiguous segments, sizes on the device.
// Algorithmic bytes: key column once more (count pass) + 2x (key + payload) (read + write).
#include <algorithm>
#include <numeric>

#include "common.cuh"

namespace {

constexpr int PART_THREADS = 256;
constexpr int PART_ITEMS = 8;
constexpr int PART_TILE = PART_THREADS * PART_ITEMS;
constexpr int PART_MAX = 1024;
constexpr int PART_MAX_COLS = 16;

struct PartCol {
    const void *src;
    const uint8_t *src_valid;
    void *dst;
    unsigned *dst_valid;
    int elem;   // 4 or 8
};

struct PartParams {
    int64_t n;
    int nparts;
    int key_is64;
    int ncols;            // col[0] is the key
    PartCol col[PART_MAX_COLS];
};

// rank = mix(key) % nparts; the 64-bit division is kept off the int32 path, a power-of-two nparts is a mask (same value)
__device__ __forceinline__ int part_of(const PartParams &p, int64_t i) {
    if (!dev_bit_valid(p.col[0].src_valid, i)) return 0;      // null keys travel to rank 0
    const unsigned np = (unsigned)p.nparts;
    if (!p.key_is64) {
        const uint32_t h = mix32((uint32_t)((const int32_t *)p.col[0].src)[i]);
        return (int)((np & (np - 1)) == 0 ? (h & (np - 1)) : (h % np));
    }
    const unsigned long long h = mix64((unsigned long long)((const long long *)p.col[0].src)[i]);
    return (int)((np & (np - 1)) == 0 ? (h & (unsigned long long)(np - 1)) : (h % (unsigned long long)np));
}

__device__ __forceinline__ int part_of_key(const PartParams &p, long long k, bool valid) {
    if (!valid) return 0;
    const unsigned np = (unsigned)p.nparts;
    if (!p.key_is64) {
        const uint32_t h = mix32((uint32_t)(int32_t)k);
        return (int)((np & (np - 1)) == 0 ? (h & (np - 1)) : (h % np));
    }
    const unsigned long long h = mix64((unsigned long long)k);
    return (int)((np & (np - 1)) == 0 ? (h & (unsigned long long)(np - 1)) : (h % (unsigned long long)np));
}

__global__ void __launch_bounds__(PART_THREADS) k_part_count(const __grid_constant__ PartParams p,
                                                             unsigned long long *__restrict__ counts) {
    __shared__ unsigned hist[PART_MAX];
    for (int i = threadIdx.x; i < p.nparts; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    const unsigned lane = threadIdx.x & 31;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t base = blockIdx.x * (int64_t)blockDim.x; base < p.n; base += stride) {   // whole warps stay converged
        const int64_t i = base + threadIdx.x;
        const int dest = i < p.n ? part_of(p, i) : p.nparts;
        const unsigned peers = __match_any_sync(0xffffffffu, dest);
        if (dest < p.nparts && lane == (unsigned)(__ffs(peers) - 1)) atomicAdd(&hist[dest], (unsigned)__popc(peers));
    }
    __syncthreads();
    for (int i = threadIdx.x; i < p.nparts; i += blockDim.x)
        if (hist[i]) atomicAdd(&counts[i], (unsigned long long)hist[i]);
}

// cursors[p] = sum of counts[q] for q < p (nparts <= 1024: one block)
__global__ void __launch_bounds__(1024) k_part_offsets(const unsigned long long *__restrict__ counts, int nparts,
                                                       unsigned long long *__restrict__ cursors) {
    __shared__ unsigned long long s[PART_MAX];
    const int t = threadIdx.x;
    s[t] = t < nparts ? counts[t] : 0ULL;
    __syncthreads();
    for (int o = 1; o < PART_MAX; o <<= 1) {
        const unsigned long long v = t >= o ? s[t - o] : 0ULL;
        __syncthreads();
        s[t] += v;
        __syncthreads();
    }
    if (t < nparts) cursors[t] = s[t] - counts[t];
}

// One tile = 2048 rows.  (a) every row gets its position inside the tile's destination-grouped order: MATCH.ANY finds the
// lanes of a warp that share a destination, one of them reserves their places in the CTA's run for that destination;
// (b) the runs are laid out (exclusive scan over the destinations) and reserved in the output with one global atomic
// per destination; (c) column by column the tile is read in row order (coalesced), parked in shared memory at the
// row's position, and written out in position order -- consecutive threads write consecutive rows of a run, so every
// column leaves the CTA as full sectors.
template <bool SMALL>   // SMALL: nparts <= 32, the scan over the destinations is one warp's shuffle
__global__ void __launch_bounds__(PART_THREADS, 6) k_part_scatter(const __grid_constant__ PartParams p,
                                                               unsigned long long *__restrict__ cursors, void *const *__restrict__ dst_tab) {
    // dst_tab (optional): one destination pointer per (column, destination), [c * nparts + d] -- the exchange over peer
    // memory hands in pointers into the receive buffers of the other GPUs, so the rows cross NVLink as this kernel's
    // stores (runs of a few hundred rows per destination and tile) and never rest in a send buffer
    __shared__ unsigned hist[SMALL ? 32 : PART_MAX];                 // rows of this tile per destination, then the tile-local run start
    __shared__ unsigned long long base[SMALL ? 32 : PART_MAX];       // first output row of this tile's run, minus the tile-local run start
    __shared__ unsigned short pdest[PART_TILE];                      // position in the tile's destination-grouped order -> destination
    __shared__ __align__(16) unsigned long long stage[PART_TILE];    // one column of the tile, in position order
    __shared__ unsigned scan_tmp[SMALL ? 1 : PART_THREADS];
    const unsigned lane = threadIdx.x & 31;
    for (int64_t tile = blockIdx.x; tile * PART_TILE < p.n; tile += gridDim.x) {
        for (int i = threadIdx.x; i < p.nparts; i += blockDim.x) hist[i] = 0;
        __syncthreads();
        const int64_t tile_base = tile * PART_TILE;
        const int rows = (int)(p.n - tile_base < PART_TILE ? p.n - tile_base : PART_TILE);
        // all key loads of the tile are issued before the first one is consumed
        long long kv[PART_ITEMS];
        bool live[PART_ITEMS];
#pragma unroll
        for (int j = 0; j < PART_ITEMS; j++) {
            const int r = j * PART_THREADS + threadIdx.x;
            live[j] = r < rows;
            kv[j] = 0;
            if (live[j]) kv[j] = p.key_is64 ? ((const long long *)p.col[0].src)[tile_base + r] : (long long)((const int32_t *)p.col[0].src)[tile_base + r];
        }
        int dest[PART_ITEMS];
        unsigned pos[PART_ITEMS];
#pragma unroll
        for (int j = 0; j < PART_ITEMS; j++) {
            dest[j] = live[j] ? part_of_key(p, kv[j], dev_bit_valid(p.col[0].src_valid, tile_base + j * PART_THREADS + threadIdx.x)) : p.nparts;
            const unsigned peers = __match_any_sync(0xffffffffu, dest[j]);
            const int leader = __ffs(peers) - 1;
            unsigned first = 0;
            if (dest[j] < p.nparts && (int)lane == leader) first = atomicAdd(&hist[dest[j]], (unsigned)__popc(peers));
            first = __shfl_sync(0xffffffffu, first, leader);
            pos[j] = first + (unsigned)__popc(peers & ((1u << lane) - 1u));
        }
        __syncthreads();
        if (SMALL) {
            if (threadIdx.x < 32) {
                const unsigned c = (int)lane < p.nparts ? hist[lane] : 0u;
                unsigned x = c;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const unsigned y = __shfl_up_sync(0xffffffffu, x, o); if ((int)lane >= o) x += y; }
                if ((int)lane < p.nparts) {
                    const unsigned run = x - c;
                    hist[lane] = run;
                    base[lane] = (c ? atomicAdd(&cursors[lane], (unsigned long long)c) : 0ULL) - (unsigned long long)run;
                }
            }
        } else {
            unsigned c[4], tot = 0;
#pragma unroll
            for (int q = 0; q < 4; q++) { const int d = threadIdx.x * 4 + q; c[q] = d < p.nparts ? hist[d] : 0u; tot += c[q]; }
            scan_tmp[threadIdx.x] = tot;
            __syncthreads();
            for (int o = 1; o < PART_THREADS; o <<= 1) {
                const unsigned v = (int)threadIdx.x >= o ? scan_tmp[threadIdx.x - o] : 0u;
                __syncthreads();
                scan_tmp[threadIdx.x] += v;
                __syncthreads();
            }
            unsigned run = scan_tmp[threadIdx.x] - tot;
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const int d = threadIdx.x * 4 + q;
                if (d < p.nparts) {
                    hist[d] = run;
                    base[d] = (c[q] ? atomicAdd(&cursors[d], (unsigned long long)c[q]) : 0ULL) - (unsigned long long)run;
                    run += c[q];
                }
            }
        }
        __syncthreads();
#pragma unroll
        for (int j = 0; j < PART_ITEMS; j++) {
            if (dest[j] >= p.nparts) continue;
            pos[j] += hist[dest[j]];
            pdest[pos[j]] = (unsigned short)dest[j];
        }
        for (int c = 0; c < p.ncols; c++) {
            const PartCol &pc = p.col[c];
            __syncthreads();                       // pdest complete (first column) / the previous column has left the stage
            if (pc.elem == 4) {
                const int32_t *src = (const int32_t *)pc.src + tile_base;
#pragma unroll
                for (int j = 0; j < PART_ITEMS; j++) {
                    const int r = j * PART_THREADS + threadIdx.x;
                    if (r < rows) ((int32_t *)stage)[pos[j]] = src[r];
                }
                __syncthreads();
#pragma unroll
                for (int j = 0; j < PART_ITEMS; j++) {
                    const int q = j * PART_THREADS + threadIdx.x;
                    if (q < rows) {
                        const int dd = pdest[q];
                        int32_t *dp = dst_tab ? (int32_t *)dst_tab[c * p.nparts + dd] : (int32_t *)pc.dst;
                        dp[base[dd] + (unsigned long long)q] = ((const int32_t *)stage)[q];
                    }
                }
            } else {
                const unsigned long long *src = (const unsigned long long *)pc.src + tile_base;
#pragma unroll
                for (int j = 0; j < PART_ITEMS; j++) {
                    const int r = j * PART_THREADS + threadIdx.x;
                    if (r < rows) stage[pos[j]] = src[r];
                }
                __syncthreads();
#pragma unroll
                for (int j = 0; j < PART_ITEMS; j++) {
                    const int q = j * PART_THREADS + threadIdx.x;
                    if (q < rows) {
                        const int dd = pdest[q];
                        unsigned long long *dp = dst_tab ? (unsigned long long *)dst_tab[c * p.nparts + dd] : (unsigned long long *)pc.dst;
                        dp[base[dd] + (unsigned long long)q] = stage[q];
                    }
                }
            }
            if (pc.dst_valid) {                    // bitmaps (rare: the exchanges ship flags as columns): bit by bit
#pragma unroll
                for (int j = 0; j < PART_ITEMS; j++) {
                    const int r = j * PART_THREADS + threadIdx.x;
                    if (r >= rows || !dev_bit_valid(pc.src_valid, tile_base + r)) continue;
                    const unsigned long long o = base[dest[j]] + (unsigned long long)pos[j];
                    atomicOr(&pc.dst_valid[o >> 5], 1u << (o & 31));
                }
            }
        }
        __syncthreads();
    }
}

} // namespace

// K6 without any host round trip: counts (device, nparts u64) receives the rows per destination, *out the columns
// regrouped by destination (column 0 = keys).  For callers already inside an ABI call (csrc/dist.cu).
int32_t fsx_internal_partition(fsx_ctx *ctx, const fsx_col *key, const fsx_col *const *payload, int32_t npayload, int32_t nparts,
                               unsigned long long *d_counts, fsx_table **out) {
    *out = nullptr;
    FSX_REQUIRE(ctx, key->dtype == FSX_I32 || key->dtype == FSX_I64, FSX_E_UNSUPPORTED,
                "fsx_partition_by_key: key dtype %d (only i32/i64 keys)", key->dtype);
    FSX_REQUIRE(ctx, nparts >= 1 && nparts <= PART_MAX, FSX_E_INVALID, "fsx_partition_by_key: nparts %d outside [1,%d]", nparts, (int)PART_MAX);
    FSX_REQUIRE(ctx, npayload >= 0 && npayload < PART_MAX_COLS && (npayload == 0 || payload), FSX_E_INVALID,
                "fsx_partition_by_key: npayload %d outside [0,%d)", npayload, (int)PART_MAX_COLS);
    PartParams p;
    memset(&p, 0, sizeof p);
    p.n = key->n; p.nparts = nparts; p.key_is64 = key->dtype == FSX_I64; p.ncols = 1 + npayload;
    fsx_table *t = new fsx_table();
    t->nrows = key->n;
    auto fail = [&](int32_t rc) { fsx_table_free(ctx, t); return rc; };
    for (int c = 0; c < p.ncols; c++) {
        const fsx_col *src = c == 0 ? key : payload[c - 1];
        if (!src || src->n != key->n) return fail(fsx_fail(ctx, FSX_E_INVALID, "fsx_partition_by_key: payload column %d is NULL or has a different length", c - 1));
        fsx_col *dst = nullptr;
        int32_t rc = fsx_col_new(ctx, src->dtype, src->n, src->validity != nullptr, &dst);
        if (rc != FSX_OK) return fail(rc);
        t->cols.push_back(dst);
        if (dst->validity) FSX_ASYNC(ctx, cudaMemsetAsync(dst->validity, 0, (size_t)((src->n + 127) / 128) * 16, ctx->stream));
        p.col[c].src = src->data; p.col[c].src_valid = src->validity;
        p.col[c].dst = dst->data; p.col[c].dst_valid = (unsigned *)dst->validity;
        p.col[c].elem = (int)fsx_dtype_size(src->dtype);
    }
    unsigned long long *d_cursors = nullptr;
    int32_t rc = fsx_dev_alloc(ctx, (size_t)nparts * 8, (void **)&d_cursors);
    if (rc != FSX_OK) return fail(rc);
    cudaError_t e = cudaMemsetAsync(d_counts, 0, (size_t)nparts * 8, ctx->stream);
    if (e == cudaSuccess && key->n > 0) {
        k_part_count<<<fsx_grid_for(ctx, key->n, PART_TILE, 8), PART_THREADS, 0, ctx->stream>>>(p, d_counts);
        k_part_offsets<<<1, 1024, 0, ctx->stream>>>(d_counts, nparts, d_cursors);
        if (nparts <= 32) k_part_scatter<true><<<fsx_grid_for(ctx, key->n, PART_TILE, 6), PART_THREADS, 0, ctx->stream>>>(p, d_cursors, nullptr);
        else k_part_scatter<false><<<fsx_grid_for(ctx, key->n, PART_TILE, 5), PART_THREADS, 0, ctx->stream>>>(p, d_cursors, nullptr);
        ctx->launches += 3;
        e = cudaGetLastError();
    }
    fsx_dev_free(ctx, d_cursors);
    if (e != cudaSuccess) { fsx_table_free(ctx, t); FSX_CUDA(ctx, e); }
    *out = t;
    return FSX_OK;
}

// The two halves of K6 for the exchange over peer memory (csrc/dist.cu): rows per destination, then -- once every rank
// knows where its rows go in the receivers' buffers -- the scatter straight to those addresses.
int32_t fsx_internal_partition_count(fsx_ctx *ctx, const fsx_col *key, int32_t nparts, unsigned long long *d_counts) {
    FSX_REQUIRE(ctx, nparts >= 1 && nparts <= PART_MAX, FSX_E_INVALID, "partition: nparts %d outside [1,%d]", nparts, (int)PART_MAX);
    PartParams p;
    memset(&p, 0, sizeof p);
    p.n = key->n; p.nparts = nparts; p.key_is64 = key->dtype == FSX_I64; p.ncols = 1;
    p.col[0].src = key->data; p.col[0].src_valid = key->validity; p.col[0].elem = (int)fsx_dtype_size(key->dtype);
    FSX_CUDA(ctx, cudaMemsetAsync(d_counts, 0, (size_t)nparts * 8, ctx->stream));
    if (key->n > 0) FSX_LAUNCH(ctx, k_part_count, fsx_grid_for(ctx, key->n, PART_TILE, 8), PART_THREADS, 0, p, d_counts);
    return FSX_OK;
}

// cols[0] is the key; dst_tab (device): (1 + npayload) x nparts pointers, [c * nparts + d] = where column c's rows for
// destination d start; d_cursors (device, nparts u64): zeroed by the caller's memset below
int32_t fsx_internal_partition_scatter_to(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols, int32_t nparts, void *const *dst_tab,
                                          unsigned long long *d_cursors) {
    FSX_REQUIRE(ctx, ncols >= 1 && ncols <= PART_MAX_COLS, FSX_E_INVALID, "partition: %d columns outside [1,%d]", ncols, (int)PART_MAX_COLS);
    PartParams p;
    memset(&p, 0, sizeof p);
    p.n = cols[0]->n; p.nparts = nparts; p.key_is64 = cols[0]->dtype == FSX_I64; p.ncols = ncols;
    for (int c = 0; c < ncols; c++) {
        p.col[c].src = cols[c]->data; p.col[c].src_valid = c == 0 ? cols[c]->validity : nullptr;
        p.col[c].dst = nullptr; p.col[c].dst_valid = nullptr;
        p.col[c].elem = (int)fsx_dtype_size(cols[c]->dtype);
    }
    FSX_CUDA(ctx, cudaMemsetAsync(d_cursors, 0, (size_t)nparts * 8, ctx->stream));
    if (p.n > 0) {
        if (nparts <= 32) FSX_LAUNCH(ctx, k_part_scatter<true>, fsx_grid_for(ctx, p.n, PART_TILE, 6), PART_THREADS, 0, p, d_cursors, dst_tab);
        else FSX_LAUNCH(ctx, k_part_scatter<false>, fsx_grid_for(ctx, p.n, PART_TILE, 5), PART_THREADS, 0, p, d_cursors, dst_tab);
    }
    return FSX_OK;
}

extern "C" int32_t fsx_partition_by_key(fsx_ctx *ctx, const fsx_col *key, const fsx_col *const *payload,
                                        int32_t npayload, int32_t nparts, int64_t *counts_out,
                                        fsx_table **out) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, key && counts_out && out, FSX_E_INVALID, "fsx_partition_by_key: NULL argument");
    *out = nullptr;
    FSX_REQUIRE(ctx, nparts >= 1 && nparts <= PART_MAX, FSX_E_INVALID, "fsx_partition_by_key: nparts %d outside [1,%d]", nparts, (int)PART_MAX);
    unsigned long long *d_counts = nullptr;
    FSX_CHECK(fsx_dev_alloc(ctx, (size_t)nparts * 8, (void **)&d_counts));
    fsx_table *t = nullptr;
    int32_t rc = fsx_internal_partition(ctx, key, payload, npayload, nparts, d_counts, &t);
    cudaError_t e = cudaSuccess;
    if (rc == FSX_OK) {
        FSX_CUDA(ctx, cudaEventRecord(ctx->ev_d, ctx->stream));
        e = cudaMemcpyAsync(counts_out, d_counts, (size_t)nparts * 8, cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    }
    fsx_dev_free(ctx, d_counts);
    if (rc != FSX_OK) return rc;
    if (e != cudaSuccess) { fsx_table_free(ctx, t); FSX_CUDA(ctx, e); }
    *out = t;
    FSX_END();
}
