// sort.cu -- least-significant-digit radix sort (8-bit digits) of (ordered-u64 key, u32 row index) pairs.
//
// Every column is first mapped to order-preserving unsigned keys (csrc/common.cuh ord_from_*: NaN above +inf, the
// way Spark orders doubles) minus the column's minimum, so only the bytes that actually vary are sorted: a key column
// with 50 distinct team ids costs one pass, a unix timestamp column three or four, not eight.  One pass = a per-CTA
// digit histogram, a device-wide exclusive scan of the (digit, CTA) count matrix and a stable scatter.  Stability
// inside a CTA comes from each warp owning a contiguous run of the tile and ranking equal digits with
// __match_any_sync in (round, lane) = row order; no atomics decide an output position, so the result is deterministic.
#include <vector>

#include "scan.cuh"
#include "sort.cuh"

namespace {

constexpr int RS_THREADS = 256, RS_WARPS = 8, RS_ROUNDS = 8;
constexpr int RS_WARP_RUN = 32 * RS_ROUNDS;            // 256 contiguous rows per warp
constexpr int RS_TILE = RS_THREADS * RS_ROUNDS;        // 2048 rows per CTA

template <typename T> __device__ __forceinline__ unsigned long long s_ord(T v);
template <> __device__ __forceinline__ unsigned long long s_ord<int32_t>(int32_t v) { return ord_from_i64((long long)v); }
template <> __device__ __forceinline__ unsigned long long s_ord<long long>(long long v) { return ord_from_i64(v); }
template <> __device__ __forceinline__ unsigned long long s_ord<float>(float v) { return ord_from_f64((double)v); }
template <> __device__ __forceinline__ unsigned long long s_ord<double>(double v) { return ord_from_f64(v); }

template <typename T>
__global__ void __launch_bounds__(256) k_sort_ord_minmax(const T *__restrict__ v, const uint8_t *__restrict__ valid, int64_t n, unsigned long long *scalars) {
    unsigned long long lo = ~0ULL, hi = 0ULL;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (!dev_bit_valid(valid, i)) continue;
        const unsigned long long o = s_ord<T>(v[i]);
        lo = o < lo ? o : lo;
        hi = o > hi ? o : hi;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long l2 = __shfl_xor_sync(0xffffffffu, lo, o), h2 = __shfl_xor_sync(0xffffffffu, hi, o);
        lo = l2 < lo ? l2 : lo;
        hi = h2 > hi ? h2 : hi;
    }
    if ((threadIdx.x & 31) == 0) { atomicMin(&scalars[2], lo); atomicMax(&scalars[3], hi); }
}

// keys[i] = ord(col[perm[i]]) - omin, nulls -> null_key (one above the largest value: nulls last)
template <typename T>
__global__ void __launch_bounds__(256) k_sort_make_keys(const T *__restrict__ v, const uint8_t *__restrict__ valid, const unsigned *__restrict__ perm,
                                                        int64_t n, unsigned long long omin, unsigned long long null_key, unsigned long long *keys) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const unsigned r = perm[i];
        keys[i] = dev_bit_valid(valid, r) ? s_ord<T>(v[r]) - omin : null_key;
    }
}

// keys[i] = 1 for a null row: a last one-bit pass when the values leave no code free for "null"
__global__ void __launch_bounds__(256) k_sort_null_keys(const uint8_t *__restrict__ valid, const unsigned *__restrict__ perm, int64_t n, unsigned long long *keys) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) keys[i] = dev_bit_valid(valid, perm[i]) ? 0ULL : 1ULL;
}

__global__ void k_sort_iota(unsigned *p, int64_t n) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = (unsigned)i;
}

// counts[digit * nblocks + block]
__global__ void __launch_bounds__(RS_THREADS) k_rs_hist(const unsigned long long *__restrict__ keys, int64_t n, int shift, unsigned *counts, unsigned nblocks) {
    __shared__ unsigned h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * RS_TILE;
#pragma unroll 4
    for (int r = 0; r < RS_ROUNDS; r++) {
        const int64_t i = base + (int64_t)r * RS_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&h[(unsigned)(keys[i] >> shift) & 255u], 1u);
    }
    __syncthreads();
    counts[(size_t)threadIdx.x * nblocks + blockIdx.x] = h[threadIdx.x];
}

// dynamic shared memory: the tile's (key, idx) pairs in digit order, so that every digit run leaves the CTA as one
// contiguous write instead of 4096 scattered 12-byte stores
constexpr size_t RS_STAGE_BYTES = (size_t)RS_TILE * 12;

__global__ void __launch_bounds__(RS_THREADS) k_rs_scatter(const unsigned long long *__restrict__ keys_in, const unsigned *__restrict__ idx_in, int64_t n,
                                                           int shift, const unsigned long long *__restrict__ offsets, unsigned nblocks,
                                                           unsigned long long *__restrict__ keys_out, unsigned *__restrict__ idx_out) {
    extern __shared__ __align__(16) unsigned char rs_stage[];
    unsigned long long *skeys = (unsigned long long *)rs_stage;
    unsigned *sidx = (unsigned *)(rs_stage + (size_t)RS_TILE * 8);
    __shared__ unsigned cnt[RS_WARPS][256];
    __shared__ unsigned long long goff[256];
    __shared__ unsigned dstart[256];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int w = 0; w < RS_WARPS; w++) cnt[w][threadIdx.x] = 0;
    goff[threadIdx.x] = offsets[(size_t)threadIdx.x * nblocks + blockIdx.x];
    __syncthreads();
    const int64_t tile0 = (int64_t)blockIdx.x * RS_TILE;
    const int64_t run = tile0 + (int64_t)warp * RS_WARP_RUN;
    unsigned long long k[RS_ROUNDS];
    unsigned short pos[RS_ROUNDS];
#pragma unroll
    for (int r = 0; r < RS_ROUNDS; r++) {
        const int64_t i = run + r * 32 + lane;
        const bool in = i < n;
        k[r] = in ? keys_in[i] : 0ULL;
        const unsigned d = (unsigned)(k[r] >> shift) & 255u;
        const unsigned tag = in ? d : (256u | (unsigned)lane);           // rows past the end match nobody
        const unsigned peers = __match_any_sync(0xffffffffu, tag);
        const int leader = __ffs(peers) - 1;
        unsigned old = 0;
        if (in && lane == leader) { old = cnt[warp][d]; cnt[warp][d] = old + __popc(peers); }
        old = __shfl_sync(0xffffffffu, old, leader);
        pos[r] = (unsigned short)(old + __popc(peers & ((1u << lane) - 1u)));
        __syncwarp();
    }
    __syncthreads();
    unsigned total = 0;
    {   // exclusive prefix over the warps, per digit
#pragma unroll
        for (int w = 0; w < RS_WARPS; w++) { const unsigned c = cnt[w][threadIdx.x]; cnt[w][threadIdx.x] = total; total += c; }
    }
    unsigned long long tile_total;
    dstart[threadIdx.x] = (unsigned)block_excl_scan_256((unsigned long long)total, &tile_total);   // where each digit run starts in the tile
    __syncthreads();
#pragma unroll
    for (int r = 0; r < RS_ROUNDS; r++) {
        const int64_t i = run + r * 32 + lane;
        if (i >= n) continue;
        const unsigned d = (unsigned)(k[r] >> shift) & 255u;
        const unsigned lp = dstart[d] + cnt[warp][d] + pos[r];
        skeys[lp] = k[r];
        sidx[lp] = idx_in[i];
    }
    __syncthreads();
    const unsigned tile_n = (unsigned)tile_total;
    for (unsigned t = threadIdx.x; t < tile_n; t += RS_THREADS) {
        const unsigned long long key = skeys[t];
        const unsigned d = (unsigned)(key >> shift) & 255u;
        const unsigned long long dst = goff[d] + (t - dstart[d]);
        keys_out[dst] = key;
        idx_out[dst] = sidx[t];
    }
}

template <typename T>
__global__ void __launch_bounds__(256) k_gather_u32(const T *__restrict__ src, const uint8_t *__restrict__ src_valid, const unsigned *__restrict__ perm,
                                                    int64_t n, T *__restrict__ dst, unsigned *__restrict__ dst_valid_words) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const unsigned r = perm[i];
        dst[i] = src[r];
        if (dst_valid_words && dev_bit_valid(src_valid, r)) atomicOr(&dst_valid_words[i >> 5], 1u << (i & 31));
    }
}

// stable sort of (keys, idx) by the low `bits` bits of the keys; buffers ping-pong, the result ends in *keys / *idx
int32_t radix_sort_pairs(fsx_ctx *ctx, unsigned long long **keys, unsigned **idx, unsigned long long **keys_alt, unsigned **idx_alt, int64_t n, int bits) {
    if (n <= 1 || bits <= 0) return FSX_OK;
    const unsigned nblocks = (unsigned)((n + RS_TILE - 1) / RS_TILE);
    unsigned *counts = nullptr;
    unsigned long long *offsets = nullptr;
    if (cudaFuncSetAttribute(k_rs_scatter, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)RS_STAGE_BYTES) != cudaSuccess)
        return fsx_fail(ctx, FSX_E_CUDA, "radix sort: cannot reserve %zu bytes of shared memory", RS_STAGE_BYTES);
    int32_t rc = fsx_dev_alloc(ctx, (size_t)256 * nblocks * 4, (void **)&counts);
    if (rc == FSX_OK) rc = fsx_dev_alloc(ctx, (size_t)256 * nblocks * 8, (void **)&offsets);
    for (int shift = 0; shift < bits && rc == FSX_OK; shift += 8) {
        k_rs_hist<<<nblocks, RS_THREADS, 0, ctx->stream>>>(*keys, n, shift, counts, nblocks);
        ctx->launches++;
        rc = exclusive_scan<unsigned>(ctx, counts, (unsigned long long)256 * nblocks, offsets, 0);
        if (rc != FSX_OK) break;
        k_rs_scatter<<<nblocks, RS_THREADS, RS_STAGE_BYTES, ctx->stream>>>(*keys, *idx, n, shift, offsets, nblocks, *keys_alt, *idx_alt);
        ctx->launches++;
        std::swap(*keys, *keys_alt);
        std::swap(*idx, *idx_alt);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) rc = fsx_fail(ctx, FSX_E_CUDA, "radix sort pass failed: %s", cudaGetErrorString(e));
    }
    fsx_dev_free(ctx, counts);
    fsx_dev_free(ctx, offsets);
    return rc;
}

} // namespace

int32_t fsx_internal_sort_pairs_u64(fsx_ctx *ctx, unsigned long long **keys, unsigned **idx, unsigned long long **keys_alt, unsigned **idx_alt,
                                    int64_t n, int bits) {
    return radix_sort_pairs(ctx, keys, idx, keys_alt, idx_alt, n, bits);
}

int32_t fsx_internal_argsort(fsx_ctx *ctx, const fsx_col *const *by, int nby, int64_t n, unsigned **perm_out) {
    *perm_out = nullptr;
    if (n >= (1LL << 32)) return fsx_fail(ctx, FSX_E_UNSUPPORTED, "sort: more than 2^32 - 1 rows");
    unsigned long long *keys = nullptr, *keys_alt = nullptr;
    unsigned *idx = nullptr, *idx_alt = nullptr;
    const size_t m = (size_t)(n > 0 ? n : 1);
    int32_t rc = fsx_dev_alloc(ctx, m * 8, (void **)&keys);
    if (rc == FSX_OK) rc = fsx_dev_alloc(ctx, m * 8, (void **)&keys_alt);
    if (rc == FSX_OK) rc = fsx_dev_alloc(ctx, m * 4, (void **)&idx);
    if (rc == FSX_OK) rc = fsx_dev_alloc(ctx, m * 4, (void **)&idx_alt);
    auto cleanup = [&]() { fsx_dev_free(ctx, keys); fsx_dev_free(ctx, keys_alt); fsx_dev_free(ctx, idx_alt); };
    if (rc != FSX_OK) { cleanup(); fsx_dev_free(ctx, idx); return rc; }
    const int grid = fsx_grid_for(ctx, n, 256 * 4, 8);
    if (n > 0) { k_sort_iota<<<grid, 256, 0, ctx->stream>>>(idx, n); ctx->launches++; }
    for (int c = nby - 1; c >= 0 && n > 1; c--) {                 // least significant column first
        const fsx_col *col = by[c];
        if (!col || col->n != n) { cleanup(); fsx_dev_free(ctx, idx); return fsx_fail(ctx, FSX_E_INVALID, "sort: key column %d missing or of a different length", c); }
        const unsigned long long init[8] = {0, 0, ~0ULL, 0, 0, 0, 0, 0};
        cudaError_t e = cudaMemcpyAsync(ctx->d_scalars, init, sizeof init, cudaMemcpyHostToDevice, ctx->stream);
        if (e != cudaSuccess) { cleanup(); fsx_dev_free(ctx, idx); return fsx_fail(ctx, FSX_E_CUDA, "sort: %s", cudaGetErrorString(e)); }
#define BY_DTYPE(KERN, ...) switch (col->dtype) { \
        case FSX_I32: KERN<int32_t><<<grid, 256, 0, ctx->stream>>>((const int32_t *)col->data, __VA_ARGS__); break; \
        case FSX_I64: KERN<long long><<<grid, 256, 0, ctx->stream>>>((const long long *)col->data, __VA_ARGS__); break; \
        case FSX_F32: KERN<float><<<grid, 256, 0, ctx->stream>>>((const float *)col->data, __VA_ARGS__); break; \
        default: KERN<double><<<grid, 256, 0, ctx->stream>>>((const double *)col->data, __VA_ARGS__); break; }
        BY_DTYPE(k_sort_ord_minmax, col->validity, n, ctx->d_scalars);
        ctx->launches++;
        rc = fsx_read_scalars(ctx, 0, 8);
        if (rc != FSX_OK) { cleanup(); fsx_dev_free(ctx, idx); return rc; }
        const bool any = ctx->h_scalars[2] <= ctx->h_scalars[3];
        const unsigned long long omin = any ? ctx->h_scalars[2] : 0ULL, span = any ? ctx->h_scalars[3] - ctx->h_scalars[2] : 0ULL;
        // nulls sort last: one code above the largest value; a column spanning all 2^64 codes gets a separate one-bit pass
        const bool null_pass = col->validity && span == ~0ULL;
        const unsigned long long null_key = span == ~0ULL ? ~0ULL : span + 1ULL;
        const unsigned long long top = col->validity ? null_key : span;
        int bits = 0;
        while (bits < 64 && (top >> bits) != 0ULL) bits++;
        BY_DTYPE(k_sort_make_keys, col->validity, idx, n, omin, null_key, keys);
#undef BY_DTYPE
        ctx->launches++;
        rc = radix_sort_pairs(ctx, &keys, &idx, &keys_alt, &idx_alt, n, bits);
        if (rc == FSX_OK && null_pass) {
            k_sort_null_keys<<<grid, 256, 0, ctx->stream>>>(col->validity, idx, n, keys);
            ctx->launches++;
            rc = radix_sort_pairs(ctx, &keys, &idx, &keys_alt, &idx_alt, n, 1);
        }
        if (rc != FSX_OK) { cleanup(); fsx_dev_free(ctx, idx); return rc; }
    }
    cudaError_t e = cudaGetLastError();
    cleanup();
    if (e != cudaSuccess) { fsx_dev_free(ctx, idx); return fsx_fail(ctx, FSX_E_CUDA, "sort: %s", cudaGetErrorString(e)); }
    *perm_out = idx;
    return FSX_OK;
}

int32_t fsx_internal_gather_u32(fsx_ctx *ctx, const fsx_col *src, const unsigned *perm, int64_t n, fsx_col **out) {
    fsx_col *dst = nullptr;
    int32_t rc = fsx_col_new(ctx, src->dtype, n, src->validity != nullptr, &dst);
    if (rc != FSX_OK) return rc;
    if (dst->validity) FSX_ASYNC(ctx, cudaMemsetAsync(dst->validity, 0, (size_t)((n + 127) / 128) * 16, ctx->stream));
    if (n > 0) {
        const int grid = fsx_grid_for(ctx, n, 256 * 4, 8);
        if (fsx_dtype_size(src->dtype) == 4)
            k_gather_u32<int32_t><<<grid, 256, 0, ctx->stream>>>((const int32_t *)src->data, src->validity, perm, n, (int32_t *)dst->data, (unsigned *)dst->validity);
        else
            k_gather_u32<long long><<<grid, 256, 0, ctx->stream>>>((const long long *)src->data, src->validity, perm, n, (long long *)dst->data, (unsigned *)dst->validity);
        ctx->launches++;
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { fsx_col_release(ctx, dst); return fsx_fail(ctx, FSX_E_CUDA, "gather: %s", cudaGetErrorString(e)); }
    *out = dst;
    return FSX_OK;
}

// ComputeFeatures.scala writes every feature group as it comes out of the typed map (row order unspecified); the tour
// job sorts its output by team_id so that files are reproducible.  Ascending, NaN above +inf, nulls last, stable.
extern "C" int32_t fsx_table_sort_by(fsx_ctx *ctx, fsx_table *t, int32_t by) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, t && by >= 0 && by < (int32_t)t->cols.size(), FSX_E_INVALID, "fsx_table_sort_by: bad table/column");
    const int64_t n = t->nrows;
    if (n <= 1) { FSX_END(); }
    const fsx_col *kc = t->cols[(size_t)by];
    unsigned *perm = nullptr;
    FSX_CHECK(fsx_internal_argsort(ctx, &kc, 1, n, &perm));
    for (size_t c = 0; c < t->cols.size(); c++) {
        fsx_col *dst = nullptr;
        int32_t rc = fsx_internal_gather_u32(ctx, t->cols[c], perm, n, &dst);
        if (rc != FSX_OK) { fsx_dev_free(ctx, perm); return rc; }
        fsx_col_release(ctx, t->cols[c]);
        t->cols[c] = dst;
    }
    fsx_dev_free(ctx, perm);
    FSX_END();
}
