// partition.cu -- K6 hash partitioner for the multi-rank exchange, and the table sort helper.
//
// K6 replaces Spark's `Exchange hashpartitioning(team_id, 200)` that sits between the partial
// and the final HashAggregate and in front of the join (implicit in ComputeFeatures.scala:
// 148-150).  Rows are routed to rank = mix(key) % nparts.  Two passes over
// the key column: (1) per-destination counts (shared-memory histogram per CTA, one global
// atomic per CTA and destination); (2) scatter: every CTA ranks its tile's rows per destination
// in shared memory, reserves one contiguous run per destination with a single global atomic,
// and writes keys + payload columns into the destination-grouped buffers.  The result is laid
// out exactly as an all-to-all wants it: nparts contiguous segments whose sizes are returned
// to the host.  Algorithmic bytes: 2x key column + 2x payload (read + write).
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

__device__ __forceinline__ int part_of(const PartParams &p, int64_t i) {
    if (!dev_bit_valid(p.col[0].src_valid, i)) return 0;      // null keys travel to rank 0
    unsigned long long h = p.key_is64 ? mix64((unsigned long long)((const long long *)p.col[0].src)[i])
                                      : (unsigned long long)mix32((uint32_t)((const int32_t *)p.col[0].src)[i]);
    return (int)(h % (unsigned long long)p.nparts);
}

__global__ void __launch_bounds__(PART_THREADS) k_part_count(const __grid_constant__ PartParams p,
                                                             unsigned long long *__restrict__ counts) {
    __shared__ unsigned hist[PART_MAX];
    for (int i = threadIdx.x; i < p.nparts; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < p.n; i += (int64_t)gridDim.x * blockDim.x)
        atomicAdd(&hist[part_of(p, i)], 1u);
    __syncthreads();
    for (int i = threadIdx.x; i < p.nparts; i += blockDim.x)
        if (hist[i]) atomicAdd(&counts[i], (unsigned long long)hist[i]);
}

__global__ void __launch_bounds__(PART_THREADS) k_part_scatter(const __grid_constant__ PartParams p,
                                                               unsigned long long *__restrict__ cursors) {
    __shared__ unsigned hist[PART_MAX];
    __shared__ unsigned long long base[PART_MAX];
    for (int64_t tile = blockIdx.x; tile * PART_TILE < p.n; tile += gridDim.x) {
        for (int i = threadIdx.x; i < p.nparts; i += blockDim.x) hist[i] = 0;
        __syncthreads();
        int dest[PART_ITEMS];
        unsigned rank[PART_ITEMS];
#pragma unroll
        for (int j = 0; j < PART_ITEMS; j++) {
            const int64_t i = tile * PART_TILE + (int64_t)j * PART_THREADS + threadIdx.x;
            dest[j] = -1;
            if (i < p.n) { dest[j] = part_of(p, i); rank[j] = atomicAdd(&hist[dest[j]], 1u); }
        }
        __syncthreads();
        for (int i = threadIdx.x; i < p.nparts; i += blockDim.x)
            if (hist[i]) base[i] = atomicAdd(&cursors[i], (unsigned long long)hist[i]);
        __syncthreads();
#pragma unroll
        for (int j = 0; j < PART_ITEMS; j++) {
            if (dest[j] < 0) continue;
            const int64_t i = tile * PART_TILE + (int64_t)j * PART_THREADS + threadIdx.x;
            const unsigned long long o = base[dest[j]] + rank[j];
            for (int c = 0; c < p.ncols; c++) {
                const PartCol &pc = p.col[c];
                if (pc.elem == 4) ((int32_t *)pc.dst)[o] = ((const int32_t *)pc.src)[i];
                else ((long long *)pc.dst)[o] = ((const long long *)pc.src)[i];
                if (pc.dst_valid && dev_bit_valid(pc.src_valid, i)) atomicOr(&pc.dst_valid[o >> 5], 1u << (o & 31));
            }
        }
        __syncthreads();
    }
}

template <typename T>
__global__ void __launch_bounds__(256) k_permute(const T *__restrict__ src, const uint8_t *__restrict__ src_valid,
                                                 const int64_t *__restrict__ perm, int64_t n, T *__restrict__ dst,
                                                 unsigned *__restrict__ dst_valid) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t j = perm[i];
        dst[i] = src[j];
        if (dst_valid && dev_bit_valid(src_valid, j)) atomicOr(&dst_valid[i >> 5], 1u << (i & 31));
    }
}

} // namespace

extern "C" int32_t fsx_partition_by_key(fsx_ctx *ctx, const fsx_col *key, const fsx_col *const *payload,
                                        int32_t npayload, int32_t nparts, int64_t *counts_out,
                                        fsx_table **out) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, key && counts_out && out, FSX_E_INVALID, "fsx_partition_by_key: NULL argument");
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
        if (dst->validity) cudaMemsetAsync(dst->validity, 0, (size_t)((src->n + 127) / 128) * 16, ctx->stream);
        p.col[c].src = src->data; p.col[c].src_valid = src->validity;
        p.col[c].dst = dst->data; p.col[c].dst_valid = (unsigned *)dst->validity;
        p.col[c].elem = (int)fsx_dtype_size(src->dtype);
    }
    unsigned long long *d_counts = nullptr;
    int32_t rc = fsx_dev_alloc(ctx, (size_t)nparts * 8, (void **)&d_counts);
    if (rc != FSX_OK) return fail(rc);
    cudaMemsetAsync(d_counts, 0, (size_t)nparts * 8, ctx->stream);
    std::vector<unsigned long long> h((size_t)nparts, 0ULL);
    cudaError_t e = cudaSuccess;
    if (key->n > 0) {
        k_part_count<<<fsx_grid_for(ctx, key->n, PART_TILE, 8), PART_THREADS, 0, ctx->stream>>>(p, d_counts);
        ctx->launches++;
        e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaMemcpyAsync(h.data(), d_counts, (size_t)nparts * 8, cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        if (e == cudaSuccess) {
            std::vector<unsigned long long> cur((size_t)nparts);
            unsigned long long acc = 0;
            for (int i = 0; i < nparts; i++) { cur[(size_t)i] = acc; acc += h[(size_t)i]; }
            e = cudaMemcpyAsync(d_counts, cur.data(), (size_t)nparts * 8, cudaMemcpyHostToDevice, ctx->stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        }
        if (e == cudaSuccess) {
            k_part_scatter<<<fsx_grid_for(ctx, key->n, PART_TILE, 4), PART_THREADS, 0, ctx->stream>>>(p, d_counts);
            ctx->launches++;
            e = cudaGetLastError();
        }
    }
    fsx_dev_free(ctx, d_counts);
    if (e != cudaSuccess) { fsx_table_free(ctx, t); FSX_CUDA(ctx, e); }
    for (int i = 0; i < nparts; i++) counts_out[i] = (int64_t)h[(size_t)i];
    *out = t;
    FSX_END();
}


