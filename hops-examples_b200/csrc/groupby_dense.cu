// groupby_dense.cu -- K1 on caller-owned dense state arrays, for the multi-rank merge over NVSwitch.
//
// Same reference step as groupby.cu (groupBy("team_id").sum()/.count() and the aggregates of
// BASELINE.json config 2; ComputeFeatures.scala:148-150), split so that the exchange between ranks
// (Spark: Exchange hashpartitioning + final HashAggregate) becomes three NCCL reduce-scatters instead of
// partition -> all-to-all -> merge:
//   when the key domain is dense ([kmin, kmin + nslots) agreed by all ranks) every rank accumulates its
//   row shard into the SAME slot layout -- sum[nslots] f64, count[nslots] i64, mm[2*nslots] i64 -- held
//   in buffers the caller owns (torch tensors).  Summing sum/count and taking the minimum of mm across
//   ranks merges the partial states; with reduce-scatter every rank ends up owning one contiguous slice
//   of the key domain, already merged (in-switch reduction when NVLS is available), and emits its groups.
// mm holds {min, max} interleaved in an encoding that a plain signed MIN merges:
//   mm[2s]   = order-preserving signed image of the minimum
//   mm[2s+1] = order-preserving signed image of the complement of the maximum  (min of complements = max)
// Identity elements: sum 0, count 0, mm INT64_MAX (fsx_groupby_dense_init).
// The accumulate kernel is the measured-best form of groupby.cu: RED f64 + RED u64 per row, one 16-byte
// load of {min,max} and conditional atomics, four rows per thread in flight, 16 x 148 CTAs.
#include <type_traits>
#include <vector>

#include "scan.cuh"
#include "fxacc.cuh"

namespace {

constexpr unsigned long long MSB = 0x8000000000000000ULL;

template <typename V>
__device__ __forceinline__ unsigned long long ord_of(V v);
template <> __device__ __forceinline__ unsigned long long ord_of<double>(double v) { return ord_from_f64(v); }
template <> __device__ __forceinline__ unsigned long long ord_of<float>(float v) { return ord_from_f64((double)v); }
template <> __device__ __forceinline__ unsigned long long ord_of<int32_t>(int32_t v) { return ord_from_i64((long long)v); }
template <> __device__ __forceinline__ unsigned long long ord_of<long long>(long long v) { return ord_from_i64(v); }

template <typename KeyT, typename ValT>
__global__ void __launch_bounds__(256) k_dense_accumulate(const KeyT *__restrict__ keys, const ValT *__restrict__ vals, int64_t n,
                                                          long long kmin, unsigned long long nslots, double *sum,
                                                          unsigned long long *cnt, long long *mm, unsigned long long *bad,
                                                          unsigned long long pass_lo, unsigned long long pass_w) {
    constexpr int U = 4;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    auto one = [&](long long k, ValT v) {
        const unsigned long long slot = (unsigned long long)(k - kmin);
        if (slot >= nslots) { *bad = 1ULL; return; }              // key outside the agreed domain
        if (slot - pass_lo >= pass_w) return;                     // state larger than L2: another launch owns this slot range
        atomicAdd(&cnt[slot], 1ULL);
        atomicAdd(&sum[slot], (double)v);
        const unsigned long long o = ord_of<ValT>(v);
        const long long lo = (long long)(o ^ MSB), hi = (long long)((~o) ^ MSB);
        long long *p = mm + 2 * slot;
        const longlong2 cur = __ldcg((const longlong2 *)p);      // a stale read only causes a redundant atomic
        if (lo < cur.x) atomicMin(p, lo);
        if (hi < cur.y) atomicMin(p + 1, hi);
    };
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i + (U - 1) * stride < n; i += U * stride) {
        long long k[U];
        ValT v[U];
#pragma unroll
        for (int u = 0; u < U; u++) { k[u] = keys[i + u * stride]; v[u] = vals[i + u * stride]; }
#pragma unroll
        for (int u = 0; u < U; u++) one(k[u], v[u]);
    }
    for (; i < n; i += stride) one(keys[i], vals[i]);
}


// Low-cardinality form (the slot range fits in shared memory): with few slots the global REDs above pile up on a
// handful of L2 addresses (100 M rows into 50 slots: 29 ms), so every CTA aggregates into a private table in shared
// memory -- 32-bit counts, sums as exact wide fixed-point accumulators of 32-bit limbs (csrc/fxacc.cuh: 64-bit shared
// atomics are CAS loops), min/max read first and CAS'd on improvement -- and adds its table to the global state
// once.  A value outside the fixed-point window (NaN, Inf, an outlier) goes to the global sum directly.
template <typename KeyT, typename ValT>
__global__ void __launch_bounds__(1024) k_dense_accumulate_smem(const KeyT *__restrict__ keys, const ValT *__restrict__ vals, int64_t n,
                                                                long long kmin, int S, int e_base, double *sum, unsigned long long *cnt,
                                                                long long *mm, unsigned long long *bad) {
    extern __shared__ __align__(16) unsigned d_smem[];
    constexpr bool is_float = std::is_floating_point<ValT>::value;
    constexpr int L = is_float ? fx::L_FLOAT : fx::L_INT;
    long long *s_mm = (long long *)d_smem;                         // 2 x S, 16-byte aligned pairs
    unsigned *s_cnt = d_smem + 4 * S, *limbs = s_cnt + S;
    for (int i = threadIdx.x; i < 2 * S; i += blockDim.x) s_mm[i] = INT64_MAX;
    for (int i = threadIdx.x; i < (L + 1) * S; i += blockDim.x) s_cnt[i] = 0;
    __syncthreads();
    auto one = [&](long long k, ValT v) {
        const unsigned long long slot64 = (unsigned long long)(k - kmin);
        if (slot64 >= (unsigned long long)S) { *bad = 1ULL; return; }
        const int slot = (int)slot64;
        atomicAdd(&s_cnt[slot], 1u);
        fx::Chunk ch;
        bool fits = true;
        if constexpr (is_float) fits = fx::place_f64((double)v, e_base, ch);
        else fx::place_i64((long long)v, ch);
        if (fits) fx::add(limbs + slot, S, L, ch);
        else atomicAdd(&sum[slot], (double)v);
        const unsigned long long o = ord_of<ValT>(v);
        const long long lo_o = (long long)(o ^ MSB), hi_o = (long long)((~o) ^ MSB);
        long long *p = s_mm + 2 * slot;
        const longlong2 cur = *(const longlong2 *)p;
        if (lo_o < cur.x) atomicMin(p, lo_o);
        if (hi_o < cur.y) atomicMin(p + 1, hi_o);
    };
    constexpr int U = 4;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i + (U - 1) * stride < n; i += U * stride) {
        long long k[U];
        ValT v[U];
#pragma unroll
        for (int u = 0; u < U; u++) { k[u] = keys[i + u * stride]; v[u] = vals[i + u * stride]; }
#pragma unroll
        for (int u = 0; u < U; u++) one(k[u], v[u]);
    }
    for (; i < n; i += stride) one(keys[i], vals[i]);
    __syncthreads();
    for (int s = threadIdx.x; s < S; s += blockDim.x) {
        const unsigned c = s_cnt[s];
        if (!c) continue;
        atomicAdd(&cnt[s], (unsigned long long)c);
        const double d = fx::to_double(limbs + s, S, L, is_float ? e_base - 1075 : 0);   // this CTA's exact partial, rounded once
        if (d != 0.0) atomicAdd(&sum[s], d);
        atomicMin(&mm[2 * s], s_mm[2 * s]);
        atomicMin(&mm[2 * s + 1], s_mm[2 * s + 1]);
    }
}

// largest finite |v| of a strided sample (non-negative doubles order like their bit patterns)
template <typename VT>
__global__ void k_dense_absmax_sample(const VT *__restrict__ v, int64_t n, int64_t step, unsigned long long *out) {
    unsigned long long m = 0;
    for (int64_t i = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) * step; i < n; i += (int64_t)gridDim.x * blockDim.x * step) {
        const double a = fabs((double)v[i]);
        if (a < 1.7976931348623157e308) { const unsigned long long b = (unsigned long long)__double_as_longlong(a); m = b > m ? b : m; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { const unsigned long long x = __shfl_xor_sync(0xffffffffu, m, o); m = x > m ? x : m; }
    if ((threadIdx.x & 31) == 0 && m) atomicMax(out, m);
}

constexpr int64_t DENSE_SMEM_MAX_SLOTS = 5000;       // 40 B per slot for float values (counts, 5 limbs, {min,max})

// bottom of the fixed-point window for a float column (csrc/fxacc.cuh), from a sample of |v|; unused for integer columns
template <typename ValT>
int32_t dense_pick_scale(fsx_ctx *ctx, const ValT *vals, int64_t n, int *e_base) {
    *e_base = 0;
    if constexpr (std::is_floating_point<ValT>::value) {
        cudaError_t e = cudaMemsetAsync(ctx->d_scalars + 10, 0, 8, ctx->stream);
        if (e != cudaSuccess) return fsx_fail(ctx, FSX_E_CUDA, "cudaMemsetAsync failed: %s", cudaGetErrorString(e));
        const int64_t step = n / 65536 > 0 ? n / 65536 : 1;
        k_dense_absmax_sample<ValT><<<64, 256, 0, ctx->stream>>>(vals, n, step, ctx->d_scalars + 10);
        ctx->launches++;
        int32_t rc = fsx_read_scalars(ctx, 10, 1);
        if (rc != FSX_OK) return rc;
        double smax;
        memcpy(&smax, &ctx->h_scalars[10], 8);
        *e_base = fx_pick_e_base(smax);
    }
    return FSX_OK;
}

// one accumulate launch over rows [0, n): shared-memory form when the slot range is small, global REDs otherwise
template <typename KeyT, typename ValT>
cudaError_t dense_launch(fsx_ctx *ctx, const KeyT *keys, const ValT *vals, int64_t n, long long kmin, int64_t nslots, int e_base,
                         double *sum, unsigned long long *cnt, long long *mm, unsigned long long *bad) {
    if (nslots <= DENSE_SMEM_MAX_SLOTS && n < (1LL << 31)) {
        const size_t bytes = (size_t)nslots * 4 * (5 + (std::is_floating_point<ValT>::value ? fx::L_FLOAT : fx::L_INT));
        int per_sm = (int)((216 * 1024) / (bytes + 1024)), threads = 512;
        if (per_sm >= 4) per_sm = 4; else if (per_sm >= 2) per_sm = 2; else { per_sm = 1; threads = 1024; }
        int grid = fsx_grid_for(ctx, n, threads * 8, per_sm);
        while (grid > 1 && n / grid < 4 * nslots) grid /= 2;
        cudaError_t e = cudaFuncSetAttribute(k_dense_accumulate_smem<KeyT, ValT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
        if (e != cudaSuccess) return e;
        k_dense_accumulate_smem<KeyT, ValT><<<grid, threads, bytes, ctx->stream>>>(keys, vals, n, kmin, (int)nslots, e_base, sum, cnt, mm, bad);
    } else {
        // 32 B of state per slot; beyond ~96 MB the rows are re-streamed once per 64 MB slot range (csrc/groupby.cu does the same)
        const unsigned long long bytes = 32ULL * (unsigned long long)nslots;
        int passes = bytes > (96ULL << 20) ? (int)((bytes + (64ULL << 20) - 1) / (64ULL << 20)) : 1;
        if (passes > 64) passes = 64;
        const unsigned long long width = ((unsigned long long)nslots + passes - 1) / passes;
        for (int ps = 0; ps < passes; ps++) {
            k_dense_accumulate<KeyT, ValT><<<fsx_grid_for(ctx, n, 256 * 4, 16), 256, 0, ctx->stream>>>(
                keys, vals, n, kmin, (unsigned long long)nslots, sum, cnt, mm, bad, (unsigned long long)ps * width, passes == 1 ? ~0ULL : width);
            if (ps) ctx->launches++;
        }
    }
    ctx->launches++;
    return cudaGetLastError();
}

__global__ void k_dense_init(double *sum, unsigned long long *cnt, long long *mm, unsigned long long nslots) {
    for (unsigned long long s = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; s < nslots;
         s += (unsigned long long)gridDim.x * blockDim.x) {
        sum[s] = 0.0; cnt[s] = 0ULL; mm[2 * s] = INT64_MAX; mm[2 * s + 1] = INT64_MAX;
    }
}

__global__ void k_dense_flags(const unsigned long long *__restrict__ cnt, unsigned long long nslots, unsigned *flags) {
    for (unsigned long long s = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; s < nslots;
         s += (unsigned long long)gridDim.x * blockDim.x)
        flags[s] = cnt[s] != 0;
}

template <typename KeyT, typename ValT>
__global__ void k_dense_fill(const double *__restrict__ sum, const unsigned long long *__restrict__ cnt,
                             const long long *__restrict__ mm, const unsigned long long *__restrict__ offsets,
                             unsigned long long nslots, long long kmin, KeyT *key_out, double *sum_out, double *avg_out,
                             long long *cnt_out, ValT *min_out, ValT *max_out) {
    for (unsigned long long s = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; s < nslots;
         s += (unsigned long long)gridDim.x * blockDim.x) {
        const unsigned long long c = cnt[s];
        if (!c) continue;
        const unsigned long long o = offsets[s];
        key_out[o] = (KeyT)(kmin + (long long)s);
        sum_out[o] = sum[s];
        avg_out[o] = sum[s] / (double)(long long)c;
        cnt_out[o] = (long long)c;
        const unsigned long long omin = (unsigned long long)mm[2 * s] ^ MSB, omax = ~((unsigned long long)mm[2 * s + 1] ^ MSB);
        if constexpr (std::is_floating_point<ValT>::value) { min_out[o] = (ValT)ord_to_f64(omin); max_out[o] = (ValT)ord_to_f64(omax); }
        else { min_out[o] = (ValT)ord_to_i64(omin); max_out[o] = (ValT)ord_to_i64(omax); }
    }
}

template <typename T>
__global__ void __launch_bounds__(256) k_col_minmax(const T *__restrict__ v, const uint8_t *__restrict__ valid, int64_t n,
                                                    unsigned long long *scalars) {
    unsigned long long lo = ~0ULL, hi = 0ULL, cntv = 0;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (!dev_bit_valid(valid, i)) continue;
        const unsigned long long o = ord_from_i64((long long)v[i]);
        lo = o < lo ? o : lo;
        hi = o > hi ? o : hi;
        cntv++;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long l2 = __shfl_xor_sync(0xffffffffu, lo, o), h2 = __shfl_xor_sync(0xffffffffu, hi, o);
        cntv += __shfl_xor_sync(0xffffffffu, cntv, o);
        lo = l2 < lo ? l2 : lo;
        hi = h2 > hi ? h2 : hi;
    }
    if ((threadIdx.x & 31) == 0) { atomicMin(&scalars[2], lo); atomicMax(&scalars[3], hi); if (cntv) atomicAdd(&scalars[4], cntv); }
}

} // namespace

extern "C" int32_t fsx_column_minmax(fsx_ctx *ctx, const fsx_col *col, int64_t *min_out, int64_t *max_out, int64_t *nvalid_out) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, col && min_out && max_out, FSX_E_INVALID, "fsx_column_minmax: NULL argument");
    FSX_REQUIRE(ctx, col->dtype == FSX_I32 || col->dtype == FSX_I64, FSX_E_UNSUPPORTED, "fsx_column_minmax: integer columns only");
    unsigned long long init[8] = {0, 0, ~0ULL, 0, 0, 0, 0, 0};
    FSX_CUDA(ctx, cudaMemcpyAsync(ctx->d_scalars, init, sizeof init, cudaMemcpyHostToDevice, ctx->stream));
    if (col->n > 0) {
        const int grid = fsx_grid_for(ctx, col->n, 256 * 8, 8);
        if (col->dtype == FSX_I32) FSX_LAUNCH(ctx, k_col_minmax<int32_t>, grid, 256, 0, (const int32_t *)col->data, col->validity, col->n, ctx->d_scalars);
        else FSX_LAUNCH(ctx, k_col_minmax<long long>, grid, 256, 0, (const long long *)col->data, col->validity, col->n, ctx->d_scalars);
    }
    FSX_CHECK(fsx_read_scalars(ctx, 0, 8));
    const bool any = ctx->h_scalars[4] != 0;
    *min_out = any ? (int64_t)(ctx->h_scalars[2] ^ MSB) : 0;
    *max_out = any ? (int64_t)(ctx->h_scalars[3] ^ MSB) : -1;
    if (nvalid_out) *nvalid_out = (int64_t)ctx->h_scalars[4];
    FSX_END();
}

extern "C" int32_t fsx_groupby_dense_init(fsx_ctx *ctx, int64_t nslots, double *sum_dev, int64_t *cnt_dev, int64_t *mm_dev) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, nslots >= 0 && (nslots == 0 || (sum_dev && cnt_dev && mm_dev)), FSX_E_INVALID, "fsx_groupby_dense_init: bad argument");
    if (nslots > 0)
        FSX_LAUNCH(ctx, k_dense_init, fsx_grid_for(ctx, nslots, 256, 8), 256, 0, sum_dev, (unsigned long long *)cnt_dev,
                   (long long *)mm_dev, (unsigned long long)nslots);
    FSX_END();
}

extern "C" int32_t fsx_groupby_dense_accumulate(fsx_ctx *ctx, const fsx_col *key, const fsx_col *val, int64_t kmin, int64_t nslots,
                                                double *sum_dev, int64_t *cnt_dev, int64_t *mm_dev) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, key && val && sum_dev && cnt_dev && mm_dev && nslots > 0, FSX_E_INVALID, "fsx_groupby_dense_accumulate: bad argument");
    FSX_REQUIRE(ctx, key->dtype == FSX_I32 || key->dtype == FSX_I64, FSX_E_UNSUPPORTED, "fsx_groupby_dense_accumulate: key must be i32/i64");
    FSX_REQUIRE(ctx, !key->validity && !val->validity, FSX_E_UNSUPPORTED,
                "fsx_groupby_dense_accumulate: null-free columns only (use fsx_groupby_agg + fsx_groupby_merge otherwise)");
    FSX_REQUIRE(ctx, key->n == val->n, FSX_E_INVALID, "fsx_groupby_dense_accumulate: key and value lengths differ");
    const int64_t n = key->n;
    if (n == 0) { FSX_END(); }
    FSX_CUDA(ctx, cudaMemsetAsync(ctx->d_scalars + 9, 0, 8, ctx->stream));
    unsigned long long *cnt = (unsigned long long *)cnt_dev;
    long long *mm = (long long *)mm_dev;
    unsigned long long *bad = ctx->d_scalars + 9;
    const bool small = nslots <= DENSE_SMEM_MAX_SLOTS;
    int e_base = 0;
    if (small && val->dtype == FSX_F32) FSX_CHECK(dense_pick_scale<float>(ctx, (const float *)val->data, n, &e_base));
    if (small && val->dtype == FSX_F64) FSX_CHECK(dense_pick_scale<double>(ctx, (const double *)val->data, n, &e_base));
    FSX_CUDA(ctx, cudaEventRecord(ctx->ev_c, ctx->stream));
    cudaError_t e;
#define DENSE(K, V) e = dense_launch<K, V>(ctx, (const K *)key->data, (const V *)val->data, n, kmin, nslots, e_base, sum_dev, cnt, mm, bad)
    const bool k32 = key->dtype == FSX_I32;
    switch (val->dtype) {
    case FSX_I32: if (k32) DENSE(int32_t, int32_t); else DENSE(long long, int32_t); break;
    case FSX_I64: if (k32) DENSE(int32_t, long long); else DENSE(long long, long long); break;
    case FSX_F32: if (k32) DENSE(int32_t, float); else DENSE(long long, float); break;
    default: if (k32) DENSE(int32_t, double); else DENSE(long long, double); break;
    }
#undef DENSE
    FSX_CUDA(ctx, e);
    FSX_CUDA(ctx, cudaEventRecord(ctx->ev_d, ctx->stream));
    FSX_CHECK(fsx_read_scalars(ctx, 9, 1));
    float ms = 0.f;
    FSX_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev_c, ctx->ev_d));
    ctx->timing.main_kernel_ms = ms;
    FSX_REQUIRE(ctx, ctx->h_scalars[9] == 0, FSX_E_DATA, "fsx_groupby_dense_accumulate: a key lies outside [%lld, %lld)",
                (long long)kmin, (long long)(kmin + nslots));
    FSX_END();
}

// defer: size the output columns for the upper bound (nslots rows) instead of reading the group count back in the
// middle of the call; the count then sits in ctx->d_scalars[0] and the caller sets the row count after its final
// synchronisation (fsx_internal_table_set_rows).  Used when nslots is small enough for the over-allocation not to matter.
static int32_t dense_emit_impl(fsx_ctx *ctx, int32_t key_dtype, int32_t val_dtype, int64_t kmin, int64_t nslots,
                               const double *sum_dev, const int64_t *cnt_dev, const int64_t *mm_dev, fsx_table **out, bool defer = false) {
    unsigned *flags = nullptr;
    unsigned long long *offsets = nullptr;
    int64_t G = 0;
    if (nslots > 0) {
        FSX_CHECK(fsx_dev_alloc(ctx, (size_t)nslots * 4, (void **)&flags));
        int32_t rc = fsx_dev_alloc(ctx, (size_t)nslots * 8, (void **)&offsets);
        if (rc != FSX_OK) { fsx_dev_free(ctx, flags); return rc; }
        k_dense_flags<<<fsx_grid_for(ctx, nslots, 256, 8), 256, 0, ctx->stream>>>((const unsigned long long *)cnt_dev, (unsigned long long)nslots, flags);
        ctx->launches++;
        rc = exclusive_scan<unsigned>(ctx, flags, (unsigned long long)nslots, offsets, 0);
        if (rc == FSX_OK && !defer) rc = fsx_read_scalars(ctx, 0, 1);
        if (rc != FSX_OK) { fsx_dev_free(ctx, flags); fsx_dev_free(ctx, offsets); return rc; }
        G = defer ? nslots : (int64_t)ctx->h_scalars[0];
    } else if (defer) {
        FSX_CUDA(ctx, cudaMemsetAsync(ctx->d_scalars, 0, 8, ctx->stream));
    }
    fsx_table *t = new fsx_table();
    t->nrows = G;
    const int32_t dts[6] = {key_dtype, FSX_F64, FSX_F64, FSX_I64, val_dtype, val_dtype};   // key, sum, avg, count, min, max
    for (int i = 0; i < 6; i++) {
        fsx_col *c = nullptr;
        int32_t rc = fsx_col_new(ctx, dts[i], G, false, &c);
        if (rc != FSX_OK) { fsx_table_free(ctx, t); fsx_dev_free(ctx, flags); fsx_dev_free(ctx, offsets); return rc; }
        t->cols.push_back(c);
    }
    if (G > 0) {
        const int grid = fsx_grid_for(ctx, nslots, 256, 8);
        const unsigned long long *cnt = (const unsigned long long *)cnt_dev;
        const long long *mm = (const long long *)mm_dev;
#define FILL(K, V) k_dense_fill<K, V><<<grid, 256, 0, ctx->stream>>>(sum_dev, cnt, mm, offsets, (unsigned long long)nslots, kmin, (K *)t->cols[0]->data, \
            (double *)t->cols[1]->data, (double *)t->cols[2]->data, (long long *)t->cols[3]->data, (V *)t->cols[4]->data, (V *)t->cols[5]->data)
        const bool k32 = key_dtype == FSX_I32;
        switch (val_dtype) {
        case FSX_I32: if (k32) FILL(int32_t, int32_t); else FILL(long long, int32_t); break;
        case FSX_I64: if (k32) FILL(int32_t, long long); else FILL(long long, long long); break;
        case FSX_F32: if (k32) FILL(int32_t, float); else FILL(long long, float); break;
        default: if (k32) FILL(int32_t, double); else FILL(long long, double); break;
        }
#undef FILL
        ctx->launches++;
    }
    cudaError_t e = cudaGetLastError();
    fsx_dev_free(ctx, flags);
    fsx_dev_free(ctx, offsets);
    if (e != cudaSuccess) { fsx_table_free(ctx, t); FSX_CUDA(ctx, e); }
    *out = t;
    return FSX_OK;
}

// ---- the same steps for callers that are already inside an ABI call and keep the stream asynchronous (csrc/dist.cu) ----
int32_t fsx_internal_dense_reset(fsx_ctx *ctx, int64_t nslots, double *sum, unsigned long long *cnt, long long *mm) {
    if (nslots > 0) FSX_LAUNCH(ctx, k_dense_init, fsx_grid_for(ctx, nslots, 256, 8), 256, 0, sum, cnt, mm, (unsigned long long)nslots);
    return FSX_OK;
}

// accumulate key/val into the state; *bad (device u64, not cleared here) is set when a key lies outside the domain.
// No read-back except the fixed-point scale sample of the small-domain route.  ev_c/ev_d bracket the kernel.
int32_t fsx_internal_dense_accumulate(fsx_ctx *ctx, const fsx_col *key, const fsx_col *val, int64_t kmin, int64_t nslots,
                                      double *sum, unsigned long long *cnt, long long *mm, unsigned long long *bad) {
    const int64_t n = key->n;
    if (n == 0) { FSX_CUDA(ctx, cudaEventRecord(ctx->ev_c, ctx->stream)); FSX_CUDA(ctx, cudaEventRecord(ctx->ev_d, ctx->stream)); return FSX_OK; }
    const bool small = nslots <= DENSE_SMEM_MAX_SLOTS;
    int e_base = 0;
    if (small && val->dtype == FSX_F32) FSX_CHECK(dense_pick_scale<float>(ctx, (const float *)val->data, n, &e_base));
    if (small && val->dtype == FSX_F64) FSX_CHECK(dense_pick_scale<double>(ctx, (const double *)val->data, n, &e_base));
    FSX_CUDA(ctx, cudaEventRecord(ctx->ev_c, ctx->stream));
    cudaError_t e;
#define DENSE(K, V) e = dense_launch<K, V>(ctx, (const K *)key->data, (const V *)val->data, n, kmin, nslots, e_base, sum, cnt, mm, bad)
    const bool k32 = key->dtype == FSX_I32;
    switch (val->dtype) {
    case FSX_I32: if (k32) DENSE(int32_t, int32_t); else DENSE(long long, int32_t); break;
    case FSX_I64: if (k32) DENSE(int32_t, long long); else DENSE(long long, long long); break;
    case FSX_F32: if (k32) DENSE(int32_t, float); else DENSE(long long, float); break;
    default: if (k32) DENSE(int32_t, double); else DENSE(long long, double); break;
    }
#undef DENSE
    FSX_CUDA(ctx, e);
    FSX_CUDA(ctx, cudaEventRecord(ctx->ev_d, ctx->stream));
    return FSX_OK;
}

int32_t fsx_internal_dense_emit(fsx_ctx *ctx, int32_t key_dtype, int32_t val_dtype, int64_t kmin, int64_t nslots, const double *sum,
                                const int64_t *cnt, const int64_t *mm, bool defer, fsx_table **out) {
    return dense_emit_impl(ctx, key_dtype, val_dtype, kmin, nslots, sum, cnt, mm, out, defer);
}

// a table whose columns were sized for an upper bound: the real row count, known after the final synchronisation
void fsx_internal_table_set_rows(fsx_table *t, int64_t rows) {
    t->nrows = rows;
    for (fsx_col *c : t->cols) c->n = rows;
}

extern "C" int32_t fsx_groupby_dense_emit(fsx_ctx *ctx, int32_t key_dtype, int32_t val_dtype, int64_t kmin, int64_t nslots,
                                          const double *sum_dev, const int64_t *cnt_dev, const int64_t *mm_dev, fsx_table **out) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, out && nslots >= 0, FSX_E_INVALID, "fsx_groupby_dense_emit: bad argument");
    *out = nullptr;
    FSX_REQUIRE(ctx, (key_dtype == FSX_I32 || key_dtype == FSX_I64) && fsx_dtype_ok(val_dtype), FSX_E_INVALID, "fsx_groupby_dense_emit: bad dtype");
    FSX_CHECK(dense_emit_impl(ctx, key_dtype, val_dtype, kmin, nslots, sum_dev, cnt_dev, mm_dev, out));
    FSX_END();
}

// ------------------------------------------------------------------------------------------------------------------
// One call for a batch that lives in HOST memory: chunked H2D copies on the copy stream overlap the accumulation of
// the chunks already on the device, so the call costs the PCIe transfer plus one chunk's kernel, not their sum.
// ------------------------------------------------------------------------------------------------------------------
namespace {

struct DenseState {
    double *sum = nullptr;
    unsigned long long *cnt = nullptr;
    long long *mm = nullptr;
    int64_t lo = 0, nslots = 0;
};

void dense_state_free(fsx_ctx *ctx, DenseState &st) {
    fsx_dev_free(ctx, st.sum); fsx_dev_free(ctx, st.cnt); fsx_dev_free(ctx, st.mm);
    st = DenseState();
}

int32_t dense_state_alloc(fsx_ctx *ctx, DenseState &st, int64_t lo, int64_t nslots) {
    st.lo = lo; st.nslots = nslots;
    int32_t rc = fsx_dev_alloc(ctx, (size_t)nslots * 8, (void **)&st.sum);
    if (rc == FSX_OK) rc = fsx_dev_alloc(ctx, (size_t)nslots * 8, (void **)&st.cnt);
    if (rc == FSX_OK) rc = fsx_dev_alloc(ctx, (size_t)nslots * 16, (void **)&st.mm);
    if (rc != FSX_OK) { dense_state_free(ctx, st); return rc; }
    k_dense_init<<<fsx_grid_for(ctx, nslots, 256, 8), 256, 0, ctx->stream>>>(st.sum, st.cnt, st.mm, (unsigned long long)nslots);
    ctx->launches++;
    return FSX_OK;
}

// make the state cover [lo, hi]; existing slots keep their contents
int32_t dense_state_cover(fsx_ctx *ctx, DenseState &st, int64_t lo, int64_t hi, int64_t max_slots) {
    if (st.nslots > 0 && lo >= st.lo && hi < st.lo + st.nslots) return FSX_OK;
    int64_t nlo = lo, nhi = hi;
    if (st.nslots > 0) { nlo = lo < st.lo ? lo : st.lo; nhi = hi > st.lo + st.nslots - 1 ? hi : st.lo + st.nslots - 1; }
    if ((unsigned long long)nhi - (unsigned long long)nlo >= (unsigned long long)max_slots) return FSX_E_UNSUPPORTED;
    // slack on both sides: later chunks of the same distribution rarely force another move
    const int64_t slack = (nhi - nlo + 1) / 16 + 64;
    nlo = nlo < INT64_MIN + slack ? INT64_MIN : nlo - slack;
    nhi = nhi > INT64_MAX - slack ? INT64_MAX : nhi + slack;
    if ((unsigned long long)nhi - (unsigned long long)nlo >= (unsigned long long)INT64_MAX) return FSX_E_UNSUPPORTED;
    DenseState nst;
    int32_t rc = dense_state_alloc(ctx, nst, nlo, nhi - nlo + 1);
    if (rc != FSX_OK) return rc;
    if (st.nslots > 0) {
        const int64_t off = st.lo - nlo;
        FSX_ASYNC(ctx, cudaMemcpyAsync(nst.sum + off, st.sum, (size_t)st.nslots * 8, cudaMemcpyDeviceToDevice, ctx->stream));
        FSX_ASYNC(ctx, cudaMemcpyAsync(nst.cnt + off, st.cnt, (size_t)st.nslots * 8, cudaMemcpyDeviceToDevice, ctx->stream));
        FSX_ASYNC(ctx, cudaMemcpyAsync(nst.mm + 2 * off, st.mm, (size_t)st.nslots * 16, cudaMemcpyDeviceToDevice, ctx->stream));
        dense_state_free(ctx, st);
    }
    st = nst;
    return FSX_OK;
}

bool host_pinned(const void *p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { (void)cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

template <typename KeyT, typename ValT>
int32_t groupby_host_impl(fsx_ctx *ctx, int32_t key_dtype, int32_t val_dtype, const KeyT *hk, const ValT *hv, int64_t n, int64_t chunk, fsx_table **out,
                          const DenseState *fixed = nullptr) {
    // fixed: accumulate into the caller's state over its agreed key range (no per-chunk range pass, no emit)
    const int64_t nchunks = (n + chunk - 1) / chunk;
    KeyT *dk = nullptr;
    ValT *dv = nullptr;
    DenseState st;
    if (fixed) st = *fixed;
    std::vector<cudaEvent_t> evs((size_t)nchunks + 2, nullptr);
    int32_t rc = fsx_dev_alloc(ctx, (size_t)n * sizeof(KeyT), (void **)&dk);
    if (rc == FSX_OK) rc = fsx_dev_alloc(ctx, (size_t)n * sizeof(ValT), (void **)&dv);
    cudaError_t e = cudaSuccess;
    auto cleanup = [&]() {
        cudaStreamSynchronize(ctx->copy_stream);                    // no copy may outlive its destination
        for (cudaEvent_t ev : evs) if (ev) cudaEventDestroy(ev);
        if (!fixed) dense_state_free(ctx, st);
        fsx_dev_free(ctx, dk); fsx_dev_free(ctx, dv);
    };
    if (rc != FSX_OK) { cleanup(); return rc; }
    for (size_t i = 0; i < evs.size() && e == cudaSuccess; i++)
        e = cudaEventCreate(&evs[i]);
    // evs[0..nchunks-1]: chunk c landed (evs[nchunks-1] is the last copy); evs[nchunks]: first copy may start / copy clock start
    cudaEvent_t ev_start = evs[(size_t)nchunks], ev_alloc = evs[(size_t)nchunks + 1];
    if (e == cudaSuccess) e = cudaEventRecord(ev_alloc, ctx->stream);          // the stream-ordered allocations are usable after this
    if (e == cudaSuccess) e = cudaStreamWaitEvent(ctx->copy_stream, ev_alloc, 0);
    if (e == cudaSuccess) e = cudaEventRecord(ev_start, ctx->copy_stream);
    for (int64_t c = 0; c < nchunks && e == cudaSuccess; c++) {
        const int64_t off = c * chunk, len = (n - off < chunk) ? n - off : chunk;
        e = cudaMemcpyAsync(dk + off, hk + off, (size_t)len * sizeof(KeyT), cudaMemcpyHostToDevice, ctx->copy_stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(dv + off, hv + off, (size_t)len * sizeof(ValT), cudaMemcpyHostToDevice, ctx->copy_stream);
        if (e == cudaSuccess) e = cudaEventRecord(evs[(size_t)c], ctx->copy_stream);
    }
    if (e != cudaSuccess) { cleanup(); FSX_CUDA(ctx, e); }
    ctx->timing.h2d_bytes += n * (int64_t)(sizeof(KeyT) + sizeof(ValT));
    const int64_t max_slots = n * 8 > (1LL << 22) ? (n * 8 < (1LL << 27) ? n * 8 : (1LL << 27)) : (1LL << 22);
    double acc_ms = 0.0;
    int e_base = 0;
    if (fixed) FSX_ASYNC(ctx, cudaMemsetAsync(ctx->d_scalars + 9, 0, 8, ctx->stream));
    for (int64_t c = 0; c < nchunks; c++) {
        const int64_t off = c * chunk, len = (n - off < chunk) ? n - off : chunk;
        e = cudaStreamWaitEvent(ctx->stream, evs[(size_t)c], 0);
        if (!fixed) {
            // key range of the chunk (it is still in L2 from the copy), then make the state cover it
            const unsigned long long init[8] = {0, 0, ~0ULL, 0, 0, 0, 0, 0};
            if (e == cudaSuccess) e = cudaMemcpyAsync(ctx->d_scalars, init, sizeof init, cudaMemcpyHostToDevice, ctx->stream);
            if (e != cudaSuccess) { cleanup(); FSX_CUDA(ctx, e); }
            k_col_minmax<KeyT><<<fsx_grid_for(ctx, len, 256 * 8, 8), 256, 0, ctx->stream>>>(dk + off, nullptr, len, ctx->d_scalars);
            ctx->launches++;
            rc = fsx_read_scalars(ctx, 0, 8);
            if (rc == FSX_OK) rc = dense_state_cover(ctx, st, (int64_t)(ctx->h_scalars[2] ^ MSB), (int64_t)(ctx->h_scalars[3] ^ MSB), max_slots);
            if (rc != FSX_OK) {
                cleanup();
                if (rc == FSX_E_UNSUPPORTED)
                    return fsx_fail(ctx, rc, "fsx_groupby_host: the key range is too sparse for the streamed dense route; use fsx_column_upload + fsx_groupby_agg");
                return rc;
            }
        } else if (e != cudaSuccess) { cleanup(); FSX_CUDA(ctx, e); }
        if (c == 0 && std::is_floating_point<ValT>::value) {       // fixed-point window of the low-cardinality form: sampled once, on the first chunk
            rc = dense_pick_scale<ValT>(ctx, dv, len, &e_base);
            if (rc != FSX_OK) { cleanup(); return rc; }
        }
        FSX_ASYNC(ctx, cudaEventRecord(ctx->ev_c, ctx->stream));
        e = dense_launch<KeyT, ValT>(ctx, dk + off, dv + off, len, (long long)st.lo, st.nslots, e_base, st.sum, st.cnt, st.mm, ctx->d_scalars + 9);
        if (e != cudaSuccess) { cleanup(); FSX_CUDA(ctx, e); }
        FSX_ASYNC(ctx, cudaEventRecord(ctx->ev_d, ctx->stream));
        if (c + 1 == nchunks || (c & 3) == 3) {                     // sample the kernel clock without stalling every chunk
            float ms = 0.f;
            if (cudaEventSynchronize(ctx->ev_d) == cudaSuccess && cudaEventElapsedTime(&ms, ctx->ev_c, ctx->ev_d) == cudaSuccess) acc_ms += ms;
        }
    }
    e = cudaGetLastError();
    if (e != cudaSuccess) { cleanup(); FSX_CUDA(ctx, e); }
    if (fixed) {
        rc = fsx_read_scalars(ctx, 9, 1);
        if (rc == FSX_OK && ctx->h_scalars[9] != 0) {
            cleanup();
            return fsx_fail(ctx, FSX_E_DATA, "fsx_groupby_dense_accumulate_host: a key lies outside [%lld, %lld)", (long long)st.lo, (long long)(st.lo + st.nslots));
        }
    } else
        rc = dense_emit_impl(ctx, key_dtype, val_dtype, st.lo, st.nslots, st.sum, (const int64_t *)st.cnt, (const int64_t *)st.mm, out);
    float h2d = 0.f;
    if (cudaEventSynchronize(evs[(size_t)nchunks - 1]) == cudaSuccess) cudaEventElapsedTime(&h2d, ev_start, evs[(size_t)nchunks - 1]);
    ctx->timing.main_kernel_ms = acc_ms;                               // sampled accumulate kernels (every 4th chunk and the last)
    cleanup();
    ctx->timing.h2d_ms = h2d;                                          // overlaps kernel_ms: see fsx.h
    return rc;
}

} // namespace

static int32_t groupby_host_dispatch(fsx_ctx *ctx, int32_t key_dtype, const void *keys_host, int32_t val_dtype, const void *vals_host,
                                     int64_t n, int64_t chunk_rows, fsx_table **out, const DenseState *fixed) {
    const int64_t chunk = chunk_rows > 0 ? chunk_rows : (8LL << 20);
    const bool k32 = key_dtype == FSX_I32, v32 = val_dtype == FSX_F32;
    if (k32 && v32) return groupby_host_impl<int32_t, float>(ctx, key_dtype, val_dtype, (const int32_t *)keys_host, (const float *)vals_host, n, chunk, out, fixed);
    if (k32) return groupby_host_impl<int32_t, double>(ctx, key_dtype, val_dtype, (const int32_t *)keys_host, (const double *)vals_host, n, chunk, out, fixed);
    if (v32) return groupby_host_impl<long long, float>(ctx, key_dtype, val_dtype, (const long long *)keys_host, (const float *)vals_host, n, chunk, out, fixed);
    return groupby_host_impl<long long, double>(ctx, key_dtype, val_dtype, (const long long *)keys_host, (const double *)vals_host, n, chunk, out, fixed);
}

int32_t fsx_internal_dense_accumulate_host(fsx_ctx *ctx, int32_t key_dtype, const void *keys_host, int32_t val_dtype, const void *vals_host,
                                           int64_t n, int64_t chunk_rows, int64_t kmin, int64_t nslots, double *sum_dev, int64_t *cnt_dev, int64_t *mm_dev) {
    FSX_REQUIRE(ctx, n >= 0 && chunk_rows >= 0 && nslots > 0 && sum_dev && cnt_dev && mm_dev && (n == 0 || (keys_host && vals_host)), FSX_E_INVALID,
                "fsx_groupby_dense_accumulate_host: bad argument");
    FSX_REQUIRE(ctx, key_dtype == FSX_I32 || key_dtype == FSX_I64, FSX_E_UNSUPPORTED, "fsx_groupby_dense_accumulate_host: key must be i32/i64");
    FSX_REQUIRE(ctx, val_dtype == FSX_F32 || val_dtype == FSX_F64, FSX_E_UNSUPPORTED, "fsx_groupby_dense_accumulate_host: float/double values only");
    if (n == 0) return FSX_OK;
    FSX_REQUIRE(ctx, host_pinned(keys_host) && host_pinned(vals_host), FSX_E_UNSUPPORTED, "fsx_groupby_dense_accumulate_host: pinned host buffers only");
    DenseState st;
    st.sum = sum_dev; st.cnt = (unsigned long long *)cnt_dev; st.mm = (long long *)mm_dev; st.lo = kmin; st.nslots = nslots;
    return groupby_host_dispatch(ctx, key_dtype, keys_host, val_dtype, vals_host, n, chunk_rows, nullptr, &st);
}

extern "C" int32_t fsx_groupby_dense_accumulate_host(fsx_ctx *ctx, int32_t key_dtype, const void *keys_host, int32_t val_dtype,
                                                     const void *vals_host, int64_t n, int64_t chunk_rows, int64_t kmin, int64_t nslots,
                                                     double *sum_dev, int64_t *cnt_dev, int64_t *mm_dev) {
    FSX_BEGIN(ctx);
    FSX_CHECK(fsx_internal_dense_accumulate_host(ctx, key_dtype, keys_host, val_dtype, vals_host, n, chunk_rows, kmin, nslots, sum_dev, cnt_dev, mm_dev));
    FSX_END();
}

extern "C" int32_t fsx_groupby_host(fsx_ctx *ctx, int32_t key_dtype, const void *keys_host, int32_t val_dtype, const void *vals_host,
                                    int64_t n, int64_t chunk_rows, fsx_table **out) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, out && n >= 0 && chunk_rows >= 0 && (n == 0 || (keys_host && vals_host)), FSX_E_INVALID, "fsx_groupby_host: bad argument");
    *out = nullptr;
    FSX_REQUIRE(ctx, key_dtype == FSX_I32 || key_dtype == FSX_I64, FSX_E_UNSUPPORTED, "fsx_groupby_host: key must be i32/i64");
    FSX_REQUIRE(ctx, val_dtype == FSX_F32 || val_dtype == FSX_F64, FSX_E_UNSUPPORTED,
                "fsx_groupby_host: float/double values only (integer SUM is a wrapping int64: use fsx_column_upload + fsx_groupby_agg)");
    if (n == 0) { FSX_CHECK(dense_emit_impl(ctx, key_dtype, val_dtype, 0, 0, nullptr, nullptr, nullptr, out)); FSX_END(); }
    FSX_REQUIRE(ctx, host_pinned(keys_host) && host_pinned(vals_host), FSX_E_UNSUPPORTED,
                "fsx_groupby_host: pinned host buffers only (fsx_host_alloc / cudaHostRegister); use fsx_column_upload + fsx_groupby_agg for pageable memory");
    int32_t rc = groupby_host_dispatch(ctx, key_dtype, keys_host, val_dtype, vals_host, n, chunk_rows, out, nullptr);
    FSX_CHECK(rc);
    FSX_END();
}
