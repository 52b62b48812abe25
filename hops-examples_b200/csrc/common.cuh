// common.cuh -- context, handles, error plumbing and device helpers shared by every
// translation unit of libfsx (sm_100a only; no CPU fallback anywhere in this tree).
#pragma once
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>   // header-only: one NVTX range per ABI call (visible in nsys / ncu, free when no tool is attached)
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <chrono>
#include <string>
#include <vector>

#include "../../include/fsx.h"

// ------------------------------------------------------------------------------------
// handles
// ------------------------------------------------------------------------------------
struct fsx_group;   // csrc/dist.cu: the devices/ranks a context exchanges data with (NCCL communicators)

// remembered plan of the last DIRECT-slot group-by on a context (csrc/groupby.cu): the next call of the same
// shape skips the key-range pre-pass; the accumulate kernel itself verifies every key against it
struct fsx_gb_plan {
    bool valid = false;
    int32_t key_dtype = 0;
    long long kmin = 0;
    unsigned long long cap = 0;
    const void *key_ptr = nullptr;   // the key buffer the plan was measured on, and its length
    int64_t n = 0;
};

// dense group-by state kept between multi-rank steps (csrc/dist.cu): sum f64[n], cnt i64[n], mm i64[2n]
struct fsx_dense_cache {
    double *sum = nullptr;
    unsigned long long *cnt = nullptr;
    long long *mm = nullptr;
    int64_t nslots = 0;
    bool borrowed = false;   // the arrays live in a peer-mapped buffer of the group (csrc/dist.cu), not in the pool
};

// Hook of the fused join + group-by (csrc/join.cu) for the multi-GPU form (csrc/dist.cu): an agreed group range, so that
// every rank aggregates into the same slot layout, and a reduction of the slot arrays across ranks between the aggregation
// and the emit -- the ranks' partial groups merge with three all-reduces instead of an all-gather and a second group-by.
struct fsx_jg_hook {
    int32_t (*reduce)(void *user, fsx_ctx *ctx, double *g_sum, unsigned long long *g_cnt, unsigned long long *g_nn, int range) = nullptr;
    void *user = nullptr;
    bool have_range = false;
    long long gmin = 0;
    unsigned long long span = 0;
};

// Optimistic scaling of the tensor-core Gram (csrc/gram_tc.cu): instead of a describe pass over the whole table, the centre
// and the half-width every value is expected within come from a sample; the conversion kernel verifies every value and sums
// q = rint((x - c) 2^s) exactly per column (that is S), *missed says whether the run has to be repeated with exact bounds.
struct fsx_gram_opt {
    const double *centre = nullptr;   // [d]
    const double *half = nullptr;     // [d]
    int missed = 0;                   // out
};

struct fsx_ctx {
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;   // compute + ordered allocation
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_a = nullptr, ev_b = nullptr, ev_c = nullptr, ev_d = nullptr;
    std::string err;
    bool poisoned = false;
    cudaError_t async_err = cudaSuccess;   // first failure of a fire-and-forget async call (FSX_ASYNC), reported by FsxCall::end
    fsx_timing timing{};
    int64_t launches = 0;            // kernels launched since creation
    int64_t launches_at_call = 0;
    // small persistent device scratch (counters, flags) and its pinned mirror
    unsigned long long *d_scalars = nullptr;  // 64 x u64
    unsigned long long *h_scalars = nullptr;  // pinned
    // pinned staging ring for pageable uploads/downloads
    void *staging[2] = {nullptr, nullptr};
    size_t staging_bytes = 0;
    cudaEvent_t staging_ev[2] = {nullptr, nullptr};
    fsx_gb_plan gb_plan;
    fsx_dense_cache dense;
    // multi-GPU: the group this context belongs to (owned by the leader = member 0), its index in it
    fsx_group *grp = nullptr;
    int member = 0;
};

struct fsx_col {
    int32_t dtype = 0;
    int64_t n = 0;
    void *data = nullptr;        // device
    uint8_t *validity = nullptr; // device, Arrow bit order, may be null
    bool owns = true;
    int device = 0;
};

struct fsx_table {
    std::vector<fsx_col *> cols;
    int64_t nrows = 0;
};

struct fsx_join {
    int64_t n = 0;
    int64_t *left = nullptr;  // device
    int64_t *right = nullptr; // device (may be null for semi/anti)
    bool may_miss = false;    // outer joins: an index may be -1 (no partner on that side)
};

static inline size_t fsx_dtype_size(int32_t dt) {
    return (dt == FSX_I32 || dt == FSX_F32) ? 4 : 8;
}
static inline bool fsx_dtype_ok(int32_t dt) { return dt >= 0 && dt <= 3; }
static inline bool fsx_dtype_is_float(int32_t dt) { return dt == FSX_F32 || dt == FSX_F64; }

// ------------------------------------------------------------------------------------
// error plumbing: every ABI function is   FSX_BEGIN(ctx) ... FSX_END(ctx)
// ------------------------------------------------------------------------------------
extern thread_local std::string g_fsx_tls_error;

int32_t fsx_fail(fsx_ctx *ctx, int32_t code, const char *fmt, ...);

#define FSX_REQUIRE(ctx, cond, code, ...)                      \
    do {                                                       \
        if (!(cond)) return fsx_fail((ctx), (code), __VA_ARGS__); \
    } while (0)

#define FSX_CUDA(ctx, expr)                                                               \
    do {                                                                                  \
        cudaError_t _e = (expr);                                                          \
        if (_e != cudaSuccess) {                                                          \
            int32_t _c = (_e == cudaErrorMemoryAllocation) ? FSX_E_OOM : FSX_E_CUDA;      \
            if (_e != cudaErrorMemoryAllocation && _e != cudaErrorInvalidValue)           \
                (ctx)->poisoned = true;                                                   \
            (void)cudaGetLastError();                                                     \
            return fsx_fail((ctx), _c, "%s failed: %s (%s:%d)", #expr,                    \
                            cudaGetErrorString(_e), __FILE__, __LINE__);                  \
        }                                                                                 \
    } while (0)

// an async memset / memcpy / event record whose failure need not unwind the caller right here: the first error is kept and
// fails the ABI call when it ends (FsxCall::end), so no such return code is dropped
#define FSX_ASYNC(ctx, expr)                                                              \
    do {                                                                                  \
        cudaError_t _ae = (expr);                                                         \
        if (_ae != cudaSuccess && (ctx)->async_err == cudaSuccess) (ctx)->async_err = _ae; \
    } while (0)

#define FSX_CHECK(expr)                      \
    do {                                     \
        int32_t _s = (expr);                 \
        if (_s != FSX_OK) return _s;         \
    } while (0)

// kernel launch with accounting; the launch error is picked up right away
#define FSX_LAUNCH(ctx, kernel, grid, block, smem, ...)                                   \
    do {                                                                                  \
        kernel<<<(grid), (block), (smem), (ctx)->stream>>>(__VA_ARGS__);                  \
        (ctx)->launches++;                                                                \
        FSX_CUDA((ctx), cudaGetLastError());                                              \
    } while (0)

// RAII scope for one ABI call: validates the context, makes its device current and
// fills fsx_timing (host wall time + device time between the two events).
struct FsxCall {
    fsx_ctx *ctx;
    std::chrono::steady_clock::time_point t0;
    bool armed = false;
    explicit FsxCall(fsx_ctx *c, const char *name = "fsx") : ctx(c) { nvtxRangePushA(name); }
    ~FsxCall() { nvtxRangePop(); }
    FsxCall(const FsxCall &) = delete;
    int32_t begin();
    int32_t end();  // records + synchronises; returns status
};

#define FSX_BEGIN(ctx)                                               \
    if (!(ctx)) return fsx_fail(nullptr, FSX_E_INVALID, "null context"); \
    FsxCall _call(ctx, __func__);                                    \
    FSX_CHECK(_call.begin())
#define FSX_END() return _call.end()

// context lifecycle pieces shared by ctx.cu and dist.cu
int32_t fsx_ctx_create_one(int dev, fsx_ctx **out);
void fsx_ctx_destroy_one(fsx_ctx *ctx);
int32_t fsx_group_create_local(fsx_ctx *const *members, int n);   // dist.cu: one process, n devices (ncclCommInitAll)
void fsx_group_destroy(fsx_ctx *leader);                          // dist.cu: communicators + members 1..n-1

// device memory (stream-ordered pool: reuse without cudaMalloc cost on the hot path)
int32_t fsx_dev_alloc(fsx_ctx *ctx, size_t bytes, void **out);
void fsx_dev_free(fsx_ctx *ctx, void *p);
int32_t fsx_col_new(fsx_ctx *ctx, int32_t dtype, int64_t n, bool with_validity, fsx_col **out);
void fsx_col_release(fsx_ctx *ctx, fsx_col *c);
int32_t fsx_read_scalars(fsx_ctx *ctx, int first, int count); // d_scalars -> h_scalars, sync
int32_t fsx_internal_null_count(fsx_ctx *ctx, const fsx_col *col, int64_t *out); // nulls of a column, inside an ABI call

static inline int fsx_grid_for(const fsx_ctx *ctx, int64_t work_items, int per_block,
                               int blocks_per_sm) {
    int64_t need = (work_items + per_block - 1) / per_block;
    int64_t cap = (int64_t)ctx->sm_count * blocks_per_sm;
    if (need < 1) need = 1;
    return (int)(need < cap ? need : cap);
}

// ------------------------------------------------------------------------------------
// device helpers
// ------------------------------------------------------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ bool dev_bit_valid(const uint8_t *__restrict__ v, int64_t i) {
    return v == nullptr || ((v[i >> 3] >> (i & 7)) & 1);
}

// murmur3 finalisers: slot / partition hashing
__device__ __host__ __forceinline__ uint32_t mix32(uint32_t h) {
    h ^= h >> 16; h *= 0x85ebca6bu;
    h ^= h >> 13; h *= 0xc2b2ae35u;
    h ^= h >> 16; return h;
}
__device__ __host__ __forceinline__ uint64_t mix64(uint64_t x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL;
    x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL;
    x ^= x >> 33; return x;
}

// Order-preserving maps onto uint64 so min/max run as native 64-bit integer atomics.
// Floating point follows Spark's ordering: NaN is greater than every other value.
__device__ __forceinline__ unsigned long long ord_from_i64(long long v) {
    return (unsigned long long)v ^ 0x8000000000000000ULL;
}
__device__ __forceinline__ long long ord_to_i64(unsigned long long u) {
    return (long long)(u ^ 0x8000000000000000ULL);
}
__device__ __forceinline__ unsigned long long ord_from_f64(double d) {
    if (d != d) return 0xFFFFFFFFFFFFFFFFULL; // canonical NaN: above +inf
    unsigned long long b = (unsigned long long)__double_as_longlong(d);
    return (b & 0x8000000000000000ULL) ? ~b : (b | 0x8000000000000000ULL);
}
__device__ __forceinline__ double ord_to_f64(unsigned long long u) {
    if (u == 0xFFFFFFFFFFFFFFFFULL) return __longlong_as_double(0x7FF8000000000000LL);
    unsigned long long b = (u & 0x8000000000000000ULL) ? (u & 0x7FFFFFFFFFFFFFFFULL) : ~u;
    return __longlong_as_double((long long)b);
}

// streaming 128-bit load that does not pollute L1
__device__ __forceinline__ int4 ld_stream_v4(const void *p) {
    int4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
#endif
