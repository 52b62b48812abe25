// stats_common.cuh -- column descriptors and helpers shared by stats.cu (K4/K5) and gram.cu (K7/K8).
#pragma once
#include <math.h>

#include "common.cuh"

namespace {

struct ColDesc {
    const void *data;
    const uint8_t *valid;
    int64_t n;
    int dtype;
    int pad;
};

constexpr int PILOT_SAMPLE = 1024;

// ---- pilot: mean of the first valid, finite elements, rounded to float ----------------------
__device__ __forceinline__ double load_f64(const ColDesc &c, int64_t i) {
    switch (c.dtype) {
    case FSX_I32: return (double)((const int32_t *)c.data)[i];
    case FSX_I64: return (double)((const long long *)c.data)[i];
    case FSX_F32: return (double)((const float *)c.data)[i];
    default: return ((const double *)c.data)[i];
    }
}

__global__ void __launch_bounds__(256) k_pilot(const ColDesc *__restrict__ cols, double *__restrict__ pilot) {
    const ColDesc c = cols[blockIdx.x];
    __shared__ double ssum[8];
    __shared__ unsigned scnt[8];
    double s = 0.0;
    unsigned k = 0;
    int64_t lim = c.n < PILOT_SAMPLE ? c.n : PILOT_SAMPLE;
    for (int64_t i = threadIdx.x; i < lim; i += blockDim.x) {
        if (!dev_bit_valid(c.valid, i)) continue;
        double v = load_f64(c, i);
        if (isfinite(v)) { s += v; k++; }
    }
    s = warp_sum(s);
    k = warp_sum(k);
    if ((threadIdx.x & 31) == 0) { ssum[threadIdx.x >> 5] = s; scnt[threadIdx.x >> 5] = k; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0; unsigned kk = 0;
        for (int w = 0; w < 8; w++) { t += ssum[w]; kk += scnt[w]; }
        double m = kk ? t / (double)kk : 0.0;
        float mf = (float)m;                      // n*c stays exact for n < 2^29
        pilot[blockIdx.x] = isfinite(mf) ? (double)mf : 0.0;
    }
}

int32_t upload_desc(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols, ColDesc **d_out) {
    std::vector<ColDesc> h((size_t)ncols);
    for (int i = 0; i < ncols; i++) {
        h[(size_t)i].data = cols[i]->data; h[(size_t)i].valid = cols[i]->validity;
        h[(size_t)i].n = cols[i]->n; h[(size_t)i].dtype = cols[i]->dtype; h[(size_t)i].pad = 0;
    }
    FSX_CHECK(fsx_dev_alloc(ctx, sizeof(ColDesc) * (size_t)ncols, (void **)d_out));
    FSX_CUDA(ctx, cudaMemcpyAsync(*d_out, h.data(), sizeof(ColDesc) * (size_t)ncols, cudaMemcpyHostToDevice, ctx->stream));
    FSX_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); // h goes out of scope
    return FSX_OK;
}

int32_t check_cols(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols, const char *who,
                   bool same_dtype, int64_t *n_out) {
    FSX_REQUIRE(ctx, cols && ncols > 0, FSX_E_INVALID, "%s: no columns", who);
    for (int i = 0; i < ncols; i++) {
        FSX_REQUIRE(ctx, cols[i], FSX_E_INVALID, "%s: column %d is NULL", who, i);
        FSX_REQUIRE(ctx, cols[i]->n == cols[0]->n, FSX_E_INVALID, "%s: column %d has %lld rows, column 0 has %lld",
                    who, i, (long long)cols[i]->n, (long long)cols[0]->n);
        if (same_dtype)
            FSX_REQUIRE(ctx, cols[i]->dtype == cols[0]->dtype, FSX_E_UNSUPPORTED,
                        "%s: mixed dtypes in one call (column %d)", who, i);
    }
    *n_out = cols[0]->n;
    return FSX_OK;
}


} // namespace
