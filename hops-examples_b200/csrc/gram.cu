// gram.cu -- K7 shifted Gram matrix and K8 Pearson normalisation.
//
// Replaces the correlation part of the statistics hsfs computes on `fg.save(df)` when the
// feature group is built with `new StatisticsConfig(true, true, true)` (third flag;
// ComputeFeatures.scala:114,129,170,226,273,303,325).  Upstream: MLlib Correlation.corr ->
// RowMatrix.computeCovariance -> PearsonCorrelation.computeCorrelationMatrixFromCovariance
// (SURVEY.md App. A.4.3).
//
// Formulation: one pass over the columnar table producing, for pilot-shifted columns,
//     S = sum_k (x_k - c)            (d)
//     G = sum_k (x_k - c)(x_k - c)^T (d x d, symmetric: only tiles with ti <= tj are computed)
// then cov = (G - S S^T / n) / (n - 1) in FP64 (benign: c ~ column mean), then MLlib's
// normalisation (sigma = sqrt(cov_ii), |cov_ii| <= 1e-12 -> 0, NaN for zero-variance pairs,
// unit diagonal).
//
// This file holds the FP64 CUDA-core kernel: the parity anchor, valid for any (mixed) dtypes, any
// column placement and any d.  Each CTA owns one 64x64 output tile and one slab of rows; a
// 64-column x 32-row panel is staged (shifted, widened to FP64, transposed) in shared memory
// and every thread keeps a 4x4 register block.  Row slabs are combined with FP64 atomics
// (RED.ADD.F64) into the d x d result.  Algorithmic flops: 2*n*d^2 (half with symmetry);
// algorithmic bytes: n*d*sizeof(T).
#include "stats_common.cuh"

int32_t fsx_internal_gram_tc(fsx_ctx *ctx, const fsx_col *const *cols, int32_t d, const double *shift_host,
                             const fsx_moments *mom_in, double *d_gram, double *d_colsum, double *d_shift, int *used, fsx_gram_opt *opt);
int32_t fsx_internal_gram_sample(fsx_ctx *ctx, const fsx_col *const *cols, int32_t d, double *smin, double *smax, double *ssum,
                                 long long *scnt, long long *sbad, int *ok);

// centre and half-width for the optimistic scaling from (possibly all-reduced) sample statistics: the centre is the sample
// mean rounded to float, the half-width 8x the sample's spread around it (three binades of headroom; q keeps 27+ bits)
bool fsx_gram_bounds_from_sample(int32_t d, const double *smin, const double *smax, const double *ssum, const long long *scnt,
                                 const long long *sbad, double *centre, double *half) {
    for (int i = 0; i < d; i++) {
        if (sbad[i] != 0 || scnt[i] <= 0) return false;
        const float cf = (float)(ssum[i] / (double)scnt[i]);
        if (!std::isfinite(cf)) return false;
        centre[i] = (double)cf;
        double spread = fmax(smax[i] - centre[i], centre[i] - smin[i]);
        if (!(spread > 0.0)) spread = fmax(fabs(centre[i]) * 1e-6, 1e-30);       // a constant sample: any other value ends the run
        half[i] = spread * 8.0;
        if (!std::isfinite(half[i])) return false;
    }
    return true;
}

namespace {

constexpr int GT = 64;       // output tile edge
constexpr int GK = 32;       // rows staged per step
constexpr int GTHREADS = 256;

__global__ void __launch_bounds__(GTHREADS) k_gram_f64(const ColDesc *__restrict__ cols, int d,
                                                       const double *__restrict__ shift,
                                                       int64_t rows_per_block,
                                                       const int2 *__restrict__ tiles,
                                                       double *__restrict__ gram,
                                                       double *__restrict__ colsum) {
    __shared__ double As[GK][GT + 2];
    __shared__ double Bs[GK][GT + 2];
    const int2 tile = tiles[blockIdx.y];
    const int ti = tile.x, tj = tile.y;
    const bool diag = ti == tj;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const int64_t n = cols[0].n;
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
    int64_t r1 = r0 + rows_per_block;
    if (r1 > n) r1 = n;
    double acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
        for (int b = 0; b < 4; b++) acc[a][b] = 0.0;
    double csum[4] = {0, 0, 0, 0};

    // staging map: 8 lanes cover 8 consecutive rows of one column (coalesced 32/64 B runs)
    const int s_col = threadIdx.x >> 3;  // 0..31, two passes cover 64 columns
    const int s_row = threadIdx.x & 7;   // 0..7, four passes cover 32 rows
    for (int64_t k0 = r0; k0 < r1; k0 += GK) {
#pragma unroll
        for (int cp = 0; cp < 2; cp++) {
            const int lc = s_col + cp * 32;
            const int ca = ti * GT + lc, cb = tj * GT + lc;
            const bool has_a = ca < d, has_b = !diag && cb < d;
            const ColDesc da = cols[has_a ? ca : 0], db = cols[has_b ? cb : 0];
            const double sa = ca < d ? shift[ca] : 0.0, sb = (!diag && cb < d) ? shift[cb] : 0.0;
#pragma unroll
            for (int rp = 0; rp < 4; rp++) {
                const int lr = s_row + rp * 8;
                const int64_t k = k0 + lr;
                const bool in = k < r1;
                As[lr][lc] = (has_a && in) ? load_f64(da, k) - sa : 0.0;
                if (!diag) Bs[lr][lc] = (has_b && in) ? load_f64(db, k) - sb : 0.0;
            }
        }
        __syncthreads();
        const double(*Bp)[GT + 2] = diag ? As : Bs;
#pragma unroll 8
        for (int k = 0; k < GK; k++) {
            double a[4], b[4];
#pragma unroll
            for (int q = 0; q < 4; q++) { a[q] = As[k][ty * 4 + q]; b[q] = Bp[k][tx * 4 + q]; }
#pragma unroll
            for (int x = 0; x < 4; x++)
#pragma unroll
                for (int y = 0; y < 4; y++) acc[x][y] = fma(a[x], b[y], acc[x][y]);
            if (diag && ty == 0) {
#pragma unroll
                for (int q = 0; q < 4; q++) csum[q] += b[q];
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int x = 0; x < 4; x++) {
        const int gi = ti * GT + ty * 4 + x;
        if (gi >= d) continue;
#pragma unroll
        for (int y = 0; y < 4; y++) {
            const int gj = tj * GT + tx * 4 + y;
            if (gj >= d) continue;
            if (diag && gj < gi) continue;         // upper triangle only
            atomicAdd(&gram[(size_t)gi * d + gj], acc[x][y]);
        }
    }
    if (diag && ty == 0) {
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const int gj = tj * GT + tx * 4 + q;
            if (gj < d) atomicAdd(&colsum[gj], csum[q]);
        }
    }
}

__global__ void k_gram_mirror(double *__restrict__ gram, int d) {
    const int i = blockIdx.y * blockDim.y + threadIdx.y, j = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < d && j < d && j < i) gram[(size_t)i * d + j] = gram[(size_t)j * d + i];
}

// ---- K8 -------------------------------------------------------------------------------------
__global__ void k_cov_sigma(const double *__restrict__ gram, const double *__restrict__ colsum,
                            int d, double n, double *__restrict__ sigma) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= d) return;
    const double cov = (gram[(size_t)i * d + i] - colsum[i] * colsum[i] / n) / (n - 1.0);
    sigma[i] = fabs(cov) <= 1e-12 ? 0.0 : sqrt(cov);
}

__global__ void k_pearson(const double *__restrict__ gram, const double *__restrict__ colsum,
                          const double *__restrict__ sigma, int d, double n,
                          double *__restrict__ corr) {
    const int i = blockIdx.y * blockDim.y + threadIdx.y, j = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= d || j >= d) return;
    double r;
    if (i == j) r = 1.0;
    else {
        // evaluate from the upper triangle so corr is exactly symmetric
        const int a = i < j ? i : j, b = i < j ? j : i;
        const double cov = (gram[(size_t)a * d + b] - colsum[a] * colsum[b] / n) / (n - 1.0);
        const double sa = sigma[a], sb = sigma[b];
        r = (sa == 0.0 || sb == 0.0) ? __longlong_as_double(0x7FF8000000000000LL) : cov / (sb * sa);
    }
    corr[(size_t)i * d + j] = r;
}

struct DevBufs {
    fsx_ctx *ctx;
    std::vector<void *> bufs;
    explicit DevBufs(fsx_ctx *c) : ctx(c) {}
    ~DevBufs() { for (void *b : bufs) fsx_dev_free(ctx, b); }
    int32_t alloc(size_t bytes, void **out) {
        int32_t rc = fsx_dev_alloc(ctx, bytes, out);
        if (rc == FSX_OK) bufs.push_back(*out);
        return rc;
    }
};

// device-side gram: d_gram (d*d), d_colsum (d), d_shift (d) stay on the device
int32_t gram_device(fsx_ctx *ctx, DevBufs &bufs, const fsx_col *const *cols, int32_t d,
                    const double *shift_host, double **d_gram, double **d_colsum,
                    double **d_shift, int64_t *n_out, const fsx_moments *mom_in = nullptr, fsx_gram_opt *opt_in = nullptr) {
    int64_t n = 0;
    FSX_CHECK(check_cols(ctx, cols, d, "fsx_stats_gram", false, &n));
    for (int i = 0; i < d; i++)
        FSX_REQUIRE(ctx, cols[i]->validity == nullptr, FSX_E_UNSUPPORTED,
                    "fsx_stats_gram: column %d has a validity bitmap; the Gram formulation needs "
                    "null-free columns (pairwise deletion is not implemented)", i);
    ColDesc *d_cols = nullptr;
    FSX_CHECK(upload_desc(ctx, cols, d, &d_cols));
    bufs.bufs.push_back(d_cols);
    FSX_CHECK(bufs.alloc(sizeof(double) * (size_t)d * (size_t)d, (void **)d_gram));
    FSX_CHECK(bufs.alloc(sizeof(double) * (size_t)d, (void **)d_colsum));
    FSX_CHECK(bufs.alloc(sizeof(double) * (size_t)d, (void **)d_shift));
    FSX_CUDA(ctx, cudaMemsetAsync(*d_gram, 0, sizeof(double) * (size_t)d * (size_t)d, ctx->stream));
    FSX_CUDA(ctx, cudaMemsetAsync(*d_colsum, 0, sizeof(double) * (size_t)d, ctx->stream));
    *n_out = n;
    if (n == 0) return FSX_OK;
    {   // float32 tables of >= 2^20 rows go to the tensor cores (gram_tc.cu); everything else stays here
        int used = 0;
        const char *env = getenv("FSX_GRAM_TC");
        const bool tc = !(env && env[0] == '0');
        if (tc && opt_in) {                                       // the caller (multi-GPU) agreed on the bounds; it also handles a miss
            FSX_CHECK(fsx_internal_gram_tc(ctx, cols, d, nullptr, nullptr, *d_gram, *d_colsum, *d_shift, &used, opt_in));
            if (!used) opt_in->missed = 1;
        } else if (tc && !mom_in && !shift_host && d > 128 && !getenv("FSX_GRAM_NO_OPTIMISTIC")) {
            // wide tables: no describe pass -- scale and centre from a sample, every value verified by the conversion kernel
            std::vector<double> smin((size_t)d), smax((size_t)d), ssum((size_t)d), centre((size_t)d), half((size_t)d);
            std::vector<long long> scnt((size_t)d), sbad((size_t)d);
            int ok = 0;
            FSX_CHECK(fsx_internal_gram_sample(ctx, cols, d, smin.data(), smax.data(), ssum.data(), scnt.data(), sbad.data(), &ok));
            if (ok && fsx_gram_bounds_from_sample(d, smin.data(), smax.data(), ssum.data(), scnt.data(), sbad.data(), centre.data(), half.data())) {
                fsx_gram_opt opt;
                opt.centre = centre.data(); opt.half = half.data();
                FSX_CHECK(fsx_internal_gram_tc(ctx, cols, d, nullptr, nullptr, *d_gram, *d_colsum, *d_shift, &used, &opt));
                if (used && opt.missed) {                         // a value outside the sampled range: start over with exact bounds
                    used = 0;
                    FSX_CUDA(ctx, cudaMemsetAsync(*d_gram, 0, sizeof(double) * (size_t)d * (size_t)d, ctx->stream));
                    FSX_CUDA(ctx, cudaMemsetAsync(*d_colsum, 0, sizeof(double) * (size_t)d, ctx->stream));
                }
            }
        }
        if (tc && !used && !opt_in) FSX_CHECK(fsx_internal_gram_tc(ctx, cols, d, shift_host, mom_in, *d_gram, *d_colsum, *d_shift, &used, nullptr));
        if (used) {
            dim3 mb2(16, 16), mg2((unsigned)((d + 15) / 16), (unsigned)((d + 15) / 16));
            FSX_LAUNCH(ctx, k_gram_mirror, mg2, mb2, 0, *d_gram, d);
            return FSX_OK;
        }
    }
    if (shift_host) {
        FSX_CUDA(ctx, cudaMemcpyAsync(*d_shift, shift_host, sizeof(double) * (size_t)d, cudaMemcpyHostToDevice, ctx->stream));
        FSX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    } else {
        FSX_LAUNCH(ctx, k_pilot, d, 256, 0, d_cols, *d_shift);
    }
    *n_out = n;
    if (n == 0) return FSX_OK;
    // upper-triangular tile list
    const int nt = (d + GT - 1) / GT;
    std::vector<int2> tiles;
    for (int i = 0; i < nt; i++)
        for (int j = i; j < nt; j++) tiles.push_back(make_int2(i, j));
    int2 *d_tiles = nullptr;
    FSX_CHECK(bufs.alloc(sizeof(int2) * tiles.size(), (void **)&d_tiles));
    FSX_CUDA(ctx, cudaMemcpyAsync(d_tiles, tiles.data(), sizeof(int2) * tiles.size(), cudaMemcpyHostToDevice, ctx->stream));
    FSX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    // row slabs: enough CTAs to fill the machine ~4x, slabs a multiple of GK rows
    int64_t want = ((int64_t)ctx->sm_count * 4 + (int64_t)tiles.size() - 1) / (int64_t)tiles.size();
    if (want < 1) want = 1;
    int64_t rpb = (n + want - 1) / want;
    rpb = ((rpb + GK - 1) / GK) * GK;
    if (rpb < 1024) rpb = 1024;
    const int64_t nslabs = (n + rpb - 1) / rpb;
    dim3 grid((unsigned)nslabs, (unsigned)tiles.size());
    FSX_LAUNCH(ctx, k_gram_f64, grid, GTHREADS, 0, d_cols, d, *d_shift, rpb, d_tiles, *d_gram, *d_colsum);
    dim3 mb(16, 16), mg((unsigned)((d + 15) / 16), (unsigned)((d + 15) / 16));
    FSX_LAUNCH(ctx, k_gram_mirror, mg, mb, 0, *d_gram, d);
    return FSX_OK;
}

int32_t pearson_device(fsx_ctx *ctx, DevBufs &bufs, const double *d_gram, const double *d_colsum,
                       int32_t d, int64_t n, double **d_corr) {
    FSX_REQUIRE(ctx, n > 1, FSX_E_DATA, "Pearson correlation needs more than one row (n=%lld)", (long long)n);
    double *d_sigma = nullptr;
    FSX_CHECK(bufs.alloc(sizeof(double) * (size_t)d, (void **)&d_sigma));
    FSX_CHECK(bufs.alloc(sizeof(double) * (size_t)d * (size_t)d, (void **)d_corr));
    FSX_LAUNCH(ctx, k_cov_sigma, (d + 127) / 128, 128, 0, d_gram, d_colsum, d, (double)n, d_sigma);
    dim3 mb(16, 16), mg((unsigned)((d + 15) / 16), (unsigned)((d + 15) / 16));
    FSX_LAUNCH(ctx, k_pearson, mg, mb, 0, d_gram, d_colsum, d_sigma, d, (double)n, *d_corr);
    return FSX_OK;
}

} // namespace

extern "C" int32_t fsx_stats_gram(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols,
                                  const double *shift, double *gram_out, double *colsum_out,
                                  double *shift_out, int64_t *n_out) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, gram_out && colsum_out && n_out, FSX_E_INVALID, "fsx_stats_gram: NULL output");
    DevBufs bufs(ctx);
    double *d_gram = nullptr, *d_colsum = nullptr, *d_shift = nullptr;
    FSX_CHECK(gram_device(ctx, bufs, cols, ncols, shift, &d_gram, &d_colsum, &d_shift, n_out));
    const size_t d = (size_t)ncols;
    FSX_CUDA(ctx, cudaMemcpyAsync(gram_out, d_gram, 8 * d * d, cudaMemcpyDeviceToHost, ctx->stream));
    FSX_CUDA(ctx, cudaMemcpyAsync(colsum_out, d_colsum, 8 * d, cudaMemcpyDeviceToHost, ctx->stream));
    if (shift_out) FSX_CUDA(ctx, cudaMemcpyAsync(shift_out, d_shift, 8 * d, cudaMemcpyDeviceToHost, ctx->stream));
    FSX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->timing.d2h_bytes += (int64_t)(8 * d * d + 16 * d);
    FSX_END();
}

extern "C" int32_t fsx_stats_pearson_from_gram(fsx_ctx *ctx, const double *gram, const double *colsum,
                                               int32_t d, int64_t n, double *corr_out) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, gram && colsum && corr_out && d > 0, FSX_E_INVALID, "fsx_stats_pearson_from_gram: bad argument");
    DevBufs bufs(ctx);
    double *d_gram = nullptr, *d_colsum = nullptr, *d_corr = nullptr;
    const size_t dd = (size_t)d;
    FSX_CHECK(bufs.alloc(8 * dd * dd, (void **)&d_gram));
    FSX_CHECK(bufs.alloc(8 * dd, (void **)&d_colsum));
    FSX_CUDA(ctx, cudaMemcpyAsync(d_gram, gram, 8 * dd * dd, cudaMemcpyHostToDevice, ctx->stream));
    FSX_CUDA(ctx, cudaMemcpyAsync(d_colsum, colsum, 8 * dd, cudaMemcpyHostToDevice, ctx->stream));
    FSX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    FSX_CHECK(pearson_device(ctx, bufs, d_gram, d_colsum, d, n, &d_corr));
    FSX_CUDA(ctx, cudaMemcpyAsync(corr_out, d_corr, 8 * dd * dd, cudaMemcpyDeviceToHost, ctx->stream));
    FSX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    FSX_END();
}

// ---- pieces of the multi-rank Pearson (csrc/dist.cu): the local shifted Gram stays on the device so that NCCL can add
// the ranks' partials in place; mom carries the bounds that scale the tensor-core digits (global min/max, so every rank
// uses the same scale and shift) and the LOCAL mean (S = n_local * (mean_local - c)).
int32_t fsx_internal_gram_partial(fsx_ctx *ctx, const fsx_col *const *cols, int32_t d, const double *shift_host, const fsx_moments *mom,
                                  double **d_gram, double **d_colsum, int64_t *n_out, fsx_gram_opt *opt) {
    DevBufs bufs(ctx);
    double *d_shift = nullptr;
    FSX_CHECK(gram_device(ctx, bufs, cols, d, shift_host, d_gram, d_colsum, &d_shift, n_out, mom, opt));
    // hand the two result buffers to the caller, release the rest
    for (void *&b : bufs.bufs) if (b == (void *)*d_gram || b == (void *)*d_colsum) b = nullptr;
    return FSX_OK;
}

int32_t fsx_internal_pearson_finish(fsx_ctx *ctx, const double *d_gram, const double *d_colsum, int32_t d, int64_t n, double *corr_out) {
    DevBufs bufs(ctx);
    double *d_corr = nullptr;
    FSX_CHECK(pearson_device(ctx, bufs, d_gram, d_colsum, d, n, &d_corr));
    FSX_CUDA(ctx, cudaMemcpyAsync(corr_out, d_corr, 8 * (size_t)d * (size_t)d, cudaMemcpyDeviceToHost, ctx->stream));
    FSX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->timing.d2h_bytes += (int64_t)(8 * (size_t)d * (size_t)d);
    return FSX_OK;
}

// correlations with a describe pass the caller already ran (fsx_stats_profile)
int32_t fsx_internal_pearson_with_moments(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols, const fsx_moments *mom,
                                          double *corr_out) {
    DevBufs bufs(ctx);
    double *d_gram = nullptr, *d_colsum = nullptr, *d_shift = nullptr, *d_corr = nullptr;
    int64_t n = 0;
    FSX_CHECK(gram_device(ctx, bufs, cols, ncols, nullptr, &d_gram, &d_colsum, &d_shift, &n, mom));
    FSX_CHECK(pearson_device(ctx, bufs, d_gram, d_colsum, ncols, n, &d_corr));
    const size_t d = (size_t)ncols;
    FSX_CUDA(ctx, cudaMemcpyAsync(corr_out, d_corr, 8 * d * d, cudaMemcpyDeviceToHost, ctx->stream));
    FSX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->timing.d2h_bytes += (int64_t)(8 * d * d);
    return FSX_OK;
}

extern "C" int32_t fsx_stats_pearson(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols,
                                     double *corr_out) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, corr_out, FSX_E_INVALID, "fsx_stats_pearson: corr_out is NULL");
    DevBufs bufs(ctx);
    double *d_gram = nullptr, *d_colsum = nullptr, *d_shift = nullptr, *d_corr = nullptr;
    int64_t n = 0;
    FSX_CHECK(gram_device(ctx, bufs, cols, ncols, nullptr, &d_gram, &d_colsum, &d_shift, &n));
    FSX_CHECK(pearson_device(ctx, bufs, d_gram, d_colsum, ncols, n, &d_corr));
    const size_t d = (size_t)ncols;
    FSX_CUDA(ctx, cudaMemcpyAsync(corr_out, d_corr, 8 * d * d, cudaMemcpyDeviceToHost, ctx->stream));
    FSX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->timing.d2h_bytes += (int64_t)(8 * d * d);
    FSX_END();
}
