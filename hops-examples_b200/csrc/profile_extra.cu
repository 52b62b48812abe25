// profile_extra.cu -- the frequency- and rank-based column metrics of the Deequ profile hsfs stores with a
// feature group (StatisticsConfig(true, ...), ComputeFeatures.scala:114) and of its data-validation rules
// (notebooks/featurestore/hsfs/data_validation/feature_validation_python.ipynb:218-243: HAS_NUMBER_OF_DISTINCT_VALUES,
// HAS_DISTINCTNESS, HAS_UNIQUENESS, HAS_UNIQUE_VALUE_RATIO, HAS_ENTROPY, HAS_APPROX_QUANTILE):
//
//   fsx_stats_distinct   value frequencies through K1 (group-by on the bit patterns, count(*)), then one
//                        deterministic reduction of the count column: #distinct, #values seen once, entropy
//   fsx_stats_quantiles  exact order statistics by most-significant-byte radix selection over the ordered bit
//                        patterns: 8 histogram passes over the column for any number of probabilities, no sort
//
// Both stream the column at HBM rate; neither materialises a sorted copy.
#include <math.h>

#include <string_view>
#include <unordered_map>

#include "common.cuh"
#include "sort.cuh"

int32_t fsx_internal_groupby(fsx_ctx *ctx, const fsx_col *key, const fsx_col *const *vals, const int32_t *aggs, int32_t nvals,
                             int64_t expected_groups, fsx_table **out);

namespace {

constexpr int RED_BLOCKS = 64, RED_THREADS = 256;

// fixed assignment of rows to threads, fixed shuffle/shared tree: the f64 partials do not depend on scheduling
__global__ void __launch_bounds__(RED_THREADS) k_freq_reduce(const long long *__restrict__ counts, const uint8_t *__restrict__ key_valid, int64_t g,
                                                             double inv_n, double *part_entropy, unsigned long long *part_distinct,
                                                             unsigned long long *part_unique) {
    double h = 0.0;
    unsigned long long d = 0, u = 0;
    for (int64_t i = blockIdx.x * (int64_t)RED_THREADS + threadIdx.x; i < g; i += (int64_t)RED_BLOCKS * RED_THREADS) {
        if (!dev_bit_valid(key_valid, i)) continue;            // the null group is not a value
        const long long c = counts[i];
        const double p = (double)c * inv_n;
        h -= p * log(p);
        d++;
        u += c == 1;
    }
    __shared__ double sh[RED_THREADS];
    __shared__ unsigned long long sd[RED_THREADS], su[RED_THREADS];
    sh[threadIdx.x] = h; sd[threadIdx.x] = d; su[threadIdx.x] = u;
    __syncthreads();
    for (int o = RED_THREADS / 2; o > 0; o >>= 1) {
        if (threadIdx.x < o) { sh[threadIdx.x] += sh[threadIdx.x + o]; sd[threadIdx.x] += sd[threadIdx.x + o]; su[threadIdx.x] += su[threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { part_entropy[blockIdx.x] = sh[0]; part_distinct[blockIdx.x] = sd[0]; part_unique[blockIdx.x] = su[0]; }
}

constexpr int Q_MAX = 16;

struct QState {
    int nq;
    unsigned long long prefix[Q_MAX];   // bits already decided (above `shift + 8`)
};

template <typename T> __device__ __forceinline__ unsigned long long q_ord(T v);
template <> __device__ __forceinline__ unsigned long long q_ord<int32_t>(int32_t v) { return ord_from_i64((long long)v); }
template <> __device__ __forceinline__ unsigned long long q_ord<long long>(long long v) { return ord_from_i64(v); }
template <> __device__ __forceinline__ unsigned long long q_ord<float>(float v) { return ord_from_f64((double)v); }
template <> __device__ __forceinline__ unsigned long long q_ord<double>(double v) { return ord_from_f64(v); }

// one radix pass: per probability, the 256-bin histogram of byte `shift/8` among the values whose higher bytes
// equal that probability's prefix.  Per-CTA histograms in shared memory (native 32-bit atomics), one flush.
template <typename T>
__global__ void __launch_bounds__(256) k_radix_hist(const T *__restrict__ v, const uint8_t *__restrict__ valid, int64_t n, int shift,
                                                    const __grid_constant__ QState st, unsigned long long *hist /* [nq][256] */) {
    __shared__ unsigned sh[Q_MAX * 256];
    for (int i = threadIdx.x; i < st.nq * 256; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    const unsigned long long himask = shift >= 56 ? 0ULL : ~0ULL << (shift + 8);
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (!dev_bit_valid(valid, i)) continue;
        const unsigned long long o = q_ord<T>(v[i]);
        const unsigned b = (unsigned)(o >> shift) & 255u;
        for (int q = 0; q < st.nq; q++)
            if ((o & himask) == st.prefix[q]) atomicAdd(&sh[q * 256 + b], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < st.nq * 256; i += blockDim.x)
        if (sh[i]) atomicAdd(&hist[i], (unsigned long long)sh[i]);
}

__global__ void k_negate_i64(const long long *__restrict__ in, int64_t n, long long *__restrict__ out) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = -in[i];
}

double ord_to_double_host(unsigned long long o, bool is_float) {
    if (!is_float) return (double)(long long)(o ^ 0x8000000000000000ULL);
    if (o == ~0ULL) return std::numeric_limits<double>::quiet_NaN();
    const unsigned long long b = (o >> 63) ? (o & 0x7FFFFFFFFFFFFFFFULL) : ~o;
    double d;
    memcpy(&d, &b, 8);
    return d;
}

} // namespace

namespace {

// ---- approximate distinct count (Deequ's approximateNumDistinctValues is a HyperLogLog++ estimate) ------------------
// 2^12 registers (relative standard error 1.6 %, tighter than the 5 % Deequ asks of Spark's HLL++), values hashed by
// bit pattern like fsx_stats_distinct.  Registers are per-CTA in shared memory (native 32-bit atomicMax) and merged with
// a global atomicMax: the register file, and with it the estimate, does not depend on scheduling.  The estimate is
// Ertl's closed-form estimator (no empirical bias tables), computed on the host by fsx_hll_estimate -- which is also
// what merges the sketches of several ranks (element-wise max of the registers first).
constexpr int AHLL_P = 12, AHLL_M = 1 << AHLL_P;

template <typename BitsT>
__global__ void __launch_bounds__(256) k_value_hll(const BitsT *__restrict__ v, const uint8_t *__restrict__ valid, int64_t n, unsigned *regs) {
    __shared__ unsigned r[AHLL_M];
    for (int i = threadIdx.x; i < AHLL_M; i += blockDim.x) r[i] = 0;
    __syncthreads();
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (!dev_bit_valid(valid, i)) continue;
        const unsigned long long bits = (unsigned long long)v[i];                // zero-extended bit pattern
        const unsigned long long h = mix64(bits * 0x9E3779B97F4A7C15ULL + 0x632BE59BD9B4E019ULL);
        const unsigned idx = (unsigned)(h >> (64 - AHLL_P));
        const unsigned rho = (unsigned)__clzll((long long)((h << AHLL_P) | (1ULL << (AHLL_P - 1)))) + 1u;   // 1 .. 64 - p + 1
        if (rho > r[idx]) atomicMax(&r[idx], rho);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < AHLL_M; i += blockDim.x) if (r[i]) atomicMax(&regs[i], r[i]);
}

} // namespace

// Ertl, "New cardinality estimation algorithms for HyperLogLog sketches" (2017), improved raw estimator: valid over the
// whole range without bias tables.  regs: 2^p registers with values in [0, 64 - p + 1].  Host code (no GPU needed).
extern "C" double fsx_hll_estimate(const uint32_t *regs, int32_t p) {
    if (!regs || p < 4 || p > 18) return -1.0;
    const int m = 1 << p, q = 64 - p;
    std::vector<double> C((size_t)q + 2, 0.0);
    for (int i = 0; i < m; i++) C[regs[i] <= (uint32_t)(q + 1) ? regs[i] : (uint32_t)(q + 1)] += 1.0;
    auto sigma = [](double x) -> double {
        if (x == 1.0) return (double)INFINITY;
        double y = 1.0, z = x, prev;
        do { x *= x; prev = z; z += x * y; y += y; } while (z != prev);
        return z;
    };
    auto tau = [](double x) -> double {
        if (x == 0.0 || x == 1.0) return 0.0;
        double y = 1.0, z = 1.0 - x, prev;
        do { x = sqrt(x); prev = z; y *= 0.5; z -= (1.0 - x) * (1.0 - x) * y; } while (prev != z);
        return z / 3.0;
    };
    double z = (double)m * tau(1.0 - C[(size_t)q + 1] / (double)m);
    for (int k = q; k >= 1; k--) z = 0.5 * (z + C[(size_t)k]);
    z += (double)m * sigma(C[0] / (double)m);
    const double alpha_inf = 0.5 / log(2.0);
    return alpha_inf * (double)m * (double)m / z;
}

// the column's sketch into host memory (AHLL_M registers); inside an ABI call
int32_t fsx_internal_hll_registers(fsx_ctx *ctx, const fsx_col *col, uint32_t *regs_host) {
    unsigned *d_regs = nullptr;
    FSX_CHECK(fsx_dev_alloc(ctx, AHLL_M * 4, (void **)&d_regs));
    cudaError_t e = cudaMemsetAsync(d_regs, 0, AHLL_M * 4, ctx->stream);
    if (e == cudaSuccess && col->n > 0) {
        const int grid = fsx_grid_for(ctx, col->n, 256 * 8, 8);
        if (fsx_dtype_size(col->dtype) == 4) k_value_hll<unsigned><<<grid, 256, 0, ctx->stream>>>((const unsigned *)col->data, col->validity, col->n, d_regs);
        else k_value_hll<unsigned long long><<<grid, 256, 0, ctx->stream>>>((const unsigned long long *)col->data, col->validity, col->n, d_regs);
        ctx->launches++;
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(regs_host, d_regs, AHLL_M * 4, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    fsx_dev_free(ctx, d_regs);
    FSX_CUDA(ctx, e);
    return FSX_OK;
}

extern "C" int32_t fsx_stats_approx_distinct(fsx_ctx *ctx, const fsx_col *col, int64_t *estimate_out, uint32_t *registers_out) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, col && estimate_out, FSX_E_INVALID, "fsx_stats_approx_distinct: NULL argument");
    std::vector<uint32_t> regs(AHLL_M, 0u);
    FSX_CHECK(fsx_internal_hll_registers(ctx, col, regs.data()));
    if (registers_out) memcpy(registers_out, regs.data(), AHLL_M * 4);
    *estimate_out = (int64_t)llround(fsx_hll_estimate(regs.data(), AHLL_P));
    FSX_END();
}

// Dictionary encoding of a string column handed over as Arrow-style buffers (UTF-8 bytes + n + 1 offsets + optional
// validity bitmap): codes_out[i] = index of row i's string among the distinct strings in order of first appearance,
// -1 for a null; first_row_out[k] = a row that holds the string with code k.  Host code (no GPU): this is how string
// keys and categorical columns (team_name, team_owner, CF:39-44; `type` in the reference's group-by golden) reach
// fsx_groupby_agg / fsx_stats_value_counts, which work on integer codes.
extern "C" int32_t fsx_dictionary_encode_utf8(const uint8_t *bytes, const int64_t *offsets, const uint8_t *validity, int64_t n,
                                              int32_t *codes_out, int32_t *first_row_out, int32_t *ndistinct_out) {
    if (n < 0 || (n > 0 && (!offsets || !codes_out)) || !ndistinct_out) return fsx_fail(nullptr, FSX_E_INVALID, "fsx_dictionary_encode_utf8: bad argument");
    std::unordered_map<std::string_view, int32_t> dict;
    int32_t next = 0;
    for (int64_t i = 0; i < n; i++) {
        if (validity && !((validity[i >> 3] >> (i & 7)) & 1)) { codes_out[i] = -1; continue; }
        if (offsets[i + 1] < offsets[i]) return fsx_fail(nullptr, FSX_E_INVALID, "fsx_dictionary_encode_utf8: offsets decrease at row %lld", (long long)i);
        const std::string_view sv((const char *)bytes + offsets[i], (size_t)(offsets[i + 1] - offsets[i]));
        auto it = dict.find(sv);
        if (it == dict.end()) {
            if (next == INT32_MAX) return fsx_fail(nullptr, FSX_E_UNSUPPORTED, "fsx_dictionary_encode_utf8: more than 2^31-1 distinct strings");
            it = dict.emplace(sv, next).first;
            if (first_row_out) first_row_out[next] = (int32_t)i;
            next++;
        }
        codes_out[i] = it->second;
    }
    *ndistinct_out = next;
    return FSX_OK;
}

extern "C" int32_t fsx_stats_distinct(fsx_ctx *ctx, const fsx_col *col, fsx_distinct_stats *out) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, col && out, FSX_E_INVALID, "fsx_stats_distinct: NULL argument");
    memset(out, 0, sizeof *out);
    out->n_rows = col->n;
    if (col->n == 0) { FSX_END(); }
    // value frequencies: K1 over the bit patterns (f32 -> i32 view, f64 -> i64 view), count(*)
    fsx_col view = *col;
    view.owns = false;
    view.dtype = fsx_dtype_size(col->dtype) == 4 ? FSX_I32 : FSX_I64;
    const int32_t agg = FSX_AGG_COUNT_STAR;
    const fsx_col *none = nullptr;
    fsx_table *t = nullptr;
    FSX_CHECK(fsx_internal_groupby(ctx, &view, &none, &agg, 1, 0, &t));
    const fsx_col *keys = t->cols[0], *counts = t->cols[1];
    int64_t nulls = 0;
    void *scratch = nullptr;
    int32_t rc = fsx_dev_alloc(ctx, RED_BLOCKS * 24, &scratch);
    if (rc != FSX_OK) { fsx_table_free(ctx, t); return rc; }
    double *pe = (double *)scratch;
    unsigned long long *pd = (unsigned long long *)(pe + RED_BLOCKS), *pu = pd + RED_BLOCKS;
    // non-null rows = all rows - rows of the null group: a first tiny pass is not needed, the null group's count
    // is found on the host from the (key validity, count) of the at most one invalid key row
    std::vector<double> he(RED_BLOCKS);
    std::vector<unsigned long long> hd(RED_BLOCKS), hu(RED_BLOCKS);
    cudaError_t e = cudaSuccess;
    if (col->validity) {
        // null count of the input column decides N before the reduction
        rc = fsx_internal_null_count(ctx, col, &nulls);
        if (rc != FSX_OK) { fsx_dev_free(ctx, scratch); fsx_table_free(ctx, t); return rc; }
    }
    out->n_nonnull = col->n - nulls;
    if (out->n_nonnull > 0) {
        k_freq_reduce<<<RED_BLOCKS, RED_THREADS, 0, ctx->stream>>>((const long long *)counts->data, keys->validity, t->nrows,
                                                                  1.0 / (double)out->n_nonnull, pe, pd, pu);
        ctx->launches++;
        e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaMemcpyAsync(he.data(), pe, RED_BLOCKS * 8, cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(hd.data(), pd, RED_BLOCKS * 8, cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(hu.data(), pu, RED_BLOCKS * 8, cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    }
    fsx_dev_free(ctx, scratch);
    fsx_table_free(ctx, t);
    FSX_CUDA(ctx, e);
    double h = 0.0;
    for (int b = 0; b < RED_BLOCKS; b++) { h += he[b]; out->n_distinct += (int64_t)hd[b]; out->n_unique += (int64_t)hu[b]; }
    out->entropy = out->n_nonnull > 0 ? h : 0.0;
    FSX_END();
}

// One radix pass over one column: hist_host[q * 256 + b] = how many values share probability q's decided prefix and have
// byte `shift / 8` equal to b.  Inside an ABI call; synchronises.
int32_t fsx_internal_quantile_pass(fsx_ctx *ctx, const fsx_col *col, int shift, const unsigned long long *prefix, int nq, unsigned long long *hist_host) {
    QState st;
    memset(&st, 0, sizeof st);
    st.nq = nq;
    for (int q = 0; q < nq; q++) st.prefix[q] = prefix[q];
    unsigned long long *hist = nullptr;
    FSX_CHECK(fsx_dev_alloc(ctx, (size_t)Q_MAX * 256 * 8, (void **)&hist));
    const int grid = fsx_grid_for(ctx, col->n, 256 * 8, 8);
    cudaError_t e = cudaMemsetAsync(hist, 0, (size_t)nq * 256 * 8, ctx->stream);
    if (e == cudaSuccess && col->n > 0) {
        switch (col->dtype) {
        case FSX_I32: k_radix_hist<int32_t><<<grid, 256, 0, ctx->stream>>>((const int32_t *)col->data, col->validity, col->n, shift, st, hist); break;
        case FSX_I64: k_radix_hist<long long><<<grid, 256, 0, ctx->stream>>>((const long long *)col->data, col->validity, col->n, shift, st, hist); break;
        case FSX_F32: k_radix_hist<float><<<grid, 256, 0, ctx->stream>>>((const float *)col->data, col->validity, col->n, shift, st, hist); break;
        default: k_radix_hist<double><<<grid, 256, 0, ctx->stream>>>((const double *)col->data, col->validity, col->n, shift, st, hist); break;
        }
        ctx->launches++;
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(hist_host, hist, (size_t)nq * 256 * 8, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    fsx_dev_free(ctx, hist);
    FSX_CUDA(ctx, e);
    return FSX_OK;
}

// 1-based target ranks among nn sorted values: ceil(p * n), at least 1 (Spark approxQuantile with relativeError = 0:
// floor(p n) <= rank <= ceil(p n))
void fsx_quantile_ranks(const double *probs, int nq, int64_t nn, unsigned long long *rank) {
    for (int q = 0; q < nq; q++) {
        double r = ceil(probs[q] * (double)nn);
        if (r < 1.0) r = 1.0;
        if (r > (double)nn) r = (double)nn;
        rank[q] = (unsigned long long)r;
    }
}

// walk the (merged) bins of one pass until each rank falls inside one: fixes byte `shift / 8` of every prefix
void fsx_quantile_step(const unsigned long long *hist, int nq, int shift, unsigned long long *rank, unsigned long long *prefix) {
    for (int q = 0; q < nq; q++) {
        unsigned long long acc = 0;
        int b = 0;
        for (; b < 255; b++) {
            if (acc + hist[(size_t)q * 256 + b] >= rank[q]) break;
            acc += hist[(size_t)q * 256 + b];
        }
        rank[q] -= acc;
        prefix[q] |= (unsigned long long)b << shift;
    }
}

double fsx_quantile_value(unsigned long long ord, bool is_float) { return ord_to_double_host(ord, is_float); }

extern "C" int32_t fsx_stats_quantiles(fsx_ctx *ctx, const fsx_col *col, const double *probs, int32_t nprobs, double *out,
                                       int64_t *n_nonnull_out) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, col && (nprobs == 0 || (probs && out)) && nprobs >= 0 && nprobs <= Q_MAX, FSX_E_INVALID,
                "fsx_stats_quantiles: bad argument (at most %d probabilities per call)", (int)Q_MAX);
    for (int q = 0; q < nprobs; q++)
        FSX_REQUIRE(ctx, probs[q] >= 0.0 && probs[q] <= 1.0, FSX_E_INVALID, "fsx_stats_quantiles: probability %d outside [0,1]", q);
    int64_t nulls = 0;
    if (col->validity && col->n > 0) FSX_CHECK(fsx_internal_null_count(ctx, col, &nulls));
    const int64_t nn = col->n - nulls;
    if (n_nonnull_out) *n_nonnull_out = nn;
    if (nprobs == 0) { FSX_END(); }
    if (nn == 0) {
        for (int q = 0; q < nprobs; q++) out[q] = std::numeric_limits<double>::quiet_NaN();
        FSX_END();
    }
    unsigned long long rank[Q_MAX], prefix[Q_MAX] = {0};
    fsx_quantile_ranks(probs, nprobs, nn, rank);
    std::vector<unsigned long long> h((size_t)nprobs * 256);
    for (int shift = 56; shift >= 0; shift -= 8) {
        FSX_CHECK(fsx_internal_quantile_pass(ctx, col, shift, prefix, nprobs, h.data()));
        fsx_quantile_step(h.data(), nprobs, shift, rank, prefix);
    }
    for (int q = 0; q < nprobs; q++) out[q] = ord_to_double_host(prefix[q], fsx_dtype_is_float(col->dtype));
    FSX_END();
}

// Exact value counts (Deequ's histogram of a low-cardinality column: value -> count): K1 on the bit patterns, then the
// device sort, most frequent value first (ties by ascending value, the null group last among equals).
extern "C" int32_t fsx_stats_value_counts(fsx_ctx *ctx, const fsx_col *col, fsx_table **out) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, col && out, FSX_E_INVALID, "fsx_stats_value_counts: NULL argument");
    *out = nullptr;
    fsx_col view = *col;
    view.owns = false;
    view.dtype = fsx_dtype_size(col->dtype) == 4 ? FSX_I32 : FSX_I64;
    const int32_t agg = FSX_AGG_COUNT_STAR;
    const fsx_col *none = nullptr;
    fsx_table *t = nullptr;
    FSX_CHECK(fsx_internal_groupby(ctx, &view, &none, &agg, 1, 0, &t));
    t->cols[0]->dtype = col->dtype;                               // the keys are the column's own bit patterns
    const int64_t g = t->nrows;
    if (g > 1) {
        // ascending sort by (-count, value): negate the counts into a scratch column
        fsx_col *neg = nullptr;
        int32_t rc = fsx_col_new(ctx, FSX_I64, g, false, &neg);
        if (rc != FSX_OK) { fsx_table_free(ctx, t); return rc; }
        k_negate_i64<<<fsx_grid_for(ctx, g, 256, 8), 256, 0, ctx->stream>>>((const long long *)t->cols[1]->data, g, (long long *)neg->data);
        ctx->launches++;
        const fsx_col *by[2] = {neg, t->cols[0]};
        unsigned *perm = nullptr;
        rc = fsx_internal_argsort(ctx, by, 2, g, &perm);
        fsx_col_release(ctx, neg);
        if (rc != FSX_OK) { fsx_table_free(ctx, t); return rc; }
        for (size_t c = 0; c < t->cols.size(); c++) {
            fsx_col *dst = nullptr;
            rc = fsx_internal_gather_u32(ctx, t->cols[c], perm, g, &dst);
            if (rc != FSX_OK) { fsx_dev_free(ctx, perm); fsx_table_free(ctx, t); return rc; }
            fsx_col_release(ctx, t->cols[c]);
            t->cols[c] = dst;
        }
        fsx_dev_free(ctx, perm);
    }
    *out = t;
    FSX_END();
}
