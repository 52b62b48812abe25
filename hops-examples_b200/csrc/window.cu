// window.cu -- the aggregation shapes of the hsfs tour's own feature engineering (SURVEY.md §8(f) rows 2 and 4;
// notebooks/featurestore/hsfs/basics/feature_engineering.ipynb):
//
//   fsx_column_combine_keys     composite keys: (store, dept), (store, date) -> one dense int64 key usable by K1, K3
//                               and the sorts (:229 `partitionBy(['store','dept'])`, feature_exploration.ipynb:391 joins)
//   fsx_groupby_count_distinct  `groupBy("store").agg(countDistinct("dept"))` (:152-159)
//   fsx_window_range_sum        `F.sum(v).over(Window.partitionBy(p).orderBy(ts).rangeBetween(lo, hi))` (:229-249)
//
// The window sum sorts (partition, order) once with the radix sort of csrc/sort.cu, takes one inclusive prefix sum
// of the values in double-double arithmetic (so that prefix[end] - prefix[start] keeps full f64 accuracy however
// long the column is) and finds each row's frame by a galloping search outwards from the row itself: frames are
// short compared with the column, so the search costs O(log frame), not O(log n), random reads.
#include <vector>

#include "scan.cuh"
#include "sort.cuh"

int32_t fsx_internal_groupby(fsx_ctx *ctx, const fsx_col *key, const fsx_col *const *vals, const int32_t *aggs, int32_t nvals,
                             int64_t expected_groups, fsx_table **out);

namespace {

constexpr unsigned long long W_MSB = 0x8000000000000000ULL;

template <typename T>
__global__ void __launch_bounds__(256) k_w_minmax(const T *__restrict__ v, const uint8_t *__restrict__ valid, int64_t n, unsigned long long *scalars) {
    unsigned long long lo = ~0ULL, hi = 0ULL;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (!dev_bit_valid(valid, i)) continue;
        const unsigned long long o = ord_from_i64((long long)v[i]);
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

// (min, span = max - min) of the valid values of an integer column; span = -1 when there is none
int32_t int_range(fsx_ctx *ctx, const fsx_col *col, long long *mn, long long *span) {
    const unsigned long long init[8] = {0, 0, ~0ULL, 0, 0, 0, 0, 0};
    cudaError_t e = cudaMemcpyAsync(ctx->d_scalars, init, sizeof init, cudaMemcpyHostToDevice, ctx->stream);
    if (e != cudaSuccess) return fsx_fail(ctx, FSX_E_CUDA, "int_range: %s", cudaGetErrorString(e));
    if (col->n > 0) {
        const int grid = fsx_grid_for(ctx, col->n, 256 * 8, 8);
        if (col->dtype == FSX_I32) k_w_minmax<int32_t><<<grid, 256, 0, ctx->stream>>>((const int32_t *)col->data, col->validity, col->n, ctx->d_scalars);
        else k_w_minmax<long long><<<grid, 256, 0, ctx->stream>>>((const long long *)col->data, col->validity, col->n, ctx->d_scalars);
        ctx->launches++;
    }
    int32_t rc = fsx_read_scalars(ctx, 0, 8);
    if (rc != FSX_OK) return rc;
    if (ctx->h_scalars[2] > ctx->h_scalars[3]) { *mn = 0; *span = -1; return FSX_OK; }
    *mn = (long long)(ctx->h_scalars[2] ^ W_MSB);
    const unsigned long long s = ctx->h_scalars[3] - ctx->h_scalars[2];
    *span = s > (unsigned long long)INT64_MAX ? INT64_MAX : (long long)s;
    return FSX_OK;
}

__device__ __forceinline__ long long load_int(const void *p, int dtype, int64_t i) {
    return dtype == FSX_I32 ? (long long)((const int32_t *)p)[i] : ((const long long *)p)[i];
}

// out = (a - amin) * bcodes + (b - bmin); a null code (aspan + 1 resp. bspan + 1) is used when null_codes is set,
// otherwise a null in either input makes the output null
__global__ void __launch_bounds__(256) k_combine(const void *a, int adt, const uint8_t *av, long long amin, long long anull,
                                                 const void *b, int bdt, const uint8_t *bv, long long bmin, long long bnull, long long bcodes,
                                                 int64_t n, bool null_codes, long long *out, unsigned *out_valid_words) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const bool va = dev_bit_valid(av, i), vb = dev_bit_valid(bv, i);
        const long long ca = va ? load_int(a, adt, i) - amin : anull, cb = vb ? load_int(b, bdt, i) - bmin : bnull;
        out[i] = ca * bcodes + cb;
        if (!null_codes && out_valid_words && va && vb) atomicOr(&out_valid_words[i >> 5], 1u << (i & 31));
    }
}

// packed distinct (key, value) codes -> key (original dtype, null when the code is the null code) and a 0/1 flag
template <typename KT>
__global__ void __launch_bounds__(256) k_unpack_pairs(const long long *__restrict__ packed, int64_t n, long long kmin, long long knull, long long vnull,
                                                      long long vcodes, KT *key_out, unsigned *key_valid_words, long long *flag_out) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const long long p = packed[i], kc = p / vcodes, vc = p - kc * vcodes;
        key_out[i] = kc == knull ? (KT)0 : (KT)(kc + kmin);
        if (kc != knull) atomicOr(&key_valid_words[i >> 5], 1u << (i & 31));
        flag_out[i] = vc != vnull;
    }
}

// ---- double-double prefix sums ---------------------------------------------------------------------------------
struct dd { double hi, lo; };
__device__ __forceinline__ dd dd_add(dd a, dd b) {
    const double s = a.hi + b.hi, bb = s - a.hi;
    double e = (a.hi - (s - bb)) + (b.hi - bb);          // two_sum error term
    e += a.lo + b.lo;
    const double hi = s + e;
    return dd{hi, e - (hi - s)};
}
struct WAcc {                                            // what one prefix entry carries
    dd sum;                                              // finite non-null values
    unsigned cnt, nan, pinf, ninf;                       // non-null values, NaNs, +Inf, -Inf
};
__device__ __forceinline__ WAcc wacc_add(const WAcc &a, const WAcc &b) {
    return WAcc{dd_add(a.sum, b.sum), a.cnt + b.cnt, a.nan + b.nan, a.pinf + b.pinf, a.ninf + b.ninf};
}
__device__ __forceinline__ WAcc wacc_zero() { return WAcc{dd{0.0, 0.0}, 0u, 0u, 0u, 0u}; }
__device__ __forceinline__ WAcc wacc_shfl_up(const WAcc &a, int o) {
    WAcc r;
    r.sum.hi = __shfl_up_sync(0xffffffffu, a.sum.hi, o); r.sum.lo = __shfl_up_sync(0xffffffffu, a.sum.lo, o);
    r.cnt = __shfl_up_sync(0xffffffffu, a.cnt, o); r.nan = __shfl_up_sync(0xffffffffu, a.nan, o);
    r.pinf = __shfl_up_sync(0xffffffffu, a.pinf, o); r.ninf = __shfl_up_sync(0xffffffffu, a.ninf, o);
    return r;
}

constexpr int WS_THREADS = 256, WS_ITEMS = 8, WS_TILE = WS_THREADS * WS_ITEMS;

template <typename VT>
__device__ __forceinline__ WAcc wacc_of(const VT *vals, const uint8_t *valid, unsigned row) {
    WAcc a = wacc_zero();
    if (!dev_bit_valid(valid, row)) return a;
    const double x = (double)vals[row];
    a.cnt = 1;
    if (x != x) a.nan = 1;
    else if (x > 1.7976931348623157e308) a.pinf = 1;
    else if (x < -1.7976931348623157e308) a.ninf = 1;
    else a.sum.hi = x;
    return a;
}

// inclusive scan of one tile in sorted order; MODE 0: write the tile total, MODE 1: write prefix[i] = carry[tile] + scan
template <typename VT, int MODE>
__global__ void __launch_bounds__(WS_THREADS) k_w_scan(const VT *__restrict__ vals, const uint8_t *__restrict__ valid, const unsigned *__restrict__ perm,
                                                       int64_t n, WAcc *tile_acc, WAcc *prefix) {
    __shared__ WAcc wtot[WS_THREADS / 32];
    const int64_t base = (int64_t)blockIdx.x * WS_TILE + (int64_t)threadIdx.x * WS_ITEMS;
    WAcc item[WS_ITEMS], run = wacc_zero();
#pragma unroll
    for (int j = 0; j < WS_ITEMS; j++) {
        item[j] = base + j < n ? wacc_of<VT>(vals, valid, perm[base + j]) : wacc_zero();
        run = wacc_add(run, item[j]);
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    WAcc incl = run;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const WAcc y = wacc_shfl_up(incl, o);
        if (lane >= o) incl = wacc_add(y, incl);
    }
    if (lane == 31) wtot[warp] = incl;
    __syncthreads();
    WAcc before = wacc_zero();
    for (int w = 0; w < warp; w++) before = wacc_add(before, wtot[w]);
    if (MODE == 0) {
        if (threadIdx.x == WS_THREADS - 1) tile_acc[blockIdx.x] = wacc_add(before, incl);
        return;
    }
    // exclusive prefix of this thread = carry + warps before + lanes before
    WAcc excl = wacc_add(tile_acc[blockIdx.x], before);
    const WAcc up = wacc_shfl_up(incl, 1);
    if (lane > 0) excl = wacc_add(excl, up);
#pragma unroll
    for (int j = 0; j < WS_ITEMS; j++) {
        excl = wacc_add(excl, item[j]);
        if (base + j < n) prefix[base + j] = excl;
    }
}

// tile totals -> exclusive carries: one warp, 32 tiles per step
__global__ void k_w_carry(WAcc *tile_acc, int64_t ntiles) {
    const int lane = threadIdx.x;
    WAcc carry = wacc_zero();
    for (int64_t base = 0; base < ntiles; base += 32) {
        const WAcc v = base + lane < ntiles ? tile_acc[base + lane] : wacc_zero();
        WAcc incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const WAcc y = wacc_shfl_up(incl, o);
            if (lane >= o) incl = wacc_add(y, incl);
        }
        const WAcc up = wacc_shfl_up(incl, 1);
        if (base + lane < ntiles) tile_acc[base + lane] = lane ? wacc_add(carry, up) : carry;
        WAcc tot;
        tot.sum.hi = __shfl_sync(0xffffffffu, incl.sum.hi, 31); tot.sum.lo = __shfl_sync(0xffffffffu, incl.sum.lo, 31);
        tot.cnt = __shfl_sync(0xffffffffu, incl.cnt, 31); tot.nan = __shfl_sync(0xffffffffu, incl.nan, 31);
        tot.pinf = __shfl_sync(0xffffffffu, incl.pinf, 31); tot.ninf = __shfl_sync(0xffffffffu, incl.ninf, 31);
        carry = wacc_add(carry, tot);
    }
}

// sorted copies of the partition and order keys (as int64; no partition column: all zero)
__global__ void __launch_bounds__(256) k_w_sorted_keys(const void *part, int pdt, const uint8_t *pvalid, const void *ord, int odt, const unsigned *__restrict__ perm,
                                                       int64_t n, long long *sp, unsigned char *sp_null, long long *so) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const unsigned r = perm[i];
        const bool pv = part == nullptr || dev_bit_valid(pvalid, r);
        sp[i] = part && pv ? load_int(part, pdt, r) : 0;
        sp_null[i] = pv ? 0 : 1;                                       // the null partition is one partition (Spark groups nulls)
        so[i] = load_int(ord, odt, r);
    }
}

__device__ __forceinline__ long long sat_add(long long a, long long b) {
    if (b > 0 && a > INT64_MAX - b) return INT64_MAX;
    if (b < 0 && a < INT64_MIN - b) return INT64_MIN;
    return a + b;
}

// frame of sorted row i: rows j of the same partition with so[j] in [so[i] + lo, so[i] + hi]; result scattered to the original row
__global__ void __launch_bounds__(256) k_w_frames(const long long *__restrict__ sp, const unsigned char *__restrict__ sp_null, const long long *__restrict__ so,
                                                  const unsigned *__restrict__ perm, const WAcc *__restrict__ prefix, int64_t n, long long lo, long long hi,
                                                  double *out, unsigned *out_valid_words) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const long long p = sp[i];
        const unsigned char pn = sp_null[i];
        const long long t_lo = sat_add(so[i], lo), t_hi = sat_add(so[i], hi);
        auto same = [&](int64_t j) { return sp[j] == p && sp_null[j] == pn; };
        // first = smallest j in this partition with so[j] >= t_lo; last1 = one past the largest j with so[j] <= t_hi.
        // Both are found by galloping away from i (the row is inside its own partition's run), then bisecting.
        auto below = [&](int64_t j) { return !same(j) ? (j < i) : so[j] < t_lo; };     // "j lies before the frame start"
        auto not_after = [&](int64_t j) { return !same(j) ? (j < i) : so[j] <= t_hi; }; // "j lies before or inside the frame end"
        int64_t first, last1;
        {   // lower bound of the predicate !below over [0, n)
            int64_t a, b;                                                               // below(a) true (or a = -1), below(b) false (or b = n)
            if (below(i)) { a = i; int64_t s = 1; b = i + s; while (b < n && below(b)) { a = b; s <<= 1; b = i + s; } if (b > n) b = n; }
            else { b = i; int64_t s = 1; a = i - s; while (a >= 0 && !below(a)) { b = a; s <<= 1; a = i - s; } if (a < -1) a = -1; }
            while (b - a > 1) { const int64_t m = a + (b - a) / 2; if (below(m)) a = m; else b = m; }
            first = b;
        }
        {
            int64_t a, b;                                                               // not_after(a) true (or a = -1), not_after(b) false (or b = n)
            if (not_after(i)) { a = i; int64_t s = 1; b = i + s; while (b < n && not_after(b)) { a = b; s <<= 1; b = i + s; } if (b > n) b = n; }
            else { b = i; int64_t s = 1; a = i - s; while (a >= 0 && !not_after(a)) { b = a; s <<= 1; a = i - s; } if (a < -1) a = -1; }
            while (b - a > 1) { const int64_t m = a + (b - a) / 2; if (not_after(m)) a = m; else b = m; }
            last1 = b;
        }
        const unsigned row = perm[i];
        double res = 0.0;
        bool valid = false;
        if (last1 > first) {
            const WAcc e = prefix[last1 - 1];
            WAcc s = wacc_zero();
            if (first > 0) s = prefix[first - 1];
            const unsigned cnt = e.cnt - s.cnt, nan = e.nan - s.nan, pinf = e.pinf - s.pinf, ninf = e.ninf - s.ninf;
            valid = cnt > 0;
            if (nan || (pinf && ninf)) res = __longlong_as_double(0x7FF8000000000000LL);
            else if (pinf) res = __longlong_as_double(0x7FF0000000000000LL);
            else if (ninf) res = __longlong_as_double((long long)0xFFF0000000000000ULL);
            else { const dd d = dd_add(e.sum, dd{-s.sum.hi, -s.sum.lo}); res = d.hi + d.lo; }
        }
        out[row] = valid ? res : 0.0;
        if (valid) atomicOr(&out_valid_words[row >> 5], 1u << (row & 31));
    }
}

// ---- combined-key route: (partition code << tsbits) | (order - min order) fits 64 bits -------------------------------
// One sort of one key, and the sorted key array itself is what the frame search runs on: no gathered copies of the
// partition / order columns, one array to bisect instead of two.
__global__ void __launch_bounds__(256) k_w_make_ckeys(const void *part, int pdt, const uint8_t *pvalid, long long pmin, unsigned long long pnull,
                                                      const void *ord, int odt, long long omin, int tsbits, int64_t n,
                                                      unsigned long long *ckeys, unsigned *idx) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        unsigned long long pc = 0;
        if (part) pc = dev_bit_valid(pvalid, i) ? (unsigned long long)(load_int(part, pdt, i) - pmin) : pnull;
        const unsigned long long tc = (unsigned long long)(load_int(ord, odt, i) - omin);
        ckeys[i] = (tsbits < 64 ? pc << tsbits : 0ULL) | tc;
        idx[i] = (unsigned)i;
    }
}

__global__ void __launch_bounds__(256) k_w_frames_ckey(const unsigned long long *__restrict__ ck, const unsigned *__restrict__ perm,
                                                       const WAcc *__restrict__ prefix, int64_t n, int tsbits, unsigned long long ospan,
                                                       long long lo, long long hi, double *out, unsigned *out_valid_words) {
    const unsigned long long tmask = tsbits >= 64 ? ~0ULL : ((1ULL << tsbits) - 1ULL);
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const unsigned long long key = ck[i], pbase = key & ~tmask, t = key & tmask;
        // frame in order codes, clamped to the codes that exist [0, ospan]; empty when it lies outside them
        bool empty = false;
        unsigned long long a_code, b_code;
        if (lo >= 0) { empty |= (unsigned long long)lo > ospan - t; a_code = t + (unsigned long long)lo; }
        else { const unsigned long long back = (unsigned long long)(-(lo + 1)) + 1ULL; a_code = back > t ? 0ULL : t - back; }
        if (hi >= 0) { const unsigned long long fwd = (unsigned long long)hi; b_code = fwd > ospan - t ? ospan : t + fwd; }
        else { const unsigned long long back = (unsigned long long)(-(hi + 1)) + 1ULL; empty |= back > t; b_code = t - back; }
        const unsigned row = perm[i];
        double res = 0.0;
        bool valid = false;
        if (!empty) {
            const unsigned long long ka = pbase | a_code, kb = pbase | b_code;
            int64_t first, last1;
            {   // first j with ck[j] >= ka, galloping away from i
                int64_t a, b;
                if (ck[i] < ka) { a = i; int64_t s = 1; b = i + s; while (b < n && ck[b] < ka) { a = b; s <<= 1; b = i + s; } if (b > n) b = n; }
                else { b = i; int64_t s = 1; a = i - s; while (a >= 0 && ck[a] >= ka) { b = a; s <<= 1; a = i - s; } if (a < -1) a = -1; }
                while (b - a > 1) { const int64_t m = a + (b - a) / 2; if (ck[m] < ka) a = m; else b = m; }
                first = b;
            }
            {   // one past the last j with ck[j] <= kb
                int64_t a, b;
                if (ck[i] <= kb) { a = i; int64_t s = 1; b = i + s; while (b < n && ck[b] <= kb) { a = b; s <<= 1; b = i + s; } if (b > n) b = n; }
                else { b = i; int64_t s = 1; a = i - s; while (a >= 0 && ck[a] > kb) { b = a; s <<= 1; a = i - s; } if (a < -1) a = -1; }
                while (b - a > 1) { const int64_t m = a + (b - a) / 2; if (ck[m] <= kb) a = m; else b = m; }
                last1 = b;
            }
            if (last1 > first) {
                const WAcc e = prefix[last1 - 1];
                WAcc s0 = wacc_zero();
                if (first > 0) s0 = prefix[first - 1];
                const unsigned cnt = e.cnt - s0.cnt, nan = e.nan - s0.nan, pinf = e.pinf - s0.pinf, ninf = e.ninf - s0.ninf;
                valid = cnt > 0;
                if (nan || (pinf && ninf)) res = __longlong_as_double(0x7FF8000000000000LL);
                else if (pinf) res = __longlong_as_double(0x7FF0000000000000LL);
                else if (ninf) res = __longlong_as_double((long long)0xFFF0000000000000ULL);
                else { const dd d = dd_add(e.sum, dd{-s0.sum.hi, -s0.sum.lo}); res = d.hi + d.lo; }
            }
        }
        out[row] = valid ? res : 0.0;
        if (valid) atomicOr(&out_valid_words[row >> 5], 1u << (row & 31));
    }
}

int bits_of(unsigned long long v) { int b = 0; while (b < 64 && (v >> b) != 0ULL) b++; return b; }

bool is_int(const fsx_col *c) { return c->dtype == FSX_I32 || c->dtype == FSX_I64; }

} // namespace

extern "C" int32_t fsx_column_combine_keys(fsx_ctx *ctx, const fsx_col *a, const fsx_col *b, fsx_col **out) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, a && b && out, FSX_E_INVALID, "fsx_column_combine_keys: NULL argument");
    *out = nullptr;
    FSX_REQUIRE(ctx, is_int(a) && is_int(b), FSX_E_UNSUPPORTED, "fsx_column_combine_keys: integer key columns only");
    FSX_REQUIRE(ctx, a->n == b->n, FSX_E_INVALID, "fsx_column_combine_keys: column lengths differ");
    long long amin, aspan, bmin, bspan;
    FSX_CHECK(int_range(ctx, a, &amin, &aspan));
    FSX_CHECK(int_range(ctx, b, &bmin, &bspan));
    if (aspan < 0) aspan = 0;
    if (bspan < 0) bspan = 0;
    FSX_REQUIRE(ctx, bspan < INT64_MAX && (unsigned long long)aspan < (unsigned long long)INT64_MAX / (unsigned long long)(bspan + 1) - 1ULL, FSX_E_UNSUPPORTED,
                "fsx_column_combine_keys: the two key ranges (%lld and %lld values) do not fit one int64 key", aspan + 1, bspan + 1);
    const bool nullable = a->validity || b->validity;
    fsx_col *o = nullptr;
    FSX_CHECK(fsx_col_new(ctx, FSX_I64, a->n, nullable, &o));
    if (o->validity) FSX_ASYNC(ctx, cudaMemsetAsync(o->validity, 0, (size_t)((a->n + 127) / 128) * 16, ctx->stream));
    if (a->n > 0) {
        k_combine<<<fsx_grid_for(ctx, a->n, 256 * 4, 8), 256, 0, ctx->stream>>>(a->data, a->dtype, a->validity, amin, 0, b->data, b->dtype, b->validity, bmin, 0,
                                                                                bspan + 1, a->n, false, (long long *)o->data, (unsigned *)o->validity);
        ctx->launches++;
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { fsx_col_release(ctx, o); FSX_CUDA(ctx, e); }
    *out = o;
    FSX_END();
}

extern "C" int32_t fsx_groupby_count_distinct(fsx_ctx *ctx, const fsx_col *key, const fsx_col *val, fsx_table **out) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, key && val && out, FSX_E_INVALID, "fsx_groupby_count_distinct: NULL argument");
    *out = nullptr;
    FSX_REQUIRE(ctx, is_int(key) && is_int(val), FSX_E_UNSUPPORTED, "fsx_groupby_count_distinct: integer key and value columns only");
    FSX_REQUIRE(ctx, key->n == val->n, FSX_E_INVALID, "fsx_groupby_count_distinct: column lengths differ");
    const int64_t n = key->n;
    long long kmin, kspan, vmin, vspan;
    FSX_CHECK(int_range(ctx, key, &kmin, &kspan));
    FSX_CHECK(int_range(ctx, val, &vmin, &vspan));
    if (kspan < 0) kspan = 0;
    if (vspan < 0) vspan = 0;
    // codes 0..span are values, span + 1 is the null code
    FSX_REQUIRE(ctx, vspan < INT64_MAX - 2 && (unsigned long long)kspan + 2ULL < (unsigned long long)INT64_MAX / (unsigned long long)(vspan + 2), FSX_E_UNSUPPORTED,
                "fsx_groupby_count_distinct: the key and value ranges (%lld x %lld) do not fit one int64 pair code", kspan + 1, vspan + 1);
    const long long knull = kspan + 1, vnull = vspan + 1, vcodes = vspan + 2;
    fsx_col *packed = nullptr;
    FSX_CHECK(fsx_col_new(ctx, FSX_I64, n, false, &packed));
    if (n > 0) {
        k_combine<<<fsx_grid_for(ctx, n, 256 * 4, 8), 256, 0, ctx->stream>>>(key->data, key->dtype, key->validity, kmin, knull, val->data, val->dtype, val->validity, vmin, vnull,
                                                                             vcodes, n, true, (long long *)packed->data, nullptr);
        ctx->launches++;
    }
    // 1. the distinct (key, value) pairs
    fsx_table *pairs = nullptr;
    int32_t rc = fsx_internal_groupby(ctx, packed, nullptr, nullptr, 0, 0, &pairs);
    fsx_col_release(ctx, packed);
    if (rc != FSX_OK) return rc;
    const int64_t np = pairs->nrows;
    // 2. back to (key, 1 if the value is not null), then count per key
    fsx_col *k2 = nullptr, *flag = nullptr;
    rc = fsx_col_new(ctx, key->dtype, np, true, &k2);
    if (rc == FSX_OK) rc = fsx_col_new(ctx, FSX_I64, np, false, &flag);
    if (rc != FSX_OK) { fsx_col_release(ctx, k2); fsx_table_free(ctx, pairs); return rc; }
    FSX_ASYNC(ctx, cudaMemsetAsync(k2->validity, 0, (size_t)((np + 127) / 128) * 16, ctx->stream));
    if (np > 0) {
        const int grid = fsx_grid_for(ctx, np, 256 * 4, 8);
        const long long *pk = (const long long *)pairs->cols[0]->data;
        if (key->dtype == FSX_I32) k_unpack_pairs<int32_t><<<grid, 256, 0, ctx->stream>>>(pk, np, kmin, knull, vnull, vcodes, (int32_t *)k2->data, (unsigned *)k2->validity, (long long *)flag->data);
        else k_unpack_pairs<long long><<<grid, 256, 0, ctx->stream>>>(pk, np, kmin, knull, vnull, vcodes, (long long *)k2->data, (unsigned *)k2->validity, (long long *)flag->data);
        ctx->launches++;
    }
    const int32_t agg = FSX_AGG_SUM;
    const fsx_col *fv = flag;
    fsx_table *res = nullptr;
    rc = fsx_internal_groupby(ctx, k2, &fv, &agg, 1, 0, &res);
    fsx_col_release(ctx, k2);
    fsx_col_release(ctx, flag);
    fsx_table_free(ctx, pairs);
    if (rc != FSX_OK) return rc;
    *out = res;
    FSX_END();
}

extern "C" int32_t fsx_window_range_sum(fsx_ctx *ctx, const fsx_col *partition, const fsx_col *order, const fsx_col *val, int64_t lo, int64_t hi, fsx_col **out) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, order && val && out, FSX_E_INVALID, "fsx_window_range_sum: NULL argument");
    *out = nullptr;
    FSX_REQUIRE(ctx, is_int(order) && !order->validity, FSX_E_UNSUPPORTED, "fsx_window_range_sum: the order column must be a null-free integer column");
    FSX_REQUIRE(ctx, !partition || is_int(partition), FSX_E_UNSUPPORTED, "fsx_window_range_sum: integer partition column only (combine several with fsx_column_combine_keys)");
    FSX_REQUIRE(ctx, lo <= hi, FSX_E_INVALID, "fsx_window_range_sum: empty frame (lo > hi)");
    const int64_t n = order->n;
    FSX_REQUIRE(ctx, val->n == n && (!partition || partition->n == n), FSX_E_INVALID, "fsx_window_range_sum: column lengths differ");
    fsx_col *o = nullptr;
    FSX_CHECK(fsx_col_new(ctx, FSX_F64, n, true, &o));
    FSX_ASYNC(ctx, cudaMemsetAsync(o->validity, 0, (size_t)((n + 127) / 128) * 16, ctx->stream));
    if (n == 0) { *out = o; FSX_END(); }
    // ---- combined-key route when (partition code, order code) fits one 64-bit key
    {
        long long pmin = 0, omin = 0, dummy = 0;
        unsigned long long pmax_code = 0, pnull = 0, ospan = 0;
        int pbits = 0;
        if (partition) {
            int32_t rc0 = int_range(ctx, partition, &pmin, &dummy);
            if (rc0 != FSX_OK) { fsx_col_release(ctx, o); return rc0; }
            const bool any = ctx->h_scalars[2] <= ctx->h_scalars[3];
            const unsigned long long pspan = any ? ctx->h_scalars[3] - ctx->h_scalars[2] : 0ULL;
            pnull = any ? pspan + 1ULL : 0ULL;
            pmax_code = partition->validity ? pnull : pspan;
            pbits = (any && pspan == ~0ULL) ? 65 : bits_of(pmax_code);
        }
        int32_t rc0 = int_range(ctx, order, &omin, &dummy);
        if (rc0 != FSX_OK) { fsx_col_release(ctx, o); return rc0; }
        ospan = ctx->h_scalars[3] - ctx->h_scalars[2];
        const int tsbits = bits_of(ospan);
        if (pbits + tsbits <= 64 && !getenv("FSX_WINDOW_TWO_KEYS")) {
            unsigned long long *ck = nullptr, *ck_alt = nullptr;
            unsigned *idx = nullptr, *idx_alt = nullptr;
            WAcc *prefix = nullptr, *tile_acc = nullptr;
            const int64_t ntiles = (n + WS_TILE - 1) / WS_TILE;
            int32_t rc = fsx_dev_alloc(ctx, (size_t)n * 8, (void **)&ck);
            if (rc == FSX_OK) rc = fsx_dev_alloc(ctx, (size_t)n * 8, (void **)&ck_alt);
            if (rc == FSX_OK) rc = fsx_dev_alloc(ctx, (size_t)n * 4, (void **)&idx);
            if (rc == FSX_OK) rc = fsx_dev_alloc(ctx, (size_t)n * 4, (void **)&idx_alt);
            if (rc == FSX_OK) rc = fsx_dev_alloc(ctx, (size_t)n * sizeof(WAcc), (void **)&prefix);
            if (rc == FSX_OK) rc = fsx_dev_alloc(ctx, (size_t)ntiles * sizeof(WAcc), (void **)&tile_acc);
            auto cleanup = [&]() { fsx_dev_free(ctx, ck); fsx_dev_free(ctx, ck_alt); fsx_dev_free(ctx, idx); fsx_dev_free(ctx, idx_alt); fsx_dev_free(ctx, prefix); fsx_dev_free(ctx, tile_acc); };
            if (rc != FSX_OK) { cleanup(); fsx_col_release(ctx, o); return rc; }
            const int grid = fsx_grid_for(ctx, n, 256 * 4, 8);
            k_w_make_ckeys<<<grid, 256, 0, ctx->stream>>>(partition ? partition->data : nullptr, partition ? partition->dtype : FSX_I64, partition ? partition->validity : nullptr,
                                                          pmin, pnull, order->data, order->dtype, omin, tsbits, n, ck, idx);
            ctx->launches++;
            rc = fsx_internal_sort_pairs_u64(ctx, &ck, &idx, &ck_alt, &idx_alt, n, pbits + tsbits);
            if (rc != FSX_OK) { cleanup(); fsx_col_release(ctx, o); return rc; }
            const unsigned *perm = idx;
#define W_SCAN(MODE) switch (val->dtype) { \
    case FSX_I32: k_w_scan<int32_t, MODE><<<(unsigned)ntiles, WS_THREADS, 0, ctx->stream>>>((const int32_t *)val->data, val->validity, perm, n, tile_acc, prefix); break; \
    case FSX_I64: k_w_scan<long long, MODE><<<(unsigned)ntiles, WS_THREADS, 0, ctx->stream>>>((const long long *)val->data, val->validity, perm, n, tile_acc, prefix); break; \
    case FSX_F32: k_w_scan<float, MODE><<<(unsigned)ntiles, WS_THREADS, 0, ctx->stream>>>((const float *)val->data, val->validity, perm, n, tile_acc, prefix); break; \
    default: k_w_scan<double, MODE><<<(unsigned)ntiles, WS_THREADS, 0, ctx->stream>>>((const double *)val->data, val->validity, perm, n, tile_acc, prefix); break; }
            W_SCAN(0);
            k_w_carry<<<1, 32, 0, ctx->stream>>>(tile_acc, ntiles);
            W_SCAN(1);
#undef W_SCAN
            k_w_frames_ckey<<<grid, 256, 0, ctx->stream>>>(ck, perm, prefix, n, tsbits, ospan, (long long)lo, (long long)hi, (double *)o->data, (unsigned *)o->validity);
            ctx->launches += 4;
            cudaError_t e = cudaGetLastError();
            cleanup();
            if (e != cudaSuccess) { fsx_col_release(ctx, o); FSX_CUDA(ctx, e); }
            *out = o;
            FSX_END();
        }
    }
    // ---- general route: two sort keys
    // 1. rows in (partition, order) order
    unsigned *perm = nullptr;
    const fsx_col *by[2] = {partition, order};
    int32_t rc = partition ? fsx_internal_argsort(ctx, by, 2, n, &perm) : fsx_internal_argsort(ctx, by + 1, 1, n, &perm);
    if (rc != FSX_OK) { fsx_col_release(ctx, o); return rc; }
    long long *sp = nullptr, *so = nullptr;
    unsigned char *sp_null = nullptr;
    WAcc *prefix = nullptr, *tile_acc = nullptr;
    const int64_t ntiles = (n + WS_TILE - 1) / WS_TILE;
    rc = fsx_dev_alloc(ctx, (size_t)n * 8, (void **)&sp);
    if (rc == FSX_OK) rc = fsx_dev_alloc(ctx, (size_t)n * 8, (void **)&so);
    if (rc == FSX_OK) rc = fsx_dev_alloc(ctx, (size_t)n, (void **)&sp_null);
    if (rc == FSX_OK) rc = fsx_dev_alloc(ctx, (size_t)n * sizeof(WAcc), (void **)&prefix);
    if (rc == FSX_OK) rc = fsx_dev_alloc(ctx, (size_t)ntiles * sizeof(WAcc), (void **)&tile_acc);
    auto cleanup = [&]() { fsx_dev_free(ctx, perm); fsx_dev_free(ctx, sp); fsx_dev_free(ctx, so); fsx_dev_free(ctx, sp_null); fsx_dev_free(ctx, prefix); fsx_dev_free(ctx, tile_acc); };
    if (rc != FSX_OK) { cleanup(); fsx_col_release(ctx, o); return rc; }
    const int grid = fsx_grid_for(ctx, n, 256 * 4, 8);
    k_w_sorted_keys<<<grid, 256, 0, ctx->stream>>>(partition ? partition->data : nullptr, partition ? partition->dtype : FSX_I64, partition ? partition->validity : nullptr,
                                                   order->data, order->dtype, perm, n, sp, sp_null, so);
    // 2. inclusive prefix sums of the values in sorted order
#define W_SCAN(MODE) switch (val->dtype) { \
    case FSX_I32: k_w_scan<int32_t, MODE><<<(unsigned)ntiles, WS_THREADS, 0, ctx->stream>>>((const int32_t *)val->data, val->validity, perm, n, tile_acc, prefix); break; \
    case FSX_I64: k_w_scan<long long, MODE><<<(unsigned)ntiles, WS_THREADS, 0, ctx->stream>>>((const long long *)val->data, val->validity, perm, n, tile_acc, prefix); break; \
    case FSX_F32: k_w_scan<float, MODE><<<(unsigned)ntiles, WS_THREADS, 0, ctx->stream>>>((const float *)val->data, val->validity, perm, n, tile_acc, prefix); break; \
    default: k_w_scan<double, MODE><<<(unsigned)ntiles, WS_THREADS, 0, ctx->stream>>>((const double *)val->data, val->validity, perm, n, tile_acc, prefix); break; }
    W_SCAN(0);
    k_w_carry<<<1, 32, 0, ctx->stream>>>(tile_acc, ntiles);
    W_SCAN(1);
#undef W_SCAN
    // 3. frames
    k_w_frames<<<grid, 256, 0, ctx->stream>>>(sp, sp_null, so, perm, prefix, n, (long long)lo, (long long)hi, (double *)o->data, (unsigned *)o->validity);
    ctx->launches += 5;
    cudaError_t e = cudaGetLastError();
    cleanup();
    if (e != cudaSuccess) { fsx_col_release(ctx, o); FSX_CUDA(ctx, e); }
    *out = o;
    FSX_END();
}
