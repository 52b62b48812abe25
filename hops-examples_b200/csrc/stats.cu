// stats.cu -- K4 single-pass column moments and K5 exact equal-width histograms.
//
// Replaces the descriptive-statistics and histogram parts of what hsfs computes on
// `fg.save(df)` for a feature group built with `.statisticsConfig(new StatisticsConfig(true,
// true, true))` (ComputeFeatures.scala:114,129,170,226,273,303,325); upstream that is Deequ's
// profiler / Spark describe() + RDD.histogram(20).
//
// K4: one streaming pass, 128-bit loads, per column (n, S1 = sum(x-c), S2 = sum(x-c)^2, min,
//     max) with a pilot shift c taken from the head of the column, all sums in FP64.  Block
//     partials are written to HBM and folded in a fixed order by a second tiny kernel, so the
//     result is deterministic; mean = c + S1/n, M2 = S2 - S1^2/n (benign: c ~ mean).
//     Algorithmic bytes: sizeof(T) per element (+1 bit when a validity bitmap exists).
// K5: Spark's bucket index  b = (int)(((x - min) / (max - min)) * nbins)  is monotone in x, so
//     for every bucket s there is a smallest representable x_s with b(x_s) >= s.  The host finds
//     those thresholds by bisection over the dtype's ordered bit patterns evaluating the exact
//     FP64 formula; the kernel then bins with native-dtype compares only (a float guess plus a
//     +-1 correction against the thresholds) -- bit-exact counts without an FP64 divide per
//     element.  Counters are private per lane in shared memory (no atomics, no bank conflicts),
//     folded into global u64 counts once per block.  Algorithmic bytes: sizeof(T) per element.
#include <math.h>

#include <type_traits>

#include "stats_common.cuh"

namespace {

struct MomPartial {
    double s1, s2, mn, mx;
    unsigned long long n, nan;
};

constexpr int MOM_THREADS = 256;
// ---- K4 main pass ------------------------------------------------------------------------------
// sentinels so min/max need no "seen anything yet" flag
template <typename T> struct Lim;
template <> struct Lim<float> { static __device__ __forceinline__ float hi() { return INFINITY; } static __device__ __forceinline__ float lo() { return -INFINITY; } };
template <> struct Lim<double> { static __device__ __forceinline__ double hi() { return INFINITY; } static __device__ __forceinline__ double lo() { return -INFINITY; } };
template <> struct Lim<int32_t> { static __device__ __forceinline__ int32_t hi() { return INT32_MAX; } static __device__ __forceinline__ int32_t lo() { return INT32_MIN; } };
template <> struct Lim<long long> { static __device__ __forceinline__ long long hi() { return INT64_MAX; } static __device__ __forceinline__ long long lo() { return INT64_MIN; } };

constexpr int MOM_U = 4;     // 128-bit loads in flight per thread

// Per-thread state: VEC independent (S1, S2) chains so consecutive DADD/DFMA never wait on each
// other, branch-free min/max (a NaN operand never wins a < or > comparison) and 32-bit counters.
template <typename T, int VEC>
struct MomAcc {
    double s1[VEC], s2[VEC];
    unsigned n = 0, nan = 0;
    T mn = Lim<T>::hi(), mx = Lim<T>::lo();
    __device__ __forceinline__ MomAcc() {
#pragma unroll
        for (int j = 0; j < VEC; j++) { s1[j] = 0.0; s2[j] = 0.0; }
    }
    __device__ __forceinline__ void add(int j, T x, bool ok, double c) {
        const double d = ok ? (double)x - c : 0.0;
        s1[j] += d;
        s2[j] = fma(d, d, s2[j]);
        n += ok ? 1u : 0u;
        nan += (ok && x != x) ? 1u : 0u;
        const T lo_c = ok ? x : Lim<T>::hi(), hi_c = ok ? x : Lim<T>::lo();
        mn = lo_c < mn ? lo_c : mn;
        mx = hi_c > mx ? hi_c : mx;
    }
};

template <typename T, bool HAS_VALID>
__global__ void __launch_bounds__(MOM_THREADS) k_moments(const ColDesc *__restrict__ cols,
                                                         const double *__restrict__ pilot,
                                                         MomPartial *__restrict__ partials,
                                                         int64_t rows_per_block) {
    constexpr int VEC = 16 / sizeof(T);
    const ColDesc cd = cols[blockIdx.y];
    const double c = pilot[blockIdx.y];
    const T *__restrict__ base = (const T *)cd.data;
    const uint8_t *__restrict__ valid = HAS_VALID ? cd.valid : nullptr;
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
    int64_t r1 = r0 + rows_per_block;
    if (r1 > cd.n) r1 = cd.n;
    MomAcc<T, VEC> a;
    if (r0 < r1) {
        int64_t done = r0;
        if (((uintptr_t)base & 15) == 0) {
            // rows_per_block is a multiple of 4096, so r0 is vector (and validity-byte) aligned
            const int64_t nvec = (r1 - r0) / VEC;
            const int64_t nbatch = nvec / (MOM_THREADS * MOM_U);
            for (int64_t b = 0; b < nbatch; b++) {
                const int64_t v0 = b * (MOM_THREADS * MOM_U) + threadIdx.x;
                int4 raw[MOM_U];
                unsigned bits[MOM_U];
#pragma unroll
                for (int u = 0; u < MOM_U; u++) raw[u] = ld_stream_v4(base + r0 + (v0 + (int64_t)u * MOM_THREADS) * VEC);
#pragma unroll
                for (int u = 0; u < MOM_U; u++) {
                    const int64_t row = r0 + (v0 + (int64_t)u * MOM_THREADS) * VEC;
                    bits[u] = valid ? ((unsigned)valid[row >> 3] >> (row & 7)) : 0xFFu;
                }
#pragma unroll
                for (int u = 0; u < MOM_U; u++) {
                    T x[VEC];
                    memcpy(x, &raw[u], 16);
#pragma unroll
                    for (int j = 0; j < VEC; j++) a.add(j, x[j], !HAS_VALID || ((bits[u] >> j) & 1u), c);
                }
            }
            for (int64_t v = nbatch * (MOM_THREADS * MOM_U) + threadIdx.x; v < nvec; v += MOM_THREADS) {
                const int64_t row = r0 + v * VEC;
                int4 raw = ld_stream_v4(base + row);
                const unsigned bits = valid ? ((unsigned)valid[row >> 3] >> (row & 7)) : 0xFFu;
                T x[VEC];
                memcpy(x, &raw, 16);
#pragma unroll
                for (int j = 0; j < VEC; j++) a.add(j, x[j], (bits >> j) & 1u, c);
            }
            done = r0 + nvec * VEC;
        }
        for (int64_t i = done + threadIdx.x; i < r1; i += MOM_THREADS) a.add(0, base[i], dev_bit_valid(valid, i), c);
    }
    // block fold: warp shuffles, then 8 warp partials through shared memory
    double ts1 = 0.0, ts2 = 0.0;
#pragma unroll
    for (int j = 0; j < VEC; j++) { ts1 += a.s1[j]; ts2 += a.s2[j]; }
    const double s1 = warp_sum(ts1), s2 = warp_sum(ts2);
    const unsigned long long n = warp_sum((unsigned long long)a.n), nan = warp_sum((unsigned long long)a.nan);
    double mn = (double)a.mn, mx = (double)a.mx;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    __shared__ MomPartial wp[MOM_THREADS / 32];
    if ((threadIdx.x & 31) == 0) wp[threadIdx.x >> 5] = MomPartial{s1, s2, mn, mx, n, nan};
    __syncthreads();
    if (threadIdx.x == 0) {
        MomPartial t = wp[0];
        for (int w = 1; w < MOM_THREADS / 32; w++) {
            t.s1 += wp[w].s1; t.s2 += wp[w].s2; t.n += wp[w].n; t.nan += wp[w].nan;
            t.mn = fmin(t.mn, wp[w].mn); t.mx = fmax(t.mx, wp[w].mx);
        }
        partials[(size_t)blockIdx.y * gridDim.x + blockIdx.x] = t;
    }
}

// one block per column: fixed-order fold of the block partials -> fsx_moments
__global__ void __launch_bounds__(256) k_moments_final(const ColDesc *__restrict__ cols,
                                                       const double *__restrict__ pilot,
                                                       const MomPartial *__restrict__ partials,
                                                       int nblocks, fsx_moments *__restrict__ out) {
    const int col = blockIdx.x;
    double s1 = 0, s2 = 0, mn = INFINITY, mx = -INFINITY;
    unsigned long long n = 0, nan = 0;
    for (int b = threadIdx.x; b < nblocks; b += blockDim.x) {
        const MomPartial p = partials[(size_t)col * nblocks + b];
        s1 += p.s1; s2 += p.s2; n += p.n; nan += p.nan;
        mn = fmin(mn, p.mn); mx = fmax(mx, p.mx);
    }
    s1 = warp_sum(s1); s2 = warp_sum(s2); n = warp_sum(n); nan = warp_sum(nan);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    __shared__ MomPartial wp[8];
    if ((threadIdx.x & 31) == 0) wp[threadIdx.x >> 5] = MomPartial{s1, s2, mn, mx, n, nan};
    __syncthreads();
    if (threadIdx.x == 0) {
        MomPartial t = wp[0];
        for (int w = 1; w < 8; w++) {
            t.s1 += wp[w].s1; t.s2 += wp[w].s2; t.n += wp[w].n; t.nan += wp[w].nan;
            t.mn = fmin(t.mn, wp[w].mn); t.mx = fmax(t.mx, wp[w].mx);
        }
        fsx_moments m;
        const double c = pilot[col];
        m.n_rows = cols[col].n;
        m.n_nonnull = (int64_t)t.n;
        if (t.n == 0) { m.mean = 0; m.m2 = 0; m.min = 0; m.max = 0; m.sum = 0; }
        else {
            const double dn = (double)t.n;
            m.mean = c + t.s1 / dn;
            double m2 = t.s2 - t.s1 * t.s1 / dn;
            m.m2 = m2 < 0 ? 0 : m2;
            m.sum = t.s1 + dn * c;
            const double qnan = __longlong_as_double(0x7FF8000000000000LL);
            // Spark ordering: NaN greatest -> max is NaN if any NaN; min is NaN only if all are
            m.max = t.nan ? qnan : t.mx;
            m.min = (t.nan == t.n) ? qnan : t.mn;
        }
        out[col] = m;
    }
}

// ---- K5 ------------------------------------------------------------------------------------------
constexpr int HIST_MAX_BINS = 1024;

template <typename T>
struct HistCol {                 // per column, native dtype
    T lo, hi;                    // smallest / largest native value inside [lo_f64, hi_f64]
    double lo_d, inv_width;      // guess only: t = (x - lo_d) * inv_width ~ bucket index + fraction
    int nb;                      // effective buckets (1 when min == max)
    int skip;                    // column has an empty range: count nothing
    int always_exact;            // guess arithmetic cannot be trusted (|int64| > 2^52, degenerate width)
};

// The guess runs in float for float columns (x - lo is a float subtraction, total error < 4e-6
// buckets) and in double for every other dtype (error ~1e-14 buckets).  A guess whose fractional
// part is at least GUESS_EPS away from a bucket boundary is provably the exact Spark bucket; the
// others (a fraction ~2*GUESS_EPS of the elements) take the exact threshold search.
template <typename T> struct Guess { using type = double; static constexpr double eps = 1e-9; };
template <> struct Guess<float> { using type = float; static constexpr float eps = 2.5e-4f; };

template <typename T, bool HAS_VALID>
__global__ void __launch_bounds__(256) k_histogram(const ColDesc *__restrict__ cols,
                                                   const HistCol<T> *__restrict__ hc,
                                                   const T *__restrict__ thresholds, // [ncols][nbins] : T[s] for s=0..nbins-1
                                                   int nbins, int64_t rows_per_block,
                                                   unsigned long long *__restrict__ counts) {
    extern __shared__ unsigned smem_u32[];
    constexpr int VEC = 16 / sizeof(T);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    unsigned *cnt = smem_u32 + (size_t)warp * nbins * 32;   // [bin][lane]
    T *thr = (T *)(smem_u32 + (size_t)nwarps * nbins * 32);  // [nbins]
    const ColDesc cd = cols[blockIdx.y];
    const HistCol<T> h = hc[blockIdx.y];
    for (int i = threadIdx.x; i < nwarps * nbins * 32; i += blockDim.x) smem_u32[i] = 0;
    for (int i = threadIdx.x; i < nbins; i += blockDim.x) thr[i] = thresholds[(size_t)blockIdx.y * nbins + i];
    __syncthreads();
    const T *__restrict__ base = (const T *)cd.data;
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
    int64_t r1 = r0 + rows_per_block;
    if (r1 > cd.n) r1 = cd.n;
    const int last = h.nb - 1;
    using GT = typename Guess<T>::type;
    // always_exact columns get an eps no fraction can clear, so every value takes the exact path without a test of its own
    const GT lo_g = (GT)h.lo_d, inv_g = (GT)h.inv_width, eps = h.always_exact ? (GT)1 : Guess<T>::eps;
    auto exact_bin = [&](T x, int g) {
        g = g < 0 ? 0 : (g > last ? last : g);
        while (g < last && x >= thr[g + 1]) g++;
        while (g > 0 && x < thr[g]) g--;
        return g;
    };
    // Fast path proof: |t - t_exact| < eps, so a guess with g in [0,last] whose fraction is more
    // than eps away from both bucket edges has floor(t_exact) == g and lo < x < hi (no range test
    // needed); NaN makes every comparison false and falls to the exact path, which drops it.
    // counters are private to a lane (bank = lane): the shared-memory RED is conflict free and one instruction where a
    // load / add / store would be three issue slots of a kernel that is bound by issue slots
    auto slow_one = [&](T x, int g) {
        if (x >= h.lo && x <= h.hi) atomicAdd(&cnt[exact_bin(x, g) * 32 + lane], 1u);
    };
    auto bin_one = [&](T x, bool ok) {
        int g;
        bool safe;
        if constexpr (std::is_same<T, float>::value) {
            // no conversion-pipe instructions: round (t - 0.5) to nearest with the 2^23 trick
            const float u = fmaf(x - lo_g, inv_g, -0.5f);
            const float r = u + 8388608.0f;
            g = __float_as_int(r) - 0x4B000000;
            const float fr = u - (r - 8388608.0f);                  // t - g - 0.5, in (-0.5, 0.5]
            safe = fabsf(fr) < 0.5f - eps && (unsigned)g <= (unsigned)last;
        } else {
            const GT t = ((GT)x - lo_g) * inv_g;
            g = (int)t;
            const GT frac = t - (GT)g;
            safe = frac > eps && frac < (GT)1 - eps && (unsigned)g <= (unsigned)last;
        }
        if (!HAS_VALID || ok) {
            if (safe) atomicAdd(&cnt[g * 32 + lane], 1u);
            else slow_one(x, g);
        }
    };
    constexpr int HU = 4;
    if (!h.skip && r0 < r1) {
        int64_t done = r0;
        if (((uintptr_t)base & 15) == 0) {
            const int64_t nvec = (r1 - r0) / VEC;
            const int64_t per_batch = (int64_t)blockDim.x * HU;
            const int64_t nbatch = nvec / per_batch;
            for (int64_t b = 0; b < nbatch; b++) {
                const int64_t v0 = b * per_batch + threadIdx.x;
                int4 raw[HU];
                unsigned bits[HU];
#pragma unroll
                for (int u = 0; u < HU; u++) raw[u] = ld_stream_v4(base + r0 + (v0 + (int64_t)u * blockDim.x) * VEC);
#pragma unroll
                for (int u = 0; u < HU; u++) {
                    const int64_t row = r0 + (v0 + (int64_t)u * blockDim.x) * VEC;
                    bits[u] = HAS_VALID ? ((unsigned)cd.valid[row >> 3] >> (row & 7)) : 0xFFu;
                }
#pragma unroll
                for (int u = 0; u < HU; u++) {
                    T x[VEC];
                    memcpy(x, &raw[u], 16);
#pragma unroll
                    for (int j = 0; j < VEC; j++) bin_one(x[j], (bits[u] >> j) & 1u);
                }
            }
            for (int64_t v = nbatch * per_batch + threadIdx.x; v < nvec; v += blockDim.x) {
                const int64_t row = r0 + v * VEC;
                int4 raw = ld_stream_v4(base + row);
                const unsigned bits = HAS_VALID ? ((unsigned)cd.valid[row >> 3] >> (row & 7)) : 0xFFu;
                T x[VEC];
                memcpy(x, &raw, 16);
#pragma unroll
                for (int j = 0; j < VEC; j++) bin_one(x[j], (bits >> j) & 1u);
            }
            done = r0 + nvec * VEC;
        }
        for (int64_t i = done + threadIdx.x; i < r1; i += blockDim.x) bin_one(base[i], !HAS_VALID || dev_bit_valid(cd.valid, i));
    }
    __syncthreads();
    // fold lanes and warps; one global atomic per (block, bin)
    for (int b = warp; b < nbins; b += nwarps) {
        unsigned long long t = 0;
        for (int w = 0; w < nwarps; w++) t += smem_u32[((size_t)w * nbins + b) * 32 + lane];
        t = warp_sum(t);
        if (lane == 0 && t) atomicAdd(&counts[(size_t)blockIdx.y * nbins + b], t);
    }
}

// ---- host side of K5: exact thresholds by bisection over ordered bit patterns ----------------
// Spark DoubleRDDFunctions.histogram fast path, literally.
inline int spark_bucket(double e, double lo, double hi, int nbins) {
    volatile double num = e - lo;
    volatile double den = hi - lo;
    volatile double q = num / den;
    volatile double scaled = q * (double)nbins;
    int b = (int)scaled;
    return b > nbins - 1 ? nbins - 1 : b;
}

template <typename T> struct Ord;
template <> struct Ord<float> {
    using U = uint32_t;
    static U to(float f) { U b; memcpy(&b, &f, 4); return (b & 0x80000000u) ? ~b : (b | 0x80000000u); }
    static float from(U u) { U b = (u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u; float f; memcpy(&f, &b, 4); return f; }
};
template <> struct Ord<double> {
    using U = uint64_t;
    static U to(double f) { U b; memcpy(&b, &f, 8); return (b >> 63) ? ~b : (b | 0x8000000000000000ULL); }
    static double from(U u) { U b = (u >> 63) ? (u & 0x7FFFFFFFFFFFFFFFULL) : ~u; double f; memcpy(&f, &b, 8); return f; }
};
template <> struct Ord<int32_t> {
    using U = uint32_t;
    static U to(int32_t v) { return (U)v ^ 0x80000000u; }
    static int32_t from(U u) { return (int32_t)(u ^ 0x80000000u); }
};
template <> struct Ord<long long> {
    using U = uint64_t;
    static U to(long long v) { return (U)v ^ 0x8000000000000000ULL; }
    static long long from(U u) { return (long long)(u ^ 0x8000000000000000ULL); }
};
template <typename T> T type_lowest();
template <> float type_lowest<float>() { return -INFINITY; }
template <> double type_lowest<double>() { return -INFINITY; }
template <> int32_t type_lowest<int32_t>() { return INT32_MIN; }
template <> long long type_lowest<long long>() { return INT64_MIN; }
template <typename T> T type_highest();
template <> float type_highest<float>() { return INFINITY; }
template <> double type_highest<double>() { return INFINITY; }
template <> int32_t type_highest<int32_t>() { return INT32_MAX; }
template <> long long type_highest<long long>() { return INT64_MAX; }

// smallest native x in [a,b] (ordered) with pred(x) true; pred monotone false->true; returns
// b+1-equivalent flag through `found`
template <typename T, typename P>
T bisect_first(T a, T b, P pred, bool &found) {
    using U = typename Ord<T>::U;
    U lo = Ord<T>::to(a), hi = Ord<T>::to(b);
    if (!pred(Ord<T>::from(hi))) { found = false; return b; }
    while (lo < hi) {
        U mid = lo + (hi - lo) / 2;
        if (pred(Ord<T>::from(mid))) hi = mid; else lo = mid + 1;
    }
    found = true;
    return Ord<T>::from(lo);
}

template <typename T>
void plan_hist_col(double lo, double hi, int nbins, HistCol<T> &h, T *thr) {
    bool found = false;
    // native clamp of the f64 bounds: lo_n = first native x with (double)x >= lo, hi_n = last with <= hi
    T lo_n = bisect_first<T>(type_lowest<T>(), type_highest<T>(), [&](T x) { return (double)x >= lo; }, found);
    bool ok = found;
    T first_above = bisect_first<T>(type_lowest<T>(), type_highest<T>(), [&](T x) { return (double)x > hi; }, found);
    T hi_n;
    if (!found) hi_n = type_highest<T>();
    else {
        using U = typename Ord<T>::U;
        U u = Ord<T>::to(first_above);
        if (u == 0) ok = false;
        hi_n = Ord<T>::from(u - 1);
    }
    h.skip = !ok || !(lo_n <= hi_n);
    h.lo = lo_n; h.hi = hi_n;
    h.nb = (lo == hi) ? 1 : nbins;
    double w = (hi - lo) / (double)nbins;
    h.lo_d = lo;
    h.inv_width = (w > 0 && isfinite(1.0 / w)) ? 1.0 / w : 0.0;
    h.always_exact = !(h.inv_width > 0.0) || h.nb == 1 || fabs(lo) > 4.5e15 || fabs(hi) > 4.5e15 ||
                     (std::is_same<T, float>::value &&
                      (!(fabs((double)(float)h.inv_width) < 3e38) || (double)(float)lo != lo));   // float guess needs an exact lo
    thr[0] = lo_n;
    for (int s = 1; s < nbins; s++) {
        if (h.skip || h.nb == 1) { thr[s] = hi_n; continue; }
        bool f2 = false;
        T t = bisect_first<T>(lo_n, hi_n, [&](T x) { return spark_bucket((double)x, lo, hi, nbins) >= s; }, f2);
        // not found: no value reaches bucket s; park the threshold above every in-range value
        thr[s] = f2 ? t : type_highest<T>();
        if (!f2) {
            // keep x >= thr false for all in-range x; for x == type_highest the bucket is < s anyway,
            // but "x >= highest" would be true: guard by making later thresholds unreachable via nb
            h.nb = h.nb > s ? s : h.nb;
        }
    }
}

// rows per block: fill ~4 waves of the machine per column set, in multiples of 4096 rows
int64_t pick_rows_per_block(const fsx_ctx *ctx, int64_t n, int ncols, int *nblocks_out) {
    int64_t target_blocks = ((int64_t)ctx->sm_count * 8 + ncols - 1) / ncols;
    if (target_blocks < 1) target_blocks = 1;
    int64_t rpb = (n + target_blocks - 1) / target_blocks;
    rpb = ((rpb + 4095) / 4096) * 4096;
    if (rpb < 4096) rpb = 4096;
    int64_t nb = (n + rpb - 1) / rpb;
    if (nb < 1) nb = 1;
    *nblocks_out = (int)nb;
    return rpb;
}

int32_t describe_impl(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols, fsx_moments *out) {
    int64_t n = 0;
    FSX_CHECK(check_cols(ctx, cols, ncols, "fsx_stats_describe", false, &n));
    FSX_REQUIRE(ctx, out, FSX_E_INVALID, "fsx_stats_describe: out is NULL");
    ColDesc *d_cols = nullptr;
    FSX_CHECK(upload_desc(ctx, cols, ncols, &d_cols));
    double *d_pilot = nullptr;
    fsx_moments *d_out = nullptr;
    MomPartial *d_part = nullptr;
    int nblocks = 1;
    int64_t rpb = pick_rows_per_block(ctx, n, ncols, &nblocks);
    int32_t rc = fsx_dev_alloc(ctx, sizeof(double) * (size_t)ncols, (void **)&d_pilot);
    if (rc == FSX_OK) rc = fsx_dev_alloc(ctx, sizeof(fsx_moments) * (size_t)ncols, (void **)&d_out);
    if (rc == FSX_OK) rc = fsx_dev_alloc(ctx, sizeof(MomPartial) * (size_t)ncols * (size_t)nblocks, (void **)&d_part);
    auto cleanup = [&]() { fsx_dev_free(ctx, d_cols); fsx_dev_free(ctx, d_pilot); fsx_dev_free(ctx, d_out); fsx_dev_free(ctx, d_part); };
    if (rc != FSX_OK) { cleanup(); return rc; }
    FSX_LAUNCH(ctx, k_pilot, ncols, 256, 0, d_cols, d_pilot);
    // launch per run of equal dtypes so each kernel instance is fully typed
    for (int c0 = 0; c0 < ncols;) {
        int c1 = c0;
        while (c1 < ncols && cols[c1]->dtype == cols[c0]->dtype) c1++;
        dim3 grid((unsigned)nblocks, (unsigned)(c1 - c0));
        MomPartial *part = d_part + (size_t)c0 * nblocks;
        bool any_valid = false;
        for (int c = c0; c < c1; c++) any_valid |= cols[c]->validity != nullptr;
#define MOM_LAUNCH(T)                                                                                                  \
    do {                                                                                                               \
        if (any_valid) k_moments<T, true><<<grid, MOM_THREADS, 0, ctx->stream>>>(d_cols + c0, d_pilot + c0, part, rpb); \
        else k_moments<T, false><<<grid, MOM_THREADS, 0, ctx->stream>>>(d_cols + c0, d_pilot + c0, part, rpb);          \
        ctx->launches++;                                                                                               \
    } while (0)
        switch (cols[c0]->dtype) {
        case FSX_I32: MOM_LAUNCH(int32_t); break;
        case FSX_I64: MOM_LAUNCH(long long); break;
        case FSX_F32: MOM_LAUNCH(float); break;
        default: MOM_LAUNCH(double); break;
        }
#undef MOM_LAUNCH
        {
            cudaError_t le = cudaGetLastError();
            if (le != cudaSuccess) { cleanup(); FSX_CUDA(ctx, le); }
        }
        c0 = c1;
    }
    FSX_LAUNCH(ctx, k_moments_final, ncols, 256, 0, d_cols, d_pilot, d_part, nblocks, d_out);
    cudaError_t e = cudaMemcpyAsync(out, d_out, sizeof(fsx_moments) * (size_t)ncols, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cleanup();
    FSX_CUDA(ctx, e);
    ctx->timing.d2h_bytes += (int64_t)(sizeof(fsx_moments) * (size_t)ncols);
    return FSX_OK;
}

template <typename T>
int32_t histogram_typed(fsx_ctx *ctx, const fsx_col *const *cols, int c0, int c1, int nbins,
                        const double *lo, const double *hi, const ColDesc *d_cols, int64_t n,
                        unsigned long long *d_counts, int32_t *nbuckets_out) {
    const int nc = c1 - c0;
    std::vector<HistCol<T>> hc((size_t)nc);
    std::vector<T> thr((size_t)nc * (size_t)nbins);
    for (int i = 0; i < nc; i++) {
        plan_hist_col<T>(lo[c0 + i], hi[c0 + i], nbins, hc[(size_t)i], thr.data() + (size_t)i * nbins);
        if (nbuckets_out) nbuckets_out[c0 + i] = (lo[c0 + i] == hi[c0 + i]) ? 1 : nbins;
        (void)cols;
    }
    HistCol<T> *d_hc = nullptr;
    T *d_thr = nullptr;
    int32_t rc = fsx_dev_alloc(ctx, sizeof(HistCol<T>) * (size_t)nc, (void **)&d_hc);
    if (rc == FSX_OK) rc = fsx_dev_alloc(ctx, sizeof(T) * (size_t)nc * (size_t)nbins, (void **)&d_thr);
    auto cleanup = [&]() { fsx_dev_free(ctx, d_hc); fsx_dev_free(ctx, d_thr); };
    if (rc != FSX_OK) { cleanup(); return rc; }
    cudaError_t e = cudaMemcpyAsync(d_hc, hc.data(), sizeof(HistCol<T>) * (size_t)nc, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_thr, thr.data(), sizeof(T) * (size_t)nc * (size_t)nbins, cudaMemcpyHostToDevice, ctx->stream);
    if (e != cudaSuccess) { cleanup(); FSX_CUDA(ctx, e); }
    // warps per block bounded by the per-lane counter footprint (nbins * 128 B per warp)
    int warps = 8;
    while (warps > 1 && (size_t)warps * nbins * 128 + (size_t)nbins * sizeof(T) > 200 * 1024) warps >>= 1;
    size_t smem = (size_t)warps * nbins * 128 + (size_t)nbins * sizeof(T);
    bool any_valid = false;                               // one grid covers the whole dtype run: the bitmap-free kernel needs all of it bitmap free
    for (int i = c0; i < c1; i++) any_valid |= cols[i]->validity != nullptr;
    auto kern = any_valid ? k_histogram<T, true> : k_histogram<T, false>;
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) { cleanup(); FSX_CUDA(ctx, e); }
    int nblocks = 1;
    int64_t rpb = pick_rows_per_block(ctx, n, nc, &nblocks);
    dim3 grid((unsigned)nblocks, (unsigned)nc);
    kern<<<grid, warps * 32, smem, ctx->stream>>>(d_cols + c0, d_hc, d_thr, nbins, rpb, d_counts + (size_t)c0 * nbins);
    ctx->launches++;
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream); // hc/thr host vectors die here
    cleanup();
    FSX_CUDA(ctx, e);
    return FSX_OK;
}

} // namespace

// internal entry for other translation units (gram_tc.cu needs min/max/mean before it can scale)
int32_t fsx_internal_describe(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols, fsx_moments *out) {
    return describe_impl(ctx, cols, ncols, out);
}

extern "C" int32_t fsx_stats_describe(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols,
                                      fsx_moments *out) {
    FSX_BEGIN(ctx);
    FSX_CHECK(describe_impl(ctx, cols, ncols, out));
    FSX_END();
}

// mom: optional precomputed moments (bounds for NaN lo/hi entries); lenient: a column without usable
// bounds (no non-null value, NaN or infinite extremes) gets zero counts and nbuckets 0 instead of an error
static int32_t histogram_impl(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols, int32_t nbins, const double *lo,
                              const double *hi, const fsx_moments *mom_in, bool lenient, int64_t *counts_out,
                              double *edges_out, int32_t *nbuckets_out) {
    int64_t n = 0;
    FSX_CHECK(check_cols(ctx, cols, ncols, "fsx_stats_histogram", false, &n));
    FSX_REQUIRE(ctx, counts_out, FSX_E_INVALID, "fsx_stats_histogram: counts_out is NULL");
    FSX_REQUIRE(ctx, nbins >= 1 && nbins <= HIST_MAX_BINS, FSX_E_INVALID,
                "fsx_stats_histogram: nbins %d outside [1,%d]", nbins, (int)HIST_MAX_BINS);
    // bounds: caller-supplied, or the column's min/max (Spark's first job)
    std::vector<double> vlo((size_t)ncols), vhi((size_t)ncols);
    bool need_data = false;
    for (int i = 0; i < ncols; i++) {
        vlo[(size_t)i] = lo ? lo[i] : NAN;
        vhi[(size_t)i] = hi ? hi[i] : NAN;
        if (isnan(vlo[(size_t)i]) || isnan(vhi[(size_t)i])) need_data = true;
    }
    std::vector<char> unusable((size_t)ncols, 0);
    if (need_data) {
        std::vector<fsx_moments> mom_local;
        const fsx_moments *mom = mom_in;
        if (!mom) {
            mom_local.resize((size_t)ncols);
            FSX_CHECK(describe_impl(ctx, cols, ncols, mom_local.data()));
            mom = mom_local.data();
        }
        for (int i = 0; i < ncols; i++) {
            const fsx_moments &m = mom[(size_t)i];
            if (!(isnan(vlo[(size_t)i]) || isnan(vhi[(size_t)i]))) continue;
            if (lenient && (m.n_nonnull == 0 || isnan(m.max) || isnan(m.min) || isinf(m.max) || isinf(m.min))) {
                unusable[(size_t)i] = 1; vlo[(size_t)i] = 1.0; vhi[(size_t)i] = -1.0;     // empty range: nothing is counted
                continue;
            }
            FSX_REQUIRE(ctx, m.n_nonnull > 0, FSX_E_DATA,
                        "fsx_stats_histogram: column %d has no non-null value to take bounds from", i);
            FSX_REQUIRE(ctx, !isnan(m.max) && !isnan(m.min), FSX_E_DATA,
                        "fsx_stats_histogram: column %d contains NaN (Spark histogram rejects NaN bounds)", i);
            if (isnan(vlo[(size_t)i])) vlo[(size_t)i] = m.min;
            if (isnan(vhi[(size_t)i])) vhi[(size_t)i] = m.max;
        }
    }
    for (int i = 0; i < ncols; i++)
        FSX_REQUIRE(ctx, unusable[(size_t)i] || (!isinf(vlo[(size_t)i]) && !isinf(vhi[(size_t)i]) && vlo[(size_t)i] <= vhi[(size_t)i]), FSX_E_DATA,
                    "fsx_stats_histogram: column %d has infinite or inverted bounds [%g,%g]", i, vlo[(size_t)i], vhi[(size_t)i]);
    if (edges_out) {
        for (int i = 0; i < ncols; i++) {
            double *e = edges_out + (size_t)i * (nbins + 1);
            const double a = vlo[(size_t)i], b = vhi[(size_t)i];
            if (a == b) { for (int s = 0; s <= nbins; s++) e[s] = a; continue; }
            volatile double span = b - a;
            for (int s = 0; s < nbins; s++) { volatile double num = (double)s * span; e[s] = a + num / (double)nbins; }
            e[nbins] = b;
        }
    }
    ColDesc *d_cols = nullptr;
    FSX_CHECK(upload_desc(ctx, cols, ncols, &d_cols));
    unsigned long long *d_counts = nullptr;
    int32_t rc = fsx_dev_alloc(ctx, 8 * (size_t)ncols * (size_t)nbins, (void **)&d_counts);
    if (rc != FSX_OK) { fsx_dev_free(ctx, d_cols); return rc; }
    FSX_ASYNC(ctx, cudaMemsetAsync(d_counts, 0, 8 * (size_t)ncols * (size_t)nbins, ctx->stream));
    for (int c0 = 0; c0 < ncols && rc == FSX_OK;) {
        int c1 = c0;
        while (c1 < ncols && cols[c1]->dtype == cols[c0]->dtype) c1++;
        switch (cols[c0]->dtype) {
        case FSX_I32: rc = histogram_typed<int32_t>(ctx, cols, c0, c1, nbins, vlo.data(), vhi.data(), d_cols, n, d_counts, nbuckets_out); break;
        case FSX_I64: rc = histogram_typed<long long>(ctx, cols, c0, c1, nbins, vlo.data(), vhi.data(), d_cols, n, d_counts, nbuckets_out); break;
        case FSX_F32: rc = histogram_typed<float>(ctx, cols, c0, c1, nbins, vlo.data(), vhi.data(), d_cols, n, d_counts, nbuckets_out); break;
        default: rc = histogram_typed<double>(ctx, cols, c0, c1, nbins, vlo.data(), vhi.data(), d_cols, n, d_counts, nbuckets_out); break;
        }
        c0 = c1;
    }
    cudaError_t e = cudaSuccess;
    if (rc == FSX_OK) {
        e = cudaMemcpyAsync(counts_out, d_counts, 8 * (size_t)ncols * (size_t)nbins, cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    }
    fsx_dev_free(ctx, d_cols);
    fsx_dev_free(ctx, d_counts);
    if (rc != FSX_OK) return rc;
    FSX_CUDA(ctx, e);
    if (nbuckets_out)
        for (int i = 0; i < ncols; i++)
            if (unusable[(size_t)i]) nbuckets_out[i] = 0;
    ctx->timing.d2h_bytes += (int64_t)(8 * (size_t)ncols * (size_t)nbins);
    return FSX_OK;
}

int32_t fsx_internal_histogram(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols, int32_t nbins, const double *lo, const double *hi,
                               const fsx_moments *mom_in, bool lenient, int64_t *counts_out, double *edges_out, int32_t *nbuckets_out) {
    return histogram_impl(ctx, cols, ncols, nbins, lo, hi, mom_in, lenient, counts_out, edges_out, nbuckets_out);
}

extern "C" int32_t fsx_stats_histogram(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols,
                                       int32_t nbins, const double *lo, const double *hi,
                                       int64_t *counts_out, double *edges_out,
                                       int32_t *nbuckets_out) {
    FSX_BEGIN(ctx);
    FSX_CHECK(histogram_impl(ctx, cols, ncols, nbins, lo, hi, nullptr, false, counts_out, edges_out, nbuckets_out));
    FSX_END();
}

int32_t fsx_internal_pearson_with_moments(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols, const fsx_moments *mom,
                                          double *corr_out);

// One call for everything hsfs computes on save with StatisticsConfig(true, true, true)
// (ComputeFeatures.scala:114): the describe pass is shared -- its min/max are the histogram bounds
// and its mean/min/max scale the tensor-core Gram.
extern "C" int32_t fsx_stats_profile(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols, int32_t nbins,
                                     fsx_moments *moments_out, int64_t *counts_out, double *edges_out,
                                     int32_t *nbuckets_out, double *corr_out) {
    FSX_BEGIN(ctx);
    FSX_REQUIRE(ctx, moments_out, FSX_E_INVALID, "fsx_stats_profile: moments_out is NULL");
    FSX_CHECK(describe_impl(ctx, cols, ncols, moments_out));
    if (counts_out)
        FSX_CHECK(histogram_impl(ctx, cols, ncols, nbins, nullptr, nullptr, moments_out, true, counts_out, edges_out, nbuckets_out));
    if (corr_out) FSX_CHECK(fsx_internal_pearson_with_moments(ctx, cols, ncols, moments_out, corr_out));
    FSX_END();
}
