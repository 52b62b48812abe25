/*
 * fsx_oracle.c -- CPU restatement of the featurestore_tour hot path (plain C11).
 *
 * TEST INFRASTRUCTURE ONLY -- see fsx_oracle.h.  Parity status per function:
 * pinned by a reference golden where one exists, otherwise
 * "parity unpinned" and following the published Spark 2.4.3 / MLlib algorithm.
 *
 * Build: make -C oracle        (gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC)
 * -ffp-contract=off matters: the histogram bin formula and the generators must
 * not be fused into FMAs, or bit-exactness against the CUDA side is lost.
 */
#include "fsx_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static inline int bit_valid(const uint8_t *v, int64_t i) {
    return v == NULL || ((v[i >> 3] >> (i & 7)) & 1);
}

static inline double load_as_double(int dtype, const void *p, int64_t i) {
    switch (dtype) {
    case ORC_I32: return (double)((const int32_t *)p)[i];
    case ORC_I64: return (double)((const int64_t *)p)[i];
    case ORC_F32: return (double)((const float *)p)[i];
    default: return ((const double *)p)[i];
    }
}

/* ------------------------------------------------------------------------- */
/* groupBy("team_id") -- CF:148-149,206-207,245-246.                          */
/* Spark semantics restated (upstream RelationalGrouped-  */
/* Dataset / HashAggregateExec): a null key forms its own group; groups are   */
/* emitted here in first-appearance order (Spark's order is unspecified, the  */
/* tests compare after sorting by key).                                       */
/* ------------------------------------------------------------------------- */
static inline uint64_t hash64(uint64_t x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL;
    x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL;
    x ^= x >> 33; return x;
}

int64_t orc_group_ids(int key_dtype, const void *keys, const uint8_t *key_valid, int64_t n,
                      int32_t *gid_out, void *uniq_keys_out, int64_t *null_gid_out) {
    int64_t cap = 16;
    while (cap < 2 * n) cap <<= 1;
    int64_t *slot_key = (int64_t *)malloc(sizeof(int64_t) * cap);
    int32_t *slot_gid = (int32_t *)malloc(sizeof(int32_t) * cap);
    for (int64_t i = 0; i < cap; i++) slot_gid[i] = -1;
    int64_t G = 0, null_gid = -1;
    for (int64_t i = 0; i < n; i++) {
        if (!bit_valid(key_valid, i)) {
            if (null_gid < 0) {
                null_gid = G++;
                if (key_dtype == ORC_I32) ((int32_t *)uniq_keys_out)[null_gid] = 0;
                else ((int64_t *)uniq_keys_out)[null_gid] = 0;
            }
            gid_out[i] = (int32_t)null_gid;
            continue;
        }
        int64_t k = key_dtype == ORC_I32 ? (int64_t)((const int32_t *)keys)[i]
                                         : ((const int64_t *)keys)[i];
        uint64_t s = hash64((uint64_t)k) & (uint64_t)(cap - 1);
        for (;;) {
            if (slot_gid[s] < 0) {
                slot_key[s] = k; slot_gid[s] = (int32_t)G;
                if (key_dtype == ORC_I32) ((int32_t *)uniq_keys_out)[G] = (int32_t)k;
                else ((int64_t *)uniq_keys_out)[G] = k;
                gid_out[i] = (int32_t)G++;
                break;
            }
            if (slot_key[s] == k) { gid_out[i] = slot_gid[s]; break; }
            s = (s + 1) & (uint64_t)(cap - 1);
        }
    }
    free(slot_key); free(slot_gid);
    if (null_gid_out) *null_gid_out = null_gid;
    return G;
}

/* groupBy(...).count() -- CF:149,207,246: count(1), every row of the group. */
void orc_group_count(const int32_t *gid, int64_t n, int64_t G, int64_t *count_out) {
    memset(count_out, 0, sizeof(int64_t) * (size_t)G);
    for (int64_t i = 0; i < n; i++) count_out[gid[i]]++;
}

/* Spark ordering for floating min/max: NaN is greater than any other value. */
static inline int dbl_lt(double a, double b) {
    if (isnan(a)) return 0;
    if (isnan(b)) return 1;
    return a < b;
}

/*
 * groupBy(...).sum() -- CF:148,206,245 -- plus min/max/non-null count for the C2
 * config.  Sum: integral input -> int64 (wraps silently, non-ANSI Spark 2.4),
 * fractional input -> float64 accumulated in row order; nulls skipped; a group
 * with no non-null value has nn == 0 (Spark: null).  Result dtypes pinned by G6
 * (notebooks/featurestore/databricks/FeatureStoreQuickStartDatabricksAzure.ipynb
 * :1380-1384: sum(long)->long, sum(double)->double).
 * min_out/max_out are typed like the input (Spark Min/Max are type preserving).
 * Pass NULL for outputs that are not wanted.
 */
void orc_group_agg(int val_dtype, const void *vals, const uint8_t *val_valid, const int32_t *gid,
                   int64_t n, int64_t G, double *sum_f64, int64_t *sum_i64, int64_t *nn_out,
                   void *min_out, void *max_out) {
    int64_t *nn = (int64_t *)calloc((size_t)(G > 0 ? G : 1), sizeof(int64_t));
    if (sum_f64) for (int64_t g = 0; g < G; g++) sum_f64[g] = 0.0;
    if (sum_i64) memset(sum_i64, 0, sizeof(int64_t) * (size_t)G);
    for (int64_t i = 0; i < n; i++) {
        if (!bit_valid(val_valid, i)) continue;
        int32_t g = gid[i];
        int first = nn[g]++ == 0;
        switch (val_dtype) {
        case ORC_I32: {
            int32_t v = ((const int32_t *)vals)[i];
            if (sum_i64) sum_i64[g] = (int64_t)((uint64_t)sum_i64[g] + (uint64_t)(int64_t)v);
            if (sum_f64) sum_f64[g] += (double)v; /* Average casts to double first */
            if (min_out && (first || v < ((int32_t *)min_out)[g])) ((int32_t *)min_out)[g] = v;
            if (max_out && (first || v > ((int32_t *)max_out)[g])) ((int32_t *)max_out)[g] = v;
        } break;
        case ORC_I64: {
            int64_t v = ((const int64_t *)vals)[i];
            if (sum_i64) sum_i64[g] = (int64_t)((uint64_t)sum_i64[g] + (uint64_t)v);
            if (sum_f64) sum_f64[g] += (double)v;
            if (min_out && (first || v < ((int64_t *)min_out)[g])) ((int64_t *)min_out)[g] = v;
            if (max_out && (first || v > ((int64_t *)max_out)[g])) ((int64_t *)max_out)[g] = v;
        } break;
        case ORC_F32: {
            float v = ((const float *)vals)[i];
            if (sum_f64) sum_f64[g] += (double)v;
            if (min_out && (first || dbl_lt(v, ((float *)min_out)[g]))) ((float *)min_out)[g] = v;
            if (max_out && (first || dbl_lt(((float *)max_out)[g], v))) ((float *)max_out)[g] = v;
        } break;
        default: {
            double v = ((const double *)vals)[i];
            if (sum_f64) sum_f64[g] += v;
            if (min_out && (first || dbl_lt(v, ((double *)min_out)[g]))) ((double *)min_out)[g] = v;
            if (max_out && (first || dbl_lt(((double *)max_out)[g], v))) ((double *)max_out)[g] = v;
        } break;
        }
    }
    if (nn_out) memcpy(nn_out, nn, sizeof(int64_t) * (size_t)G);
    free(nn);
}

/*
 * Multithreaded baseline for BASELINE.json config 2 (int32 key, float64 value ->
 * sum, count, min, max).  Mirrors Spark's physical plan for CF:148-150: a partial
 * HashAggregate per partition (thread), then a final merge -- the stand-in for
 * "Spark local[N]" that cannot run on this image.
 */
typedef struct { int32_t key; int32_t used; double sum; int64_t cnt; double mn, mx; } mt_slot;
typedef struct { int32_t key; int32_t pad; double val; } mt_row;

int64_t orc_groupby_i32_f64_mt(const int32_t *keys, const double *vals, int64_t n, int threads,
                               int32_t *keys_out, double *sum_out, int64_t *count_out,
                               double *min_out, double *max_out, int64_t cap_out) {
    if (threads < 1) threads = 1;
    const int T = threads;
    /* Phase 1 ("Exchange hashpartitioning"): every thread scatters its row range into T
     * partitions of a shared (key,value) array.  Phase 2 ("final HashAggregate"): thread p
     * aggregates partition p in a private hash table; partitions own disjoint key sets. */
    int64_t *cnt = (int64_t *)calloc((size_t)T * (size_t)T, sizeof(int64_t)); /* [t][p] */
    int64_t *off = (int64_t *)calloc((size_t)T * (size_t)T + 1, sizeof(int64_t));
    mt_row *rows = (mt_row *)malloc(sizeof(mt_row) * (size_t)(n > 0 ? n : 1));
    mt_slot **tabs = (mt_slot **)calloc((size_t)T, sizeof(mt_slot *));
    int64_t *caps = (int64_t *)calloc((size_t)T, sizeof(int64_t));
    int64_t *pstart = (int64_t *)calloc((size_t)T + 1, sizeof(int64_t));
#pragma omp parallel num_threads(T)
    {
#ifdef _OPENMP
        int t = omp_get_thread_num();
#else
        int t = 0;
#endif
        int64_t r0 = n * t / T, r1 = n * (t + 1) / T;
        for (int64_t i = r0; i < r1; i++)
            cnt[(size_t)t * T + (hash64((uint64_t)(uint32_t)keys[i]) % (uint64_t)T)]++;
#pragma omp barrier
#pragma omp single
        {
            int64_t acc = 0;
            for (int p = 0; p < T; p++) {
                pstart[p] = acc;
                for (int tt = 0; tt < T; tt++) { off[(size_t)tt * T + p] = acc; acc += cnt[(size_t)tt * T + p]; }
            }
            pstart[T] = acc;
        }
        for (int64_t i = r0; i < r1; i++) {
            int p = (int)(hash64((uint64_t)(uint32_t)keys[i]) % (uint64_t)T);
            mt_row *r = &rows[off[(size_t)t * T + p]++];
            r->key = keys[i]; r->val = vals[i];
        }
#pragma omp barrier
        int64_t cap = 1 << 12, used = 0;
        mt_slot *tab = (mt_slot *)calloc((size_t)cap, sizeof(mt_slot));
        for (int64_t i = pstart[t]; i < pstart[t + 1]; i++) {
            int32_t k = rows[i].key;
            double v = rows[i].val;
            if (2 * (used + 1) > cap) {
                int64_t ncap = cap * 2;
                mt_slot *nt = (mt_slot *)calloc((size_t)ncap, sizeof(mt_slot));
                for (int64_t s = 0; s < cap; s++) if (tab[s].used) {
                    uint64_t q = (hash64((uint64_t)(uint32_t)tab[s].key) >> 20) & (uint64_t)(ncap - 1);
                    while (nt[q].used) q = (q + 1) & (uint64_t)(ncap - 1);
                    nt[q] = tab[s];
                }
                free(tab); tab = nt; cap = ncap;
            }
            uint64_t s = (hash64((uint64_t)(uint32_t)k) >> 20) & (uint64_t)(cap - 1);
            for (;;) {
                if (!tab[s].used) {
                    tab[s].used = 1; tab[s].key = k; tab[s].sum = v; tab[s].cnt = 1;
                    tab[s].mn = v; tab[s].mx = v; used++;
                    break;
                }
                if (tab[s].key == k) {
                    tab[s].sum += v; tab[s].cnt++;
                    if (dbl_lt(v, tab[s].mn)) tab[s].mn = v;
                    if (dbl_lt(tab[s].mx, v)) tab[s].mx = v;
                    break;
                }
                s = (s + 1) & (uint64_t)(cap - 1);
            }
        }
        tabs[t] = tab; caps[t] = cap;
    }
    int64_t G_total = 0;
    for (int t = 0; t < T; t++) {
        for (int64_t s = 0; s < caps[t]; s++) if (tabs[t][s].used) {
            if (G_total < cap_out) {
                keys_out[G_total] = tabs[t][s].key; sum_out[G_total] = tabs[t][s].sum;
                count_out[G_total] = tabs[t][s].cnt; min_out[G_total] = tabs[t][s].mn;
                max_out[G_total] = tabs[t][s].mx;
            }
            G_total++;
        }
        free(tabs[t]);
    }
    free(tabs); free(caps); free(cnt); free(off); free(rows); free(pstart);
    return G_total;
}

/* ------------------------------------------------------------------------- */
/* .map(row => Feature) -- CF:151-160 (season scores), CF:248-265 (players):   */
/*   avg = sum.toFloat / count.toFloat   (JVM l2f, l2f, fdiv -- all RNE)       */
/*   sum_f = sum (Long -> Float widening, CF:159,263)                          */
/* C: (float)int64 and float/float are RNE under the default rounding mode.   */
/* ------------------------------------------------------------------------- */
void orc_finalize_avg_i64(const int64_t *sum, const int64_t *count, int64_t G, float *avg_out,
                          float *sum_out) {
    for (int64_t g = 0; g < G; g++) {
        volatile float s = (float)sum[g];
        volatile float c = (float)count[g];
        if (avg_out) avg_out[g] = s / c;
        if (sum_out) sum_out[g] = s;
    }
}

/* CF:209-218 (attendance), CF:250-264 (rating, worth): sum.toFloat is JVM d2f. */
void orc_finalize_avg_f64(const double *sum, const int64_t *count, int64_t G, float *avg_out,
                          float *sum_out) {
    for (int64_t g = 0; g < G; g++) {
        volatile float s = (float)sum[g];
        volatile float c = (float)count[g];
        if (avg_out) avg_out[g] = s / c;
        if (sum_out) sum_out[g] = s;
    }
}

/* CF:292: team_budget.toFloat */
void orc_d2f(const double *x, int64_t n, float *out) {
    for (int64_t i = 0; i < n; i++) out[i] = (float)x[i];
}

/* ------------------------------------------------------------------------- */
/* sum.join(count, "team_id") -- CF:150,208,247; createTrainingDataset joins   */
/* CF:317-319.  Inner equi-join USING(key): null keys never match; every       */
/* (left,right) pair with equal keys is emitted (M:N).  Pinned by G3/G4/G5.    */
/* Output: index pairs, left-major order.  Returns the pair count (may exceed  */
/* cap, in which case only the first cap pairs were written).                  */
/* ------------------------------------------------------------------------- */
int64_t orc_inner_join(int key_dtype, const void *lkeys, const uint8_t *lvalid, int64_t nl,
                       const void *rkeys, const uint8_t *rvalid, int64_t nr, int64_t *lidx_out,
                       int64_t *ridx_out, int64_t cap) {
    /* build on the right side: chained hash table head[]/next[] */
    int64_t hcap = 16;
    while (hcap < 2 * nr) hcap <<= 1;
    int64_t *head = (int64_t *)malloc(sizeof(int64_t) * (size_t)hcap);
    int64_t *next = (int64_t *)malloc(sizeof(int64_t) * (size_t)(nr > 0 ? nr : 1));
    for (int64_t i = 0; i < hcap; i++) head[i] = -1;
#define KEY_AT(p, i) (key_dtype == ORC_I32 ? (int64_t)((const int32_t *)(p))[i] : ((const int64_t *)(p))[i])
    for (int64_t j = nr - 1; j >= 0; j--) { /* reverse so chains are in ascending j */
        if (!bit_valid(rvalid, j)) continue;
        uint64_t s = hash64((uint64_t)KEY_AT(rkeys, j)) & (uint64_t)(hcap - 1);
        next[j] = head[s]; head[s] = j;
    }
    int64_t out = 0;
    for (int64_t i = 0; i < nl; i++) {
        if (!bit_valid(lvalid, i)) continue;
        int64_t k = KEY_AT(lkeys, i);
        uint64_t s = hash64((uint64_t)k) & (uint64_t)(hcap - 1);
        for (int64_t j = head[s]; j >= 0; j = next[j]) {
            if (KEY_AT(rkeys, j) == k) {
                if (out < cap) { lidx_out[out] = i; ridx_out[out] = j; }
                out++;
            }
        }
    }
#undef KEY_AT
    free(head); free(next);
    return out;
}

/* ------------------------------------------------------------------------- */
/* Descriptive statistics -- enabled by StatisticsConfig(true, ...) CF:114.    */
/* Parity unpinned except min/max (G2).  Follows Spark 2.4 CentralMomentAgg    */
/* (Welford update, Chan merge); min/max with NaN       */
/* greatest; all columns cast to double first (Deequ).                        */
/* m must be zero-initialised by the caller before the first update.          */
/* ------------------------------------------------------------------------- */
void orc_moments_update(int dtype, const void *x, const uint8_t *valid, int64_t n,
                        orc_moments *m) {
    for (int64_t i = 0; i < n; i++) {
        m->n_rows++;
        if (!bit_valid(valid, i)) continue;
        double v = load_as_double(dtype, x, i);
        int first = m->n_nonnull == 0;
        m->n_nonnull++;
        double delta = v - m->mean;
        double deltaN = delta / (double)m->n_nonnull;
        m->mean += deltaN;
        m->m2 += delta * (delta - deltaN);
        m->sum += v;
        if (first || dbl_lt(v, m->min)) m->min = v;
        if (first || dbl_lt(m->max, v)) m->max = v;
    }
}

void orc_moments_merge(orc_moments *a, const orc_moments *b) {
    if (b->n_nonnull == 0) { a->n_rows += b->n_rows; return; }
    if (a->n_nonnull == 0) { int64_t r = a->n_rows; *a = *b; a->n_rows += r; return; }
    double n1 = (double)a->n_nonnull, n2 = (double)b->n_nonnull, n = n1 + n2;
    double delta = b->mean - a->mean;
    double deltaN = delta / n;
    a->mean = a->mean + deltaN * n2;
    a->m2 = a->m2 + b->m2 + delta * deltaN * n1 * n2;
    a->sum += b->sum;
    if (dbl_lt(b->min, a->min)) a->min = b->min;
    if (dbl_lt(a->max, b->max)) a->max = b->max;
    a->n_nonnull += b->n_nonnull;
    a->n_rows += b->n_rows;
}

/* ------------------------------------------------------------------------- */
/* 20-bin histogram -- second flag of StatisticsConfig, CF:114.  Parity        */
/* unpinned.  Follows Spark DoubleRDDFunctions.histogram(bucketCount), even-   */
/* bucket fast path, literally and in float64:          */
/*   edges e_s = min + (s*(max-min))/steps, s=0..steps-1, then max             */
/*   skip x if isNaN(x) || x < min || x > max                                  */
/*   b = (int)(((x - min) / (max - min)) * count); b = min(b, count-1)         */
/* lo/hi NaN => take them from the data (first Spark job); NaN/Inf bounds are  */
/* an error (-1).  min == max => one bucket [min,min] (returns 1).            */
/* Returns the number of buckets written, or a negative error.                */
/* ------------------------------------------------------------------------- */
int orc_histogram(int dtype, const void *x, const uint8_t *valid, int64_t n, int nbins, double lo,
                  double hi, int64_t *counts_out, double *edges_out) {
    if (nbins < 1) return -2;
    if (isnan(lo) || isnan(hi)) {
        int any = 0;
        double mn = 0, mx = 0;
        for (int64_t i = 0; i < n; i++) {
            if (!bit_valid(valid, i)) continue;
            double v = load_as_double(dtype, x, i);
            if (isnan(v)) return -1; /* StatCounter min/max propagate NaN -> Spark throws */
            if (!any) { mn = mx = v; any = 1; }
            else { if (v < mn) mn = v; if (v > mx) mx = v; }
        }
        if (!any) return -3;
        if (isnan(lo)) lo = mn;
        if (isnan(hi)) hi = mx;
    }
    if (isnan(lo) || isnan(hi) || isinf(lo) || isinf(hi) || lo > hi) return -1;
    for (int b = 0; b < nbins; b++) counts_out[b] = 0;
    if (lo == hi) {
        int64_t c = 0;
        for (int64_t i = 0; i < n; i++) {
            if (!bit_valid(valid, i)) continue;
            double v = load_as_double(dtype, x, i);
            if (v == lo) c++;
        }
        counts_out[0] = c;
        if (edges_out) { edges_out[0] = lo; edges_out[1] = lo; }
        return 1;
    }
    double span = hi - lo;
    if (edges_out) {
        for (int s = 0; s < nbins; s++) edges_out[s] = lo + ((double)s * span) / (double)nbins;
        edges_out[nbins] = hi;
    }
    for (int64_t i = 0; i < n; i++) {
        if (!bit_valid(valid, i)) continue;
        double e = load_as_double(dtype, x, i);
        if (isnan(e) || e < lo || e > hi) continue;
        double q = (e - lo) / (hi - lo);
        int b = (int)(q * (double)nbins);
        if (b > nbins || b < 0) continue;
        if (b > nbins - 1) b = nbins - 1;
        counts_out[b]++;
    }
    return nbins;
}

/* ------------------------------------------------------------------------- */
/* Pearson correlation matrix -- third flag of StatisticsConfig, CF:114.       */
/* Parity unpinned.  Follows MLlib Correlation.corr -> RowMatrix.computeCo-    */
/* variance (dense path: mean first, then the centred Gram / (m-1)) ->         */
/* PearsonCorrelation.computeCorrelationMatrixFromCovariance (SURVEY.md App.   */
/* A.4.3): sigma_i = sqrt(cov_ii), cov_ii treated as 0 when |cov_ii| <= 1e-12; */
/* corr_ij = cov_ij / (sigma_j * sigma_i), NaN if either sigma is 0; diag 1.0. */
/* No nulls (all BASELINE.json configs).  corr_out is d*d row-major.          */
/* Returns 0, or -1 when n <= 1 (MLlib requires more than one row).           */
/* ------------------------------------------------------------------------- */
static void corr_from_cov(double *cov, int d) {
    for (int i = 0; i < d; i++) {
        double c = cov[(size_t)i * d + i];
        cov[(size_t)i * d + i] = fabs(c) <= 1e-12 ? 0.0 : sqrt(c);
    }
    for (int j = 0; j < d; j++) {
        double sigma = cov[(size_t)j * d + j];
        for (int i = 0; i < j; i++) {
            double si = cov[(size_t)i * d + i];
            double corr = (sigma == 0.0 || si == 0.0) ? NAN : cov[(size_t)i * d + j] / (sigma * si);
            cov[(size_t)i * d + j] = corr;
            cov[(size_t)j * d + i] = corr;
        }
    }
    for (int i = 0; i < d; i++) cov[(size_t)i * d + i] = 1.0;
}

int orc_pearson(int dtype, const void *const *cols, int d, int64_t n, double *corr_out) {
    if (n <= 1) return -1;
    double *mean = (double *)calloc((size_t)d, sizeof(double));
    for (int j = 0; j < d; j++) {
        double m = 0.0;
        for (int64_t i = 0; i < n; i++) m += (load_as_double(dtype, cols[j], i) - m) / (double)(i + 1);
        mean[j] = m;
    }
    memset(corr_out, 0, sizeof(double) * (size_t)d * (size_t)d);
    /* centred Gram, upper triangle, row blocks for cache friendliness */
    enum { BLK = 1024 };
    double *buf = (double *)malloc(sizeof(double) * (size_t)d * BLK);
    for (int64_t i0 = 0; i0 < n; i0 += BLK) {
        int64_t len = n - i0 < BLK ? n - i0 : BLK;
        for (int j = 0; j < d; j++)
            for (int64_t i = 0; i < len; i++)
                buf[(size_t)j * BLK + i] = load_as_double(dtype, cols[j], i0 + i) - mean[j];
        for (int a = 0; a < d; a++)
            for (int b = a; b < d; b++) {
                double acc = 0.0;
                const double *pa = buf + (size_t)a * BLK, *pb = buf + (size_t)b * BLK;
                for (int64_t i = 0; i < len; i++) acc += pa[i] * pb[i];
                corr_out[(size_t)a * d + b] += acc;
            }
    }
    for (int a = 0; a < d; a++)
        for (int b = a; b < d; b++) {
            double c = corr_out[(size_t)a * d + b] / (double)(n - 1);
            corr_out[(size_t)a * d + b] = c;
            corr_out[(size_t)b * d + a] = c;
        }
    corr_from_cov(corr_out, d);
    free(buf); free(mean);
    return 0;
}

/*
 * The raw-Gram formulation (MLlib's older/sparse path):
 *   cov = (G - S S^T / n) / (n - 1)  with G = sum (x-c)(x-c)^T, S = sum (x-c)
 * (covariance is shift invariant, so any per-column constant c may be used).
 * Used by the CPU tests of the multi-rank merge (partial G and S add up).
 */
int orc_pearson_from_gram(const double *gram, const double *colsum, const double *shift, int d,
                          int64_t n, double *corr_out) {
    (void)shift;
    if (n <= 1) return -1;
    for (int a = 0; a < d; a++)
        for (int b = 0; b < d; b++)
            corr_out[(size_t)a * d + b] =
                (gram[(size_t)a * d + b] - colsum[a] * colsum[b] / (double)n) / (double)(n - 1);
    corr_from_cov(corr_out, d);
    return 0;
}

/*
 * Multithreaded baseline for BASELINE.json config 3 (columnar float32 table):
 * describe + nbins histogram + Pearson, one thread team over column / column-pair
 * work items.  Stand-in for the Spark/Deequ statistics passes.
 */
void orc_stats_f32_mt(const float *const *cols, int d, int64_t n, int nbins, int threads,
                      orc_moments *mom_out, int64_t *hist_out, double *corr_out) {
    if (threads < 1) threads = 1;
#pragma omp parallel for num_threads(threads) schedule(dynamic, 1)
    for (int j = 0; j < d; j++) {
        orc_moments m;
        memset(&m, 0, sizeof m);
        orc_moments_update(ORC_F32, cols[j], NULL, n, &m);
        mom_out[j] = m;
        orc_histogram(ORC_F32, cols[j], NULL, n, nbins, m.min, m.max, hist_out + (size_t)j * nbins,
                      NULL);
    }
    if (!corr_out || n <= 1) return;
    double *cov = corr_out;
#pragma omp parallel for num_threads(threads) schedule(dynamic, 1) collapse(1)
    for (int a = 0; a < d; a++) {
        for (int b = a; b < d; b++) {
            double acc = 0.0, ma = mom_out[a].mean, mb = mom_out[b].mean;
            const float *pa = cols[a], *pb = cols[b];
            for (int64_t i = 0; i < n; i++) acc += ((double)pa[i] - ma) * ((double)pb[i] - mb);
            cov[(size_t)a * d + b] = acc / (double)(n - 1);
            cov[(size_t)b * d + a] = cov[(size_t)a * d + b];
        }
    }
    corr_from_cov(cov, d);
}

/* ------------------------------------------------------------------------- */
/* Synthetic generators.  Counter based, integer mixing plus    */
/* exactly-rounded IEEE adds/multiplies only, so the CUDA generator in         */
/* csrc/generate.cu produces bit-identical columns (no libm calls).            */
/* ------------------------------------------------------------------------- */
uint64_t orc_mix64(uint64_t seed, uint64_t idx) {
    uint64_t z = seed + (idx + 1) * 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

void orc_gen_uniform_i32(uint64_t seed, int64_t start, int64_t n, int32_t lo, int32_t hi_excl,
                         int32_t *out) {
    uint64_t range = (uint64_t)((int64_t)hi_excl - (int64_t)lo);
    for (int64_t i = 0; i < n; i++) {
        uint64_t r = orc_mix64(seed, (uint64_t)(start + i)) >> 32;
        out[i] = (int32_t)((int64_t)lo + (int64_t)((r * range) >> 32));
    }
}

/* unit-variance bell curve: (u1+u2+u3+u4 - 2) * sqrt(3), u ~ U[0,1) (Irwin-Hall, 4 terms) */
static inline double normalish_unit(uint64_t seed, uint64_t idx) {
    const double k = 1.0 / 9007199254740992.0; /* 2^-53 */
    double u1 = (double)(orc_mix64(seed, 4 * idx + 0) >> 11) * k;
    double u2 = (double)(orc_mix64(seed, 4 * idx + 1) >> 11) * k;
    double u3 = (double)(orc_mix64(seed, 4 * idx + 2) >> 11) * k;
    double u4 = (double)(orc_mix64(seed, 4 * idx + 3) >> 11) * k;
    double s = (u1 + u2) + (u3 + u4);
    return (s - 2.0) * 1.7320508075688772;
}

void orc_gen_normalish_f64(uint64_t seed, int64_t start, int64_t n, double mean, double scale,
                           double *out) {
    for (int64_t i = 0; i < n; i++)
        out[i] = mean + scale * normalish_unit(seed, (uint64_t)(start + i));
}

void orc_gen_normalish_f32(uint64_t seed, int64_t start, int64_t n, float mean, float scale,
                           float *out) {
    for (int64_t i = 0; i < n; i++)
        out[i] = (float)((double)mean + (double)scale * normalish_unit(seed, (uint64_t)(start + i)));
}

/* index drawn from a discrete distribution given by its cumulative table (config 4's Zipf keys):
 * first k with cdf[k] > u, u = mix(seed, idx) * 2^-53 -- the same search as csrc/ctx.cu */
void orc_gen_from_cdf(uint64_t seed, int64_t start, int64_t n, const double *cdf, int64_t m, int32_t *out) {
    for (int64_t i = 0; i < n; i++) {
        double u = (double)(orc_mix64(seed, (uint64_t)(start + i)) >> 11) * (1.0 / 9007199254740992.0);
        int64_t lo = 0, hi = m - 1;
        while (lo < hi) {
            int64_t mid = (lo + hi) >> 1;
            if (cdf[mid] > u) hi = mid; else lo = mid + 1;
        }
        out[i] = (int32_t)lo;
    }
}
