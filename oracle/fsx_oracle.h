/*
 * fsx_oracle.h -- CPU restatement of the featurestore_tour hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load it, and only as the checker or the reported CPU baseline.
 *
 * PARITY STATUS: the arithmetic of this path lives in un-vendored dependencies of
 * the reference (org.apache.spark:spark-sql_2.11:2.4.3.2, com.logicalclocks:hsfs:
 * 2.3.0-SNAPSHOT and, transitively, Deequ -- /root/reference/pom.xml:25-28,40-69).
 * The reference has no tests.  The oracle is pinned against the frozen notebook
 * outputs catalogued in SURVEY.md App. B (G1 group-by avg, G2 min/max widening,
 * G3 three-way join, G4 upsert accounting, G5 N:1 lookups); statistics JSON,
 * histograms and correlations are "parity unpinned" (no golden exists anywhere in
 * the reference tree) and follow the published Spark 2.4 algorithms named per
 * function below.
 *
 * Citations "CF:n" = featurestore_tour/src/main/scala/io/hops/examples/
 * featurestore_tour/featuregroups/ComputeFeatures.scala line n.
 */
#ifndef FSX_ORACLE_H
#define FSX_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* dtype codes shared with include/fsx.h */
enum { ORC_I32 = 0, ORC_I64 = 1, ORC_F32 = 2, ORC_F64 = 3 };

typedef struct {
    int64_t n_rows;    /* Deequ Size                                  */
    int64_t n_nonnull; /* Spark describe() "count"                    */
    double mean;       /* running Welford mean                        */
    double m2;         /* sum of squared deviations                   */
    double min;        /* cast to double (Deequ Minimum)              */
    double max;
    double sum;        /* plain left-to-right f64 sum (Deequ Sum)     */
} orc_moments;

/* ---- groupBy (CF:148-149,206-207,245-246) ---- */
int64_t orc_group_ids(int key_dtype, const void *keys, const uint8_t *key_valid, int64_t n,
                      int32_t *gid_out, void *uniq_keys_out, int64_t *null_gid_out);
void orc_group_count(const int32_t *gid, int64_t n, int64_t G, int64_t *count_out);
void orc_group_agg(int val_dtype, const void *vals, const uint8_t *val_valid, const int32_t *gid,
                   int64_t n, int64_t G, double *sum_f64, int64_t *sum_i64, int64_t *nn_out,
                   void *min_out, void *max_out);
/* multithreaded C2-shaped baseline: int32 key, f64 value -> sum,count,min,max */
int64_t orc_groupby_i32_f64_mt(const int32_t *keys, const double *vals, int64_t n, int threads,
                               int32_t *keys_out, double *sum_out, int64_t *count_out,
                               double *min_out, double *max_out, int64_t cap_out);

/* ---- typed map finalisation (CF:151-160,209-218,248-265,292) ---- */
void orc_finalize_avg_i64(const int64_t *sum, const int64_t *count, int64_t G, float *avg_out,
                          float *sum_out);
void orc_finalize_avg_f64(const double *sum, const int64_t *count, int64_t G, float *avg_out,
                          float *sum_out);
void orc_d2f(const double *x, int64_t n, float *out);

/* ---- inner equi-join (CF:150,208,247,317-319) ---- */
int64_t orc_inner_join(int key_dtype, const void *lkeys, const uint8_t *lvalid, int64_t nl,
                       const void *rkeys, const uint8_t *rvalid, int64_t nr, int64_t *lidx_out,
                       int64_t *ridx_out, int64_t cap);

/* ---- statistics on save (CF:114 StatisticsConfig(true,true,true)) ---- */
void orc_moments_update(int dtype, const void *x, const uint8_t *valid, int64_t n,
                        orc_moments *m);
void orc_moments_merge(orc_moments *a, const orc_moments *b);
int orc_histogram(int dtype, const void *x, const uint8_t *valid, int64_t n, int nbins, double lo,
                  double hi, int64_t *counts_out, double *edges_out);
int orc_pearson(int dtype, const void *const *cols, int d, int64_t n, double *corr_out);
int orc_pearson_from_gram(const double *gram, const double *colsum, const double *shift, int d,
                          int64_t n, double *corr_out);
/* multithreaded C3-shaped baseline: columnar f32 -> moments + histogram + pearson */
void orc_stats_f32_mt(const float *const *cols, int d, int64_t n, int nbins, int threads,
                      orc_moments *mom_out, int64_t *hist_out, double *corr_out);

/* ---- deterministic synthetic generators (shared with the CUDA side) ---- */
uint64_t orc_mix64(uint64_t seed, uint64_t idx);
void orc_gen_uniform_i32(uint64_t seed, int64_t start, int64_t n, int32_t lo, int32_t hi_excl,
                         int32_t *out);
void orc_gen_normalish_f64(uint64_t seed, int64_t start, int64_t n, double mean, double scale,
                           double *out);
void orc_gen_normalish_f32(uint64_t seed, int64_t start, int64_t n, float mean, float scale,
                           float *out);

void orc_gen_from_cdf(uint64_t seed, int64_t start, int64_t n, const double *cdf, int64_t m, int32_t *out);

#ifdef __cplusplus
}
#endif
#endif
