/*
 * fsx.h -- C ABI of libfsx: the B200-native engine behind the featurestore_tour job.
 *
 * This is the drop-in boundary.  The reference (moritzmeister/hops-examples) reaches
 * its aggregation / join / statistics engines through JVM calls into Spark SQL and
 * hsfs; a maintainer replaces those calls with a thin JNI/Panama binding onto the
 * functions below (see INTEGRATION.md for the Scala `object Native` stub).
 *
 * Citations: "CF:n"  = featurestore_tour/src/main/scala/io/hops/examples/
 *                      featurestore_tour/featuregroups/ComputeFeatures.scala line n
 *            "MAIN:n" = .../featurestore_tour/Main.scala line n
 *
 * Conventions
 *  - plain C types only; every function returns an int32 status (0 = FSX_OK, negative =
 *    error) and never throws.  fsx_last_error() returns the message for the last failing
 *    call on that context (or, for a NULL context, on the calling thread).
 *  - the caller owns every host buffer it passes in or asks results to be written to;
 *    the library owns device memory and the opaque handles until the matching *_free.
 *  - a context is used by one thread at a time.  fsx_ctx_create(devices, 1) binds one device;
 *    fsx_ctx_create(devices, n > 1) lets ONE driver thread run the fsx_dist_* operators on n devices
 *    of this process (the Spark driver calling the engine once per operator, MAIN:37-48);
 *    fsx_comm_init_rank joins a one-device context to the ranks of other processes (one process per
 *    GPU).  Either way the exchange between devices (NCCL over NVLink) happens inside the library.
 *  - there is NO CPU fallback: without a usable CUDA device fsx_ctx_create fails.
 *  - validity bitmaps use Arrow bit order (bit i of byte i/8, LSB first; 1 = valid);
 *    NULL means "no nulls".
 */
#ifndef FSX_H
#define FSX_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FSX_ABI_VERSION 1

typedef enum {
    FSX_OK = 0,
    FSX_E_INVALID = -1,     /* bad argument / NULL handle / dtype mismatch          */
    FSX_E_CUDA = -2,        /* CUDA runtime error (message has the CUDA string)     */
    FSX_E_OOM = -3,         /* device or pinned-host allocation failed              */
    FSX_E_UNSUPPORTED = -4, /* valid request this build does not implement          */
    FSX_E_STATE = -5,       /* context poisoned by an earlier asynchronous error    */
    FSX_E_DATA = -6,        /* data-dependent failure (NaN histogram bounds, n<=1)  */
    FSX_E_NCCL = -7         /* NCCL could not be loaded or a collective failed      */
} fsx_status;

typedef enum { FSX_I32 = 0, FSX_I64 = 1, FSX_F32 = 2, FSX_F64 = 3 } fsx_dtype;

/* aggregate kinds for fsx_groupby_agg; output dtypes follow Spark 2.4 */
typedef enum {
    FSX_AGG_SUM = 0,        /* integral -> i64 (wraps), fractional -> f64; all-null group -> null */
    FSX_AGG_COUNT = 1,      /* non-null values of the column -> i64                             */
    FSX_AGG_COUNT_STAR = 2, /* rows of the group (value column ignored, may be NULL) -> i64      */
    FSX_AGG_MIN = 3,        /* type preserving; nulls skipped; NaN greatest                      */
    FSX_AGG_MAX = 4,
    FSX_AGG_AVG = 5         /* sum(cast double)/count(non-null) -> f64                           */
} fsx_agg;

typedef enum {
    FSX_JOIN_INNER = 0, FSX_JOIN_LEFT_SEMI = 1, FSX_JOIN_LEFT_ANTI = 2,
    /* outer joins (notebooks/featurestore/hsfs/basics/feature_exploration.ipynb:521,541 `join(..., join_type="left")`): a row without a partner is paired with
     * index -1, which fsx_join_gather turns into a null */
    FSX_JOIN_LEFT_OUTER = 3, FSX_JOIN_RIGHT_OUTER = 4, FSX_JOIN_FULL_OUTER = 5
} fsx_join_type;

/* synthetic column kinds for fsx_column_generate (bit-identical to oracle/) */
typedef enum {
    FSX_GEN_UNIFORM_INT = 0, /* p0 = lo, p1 = hi (exclusive); i32/i64                         */
    FSX_GEN_NORMALISH = 1,   /* p0 = mean, p1 = scale; f32/f64; Irwin-Hall(4) bell curve       */
    FSX_GEN_IOTA = 2,        /* p0 = start, p1 = step; any dtype                               */
    FSX_GEN_MIX = 3          /* p0 = mean, p1 = a, p2 = b, seed2: mean + (a*z(seed) + b*z(seed2)) */
} fsx_gen_kind;

typedef struct fsx_ctx fsx_ctx;     /* one device: streams, memory pool, scratch, error state */
typedef struct fsx_col fsx_col;     /* one typed column resident in HBM (+ optional validity)  */
typedef struct fsx_table fsx_table; /* ordered set of equally long columns                     */
typedef struct fsx_join fsx_join;   /* matched (left row, right row) index pairs               */

/* per-column sufficient statistics; stddev is derived on the host:
 * sample = sqrt(m2/(n_nonnull-1)) [Spark describe], population = sqrt(m2/n_nonnull) [Deequ] */
typedef struct {
    int64_t n_rows;
    int64_t n_nonnull;
    double mean;
    double m2;
    double min;
    double max;
    double sum;
} fsx_moments;

/* wall-clock (host) and device (CUDA event) split of the last call, milliseconds */
typedef struct {
    double h2d_ms;
    double kernel_ms;
    double d2h_ms;
    double total_ms;
    double main_kernel_ms;   /* device time of the call's dominant kernel (0 if not instrumented) */
    int64_t kernel_launches; /* kernels of this library launched by the call */
    int64_t h2d_bytes;
    int64_t d2h_bytes;
} fsx_timing;

/* ---- context ------------------------------------------------------------------------
 * Replaces SparkSession.builder()...getOrCreate() (MAIN:37-39) for the native operators.
 * devices/n: the CUDA ordinals to bind (NULL/0 = the current device).  n > 1 creates one context per device
 * and an NCCL communicator over them (ncclCommInitAll); the returned context is member 0, it alone is passed
 * to the fsx_dist_* calls and to fsx_ctx_destroy (which also destroys the other members). */
int32_t fsx_ctx_create(const int32_t *devices, int32_t n, fsx_ctx **out);
int32_t fsx_ctx_destroy(fsx_ctx *ctx); /* replaces spark.close (MAIN:51) */
const char *fsx_last_error(const fsx_ctx *ctx);
int32_t fsx_last_timing(const fsx_ctx *ctx, fsx_timing *out);
int32_t fsx_ctx_synchronize(fsx_ctx *ctx);
int32_t fsx_abi_version(void);
/* total launches of this library's kernels on the context since creation (bench.py reads it) */
int64_t fsx_ctx_kernel_launches(const fsx_ctx *ctx);
/* raw CUDA stream of the context (cudaStream_t as void*) so callers can record events on it */
void *fsx_ctx_stream(const fsx_ctx *ctx);

/* ---- K0 columnar loader ----------------------------------------------------------------
 * Replaces the typed result of spark.read.format("csv")...load(input).as[T]
 * (CF:104-106,145-147,203-205,242-244,289-291): CSV parsing stays on the host, the typed
 * columns are moved into HBM here.  Pinned buffers make the copy a single DMA. */
int32_t fsx_host_alloc(fsx_ctx *ctx, size_t bytes, void **out);
int32_t fsx_host_free(fsx_ctx *ctx, void *p);
int32_t fsx_column_upload(fsx_ctx *ctx, int32_t dtype, int64_t n, const void *data,
                          const uint8_t *validity, fsx_col **out);
/* non-owning view over device memory the caller allocated (torch tensor, peer buffer) */
int32_t fsx_column_wrap_device(fsx_ctx *ctx, int32_t dtype, int64_t n, void *dev_data,
                               void *dev_validity, fsx_col **out);
/* synthetic columns generated in HBM (BASELINE.json configs 2-5 never cross PCIe) */
int32_t fsx_column_generate(fsx_ctx *ctx, int32_t dtype, int64_t n, int32_t kind, uint64_t seed,
                            uint64_t seed2, int64_t start, double p0, double p1, double p2,
                            fsx_col **out);
/* i32 column of indices drawn from a discrete distribution given by its cumulative table `cdf`
 * (f64, non-decreasing, last entry 1.0): out[i] = first k with cdf[k] > u_i, u_i = mix(seed, start+i)
 * in [0,1).  Used for the Zipf-skewed fact keys of BASELINE.json config 4; bit-identical to oracle/. */
int32_t fsx_column_generate_cdf(fsx_ctx *ctx, int64_t n, uint64_t seed, int64_t start,
                                const fsx_col *cdf, fsx_col **out);
/* CSV decode on the GPU: the scan + schema inference + typed conversion that
 * spark.read.format("csv").option("header", true).option("inferSchema", "true").load(input) performs before .as[T]
 * (CF:104,145,203,242,289).  text/nbytes: the file's bytes AFTER the header line (host memory); ncols: the number of
 * header fields (1..64).  types_out[c] = 1 int, 2 bigint, 3 double, 4 string -- Spark 2.4's inference over the non-empty
 * cells (empty = null; an all-null column is a string column).  *out: one column per CSV column with a validity bitmap
 * (i32 / i64 / f64; a string column comes back as the i64 byte offsets of its cells in `text`, for the host to slice).
 * Blank lines are skipped, a trailing \r is dropped.  Doubles are correctly rounded (Eisel-Lemire), i.e. the value
 * Double.parseDouble returns; every cell is bit-identical to the host reader (host/csv.hpp).  FSX_E_UNSUPPORTED (parse on
 * the host instead) for quoted fields, bare carriage returns, and decimal literals of more than 19 digits that sit on a
 * rounding boundary; FSX_E_DATA when a row has the wrong number of fields. */
int32_t fsx_csv_decode(fsx_ctx *ctx, const char *text, int64_t nbytes, int32_t ncols, int32_t *types_out,
                       fsx_table **out, int64_t *nrows_out);
/* Parquet decode on the GPU (fixture notebooks/featurestore/aws/s3/data/telco-delta/part-0000N-….snappy.parquet):
 * the column scan behind spark.read.format("delta" | "parquet").load(path) for flat schemas.  file_bytes/nbytes: one whole
 * .parquet file in host memory.  fsx_parquet_schema (host only, no GPU): "name:type,..." from the footer, type one of
 * boolean int32 int64 int96 float double byte_array fixed_len_byte_array, a trailing '?' marking an OPTIONAL column; and the
 * row count.  fsx_parquet_decode_column: the named column of every row group as one K0 column (validity bitmap when the
 * column is OPTIONAL): int32 / int64 / float / double as they are, int96 (Spark 2.4's timestamps) as i64 microseconds since
 * the epoch -- Spark's TimestampType --, boolean as i32 0 / 1.  Snappy (or no) compression, data page v1, PLAIN and
 * PLAIN_DICTIONARY / RLE_DICTIONARY values, RLE definition levels -- what Spark 2.4 / parquet-mr 1.10 writes.
 * FSX_E_UNSUPPORTED (read that column with the host's reader) for nested schemas, byte arrays (see below), data page v2,
 * other codecs or encodings; FSX_E_DATA for a corrupt page; FSX_E_INVALID for a file that is not Parquet. */
int32_t fsx_parquet_schema(const void *file_bytes, int64_t nbytes, char *out, size_t outlen, int64_t *nrows_out);
int32_t fsx_parquet_decode_column(fsx_ctx *ctx, const void *file_bytes, int64_t nbytes, const char *column,
                                  fsx_col **out);
/* The same over several files that hold one table (the part files of a Delta / Parquet directory, in the caller's order):
 * ONE column whose rows run through the files in order.  All pages of all files form one launch, so small part files do not
 * cost a launch each.  The column must have the same physical type and repetition in every file. */
int32_t fsx_parquet_decode_column_files(fsx_ctx *ctx, const void *const *files, const int64_t *nbytes, int32_t nfiles,
                                        const char *column, fsx_col **out);
/* A dictionary-encoded BYTE_ARRAY (string) column of the same files: *codes_out = an i32 column of codes (validity when the
 * column is OPTIONAL), code c naming the c-th distinct string of the file -- the chunks' dictionary pages are read on the
 * host (a few KB each) and merged in order of first appearance, each chunk's dictionary then is a table of codes that the
 * device gathers from.  The dictionary comes back as concatenated UTF-8 in dict_bytes with dict_offsets[0..ndistinct]
 * (dict_cap must be > ndistinct for the closing offset).  *ndistinct_out and *dict_bytes_needed are always set: when a
 * buffer is too small the call fails with FSX_E_INVALID and the caller retries with those sizes (NULL buffers and zero
 * capacities ask for the sizes only).  FSX_E_UNSUPPORTED when a chunk stores the strings PLAIN (the writer's fallback for
 * high-cardinality columns): read such a column on the host.  The codes feed fsx_groupby_agg / fsx_hash_join /
 * fsx_stats_value_counts like the codes of fsx_dictionary_encode_utf8. */
int32_t fsx_parquet_decode_strings(fsx_ctx *ctx, const void *file_bytes, int64_t nbytes, const char *column,
                                   fsx_col **codes_out, char *dict_bytes, int64_t dict_bytes_cap, int64_t *dict_offsets,
                                   int32_t dict_cap, int32_t *ndistinct_out, int64_t *dict_bytes_needed);
/* ... over several files: one code space and one dictionary for the whole table (see fsx_parquet_decode_column_files). */
int32_t fsx_parquet_decode_strings_files(fsx_ctx *ctx, const void *const *files, const int64_t *nbytes, int32_t nfiles,
                                         const char *column, fsx_col **codes_out, char *dict_bytes,
                                         int64_t dict_bytes_cap, int64_t *dict_offsets, int32_t dict_cap,
                                         int32_t *ndistinct_out, int64_t *dict_bytes_needed);
int32_t fsx_column_free(fsx_ctx *ctx, fsx_col *col);
int32_t fsx_column_info(const fsx_col *col, int32_t *dtype, int64_t *n, int32_t *has_validity);
int32_t fsx_column_null_count(fsx_ctx *ctx, const fsx_col *col, int64_t *out);
int32_t fsx_column_download(fsx_ctx *ctx, const fsx_col *col, void *data_dst,
                            uint8_t *validity_dst);
int32_t fsx_column_device_ptr(const fsx_col *col, void **data, void **validity);
/* slice [offset, offset+n) as a non-owning view (offset must be a multiple of 8 if the
 * column has a validity bitmap) */
int32_t fsx_column_slice(fsx_ctx *ctx, const fsx_col *col, int64_t offset, int64_t n,
                         fsx_col **out);

/* ---- K1 hash group-by ------------------------------------------------------------------
 * Replaces rawDs.groupBy("team_id").sum() + .groupBy("team_id").count() + the 1:1 join of
 * the two (CF:148-150,206-208,245-247) in ONE pass: output column 0 is the key, then one
 * column per (vals[i], aggs[i]) request.  A null key forms its own group (validity bit of
 * the key column cleared).  Row order of the result is unspecified (as in Spark).
 * key: i32 or i64.  expected_groups: capacity hint (<=0 = derive from the data). */
int32_t fsx_groupby_agg(fsx_ctx *ctx, const fsx_col *key, const fsx_col *const *vals,
                        const int32_t *aggs, int32_t nvals, int64_t expected_groups,
                        fsx_table **out);
/* merge partial group-by states produced on other ranks (rows = (key, state columns)):
 * same request list, inputs are the tables' columns; SUM/COUNT add, MIN/MAX fold.
 * Used by the multi-rank path after the all-to-all of partial states. */
int32_t fsx_groupby_merge(fsx_ctx *ctx, const fsx_col *key, const fsx_col *const *states,
                          const int32_t *aggs, int32_t nvals, int64_t expected_groups,
                          fsx_table **out);

/* Dense-domain form of K1 for the multi-rank merge (no counterpart call in the reference: Spark does
 * this with Exchange hashpartitioning + a final HashAggregate, CF:148-150).  When the keys of all ranks
 * lie in one agreed range [kmin, kmin + nslots), every rank accumulates
 *     sum(val), count(*), min(val), max(val)  GROUP BY key          (null-free i32/i64 key, any value dtype)
 * into caller-owned device arrays with one common slot layout: sum[nslots] f64, cnt[nslots] i64,
 * mm[2*nslots] i64 ({min, max} interleaved, encoded so that a signed MIN merges both).  Adding sum/cnt and
 * taking the minimum of mm across ranks (NCCL reduce-scatter: SUM, SUM, MIN) merges the partial states and
 * leaves every rank with one contiguous, fully merged slice of the key domain, which
 * fsx_groupby_dense_emit turns into the result table (key, sum, avg, count, min, max; sorted by key).
 * fsx_column_minmax gives the local key range to agree on. */
int32_t fsx_column_minmax(fsx_ctx *ctx, const fsx_col *col, int64_t *min_out, int64_t *max_out,
                          int64_t *nvalid_out);
int32_t fsx_groupby_dense_init(fsx_ctx *ctx, int64_t nslots, double *sum_dev, int64_t *cnt_dev,
                               int64_t *mm_dev);
int32_t fsx_groupby_dense_accumulate(fsx_ctx *ctx, const fsx_col *key, const fsx_col *val,
                                     int64_t kmin, int64_t nslots, double *sum_dev, int64_t *cnt_dev,
                                     int64_t *mm_dev);
int32_t fsx_groupby_dense_emit(fsx_ctx *ctx, int32_t key_dtype, int32_t val_dtype, int64_t kmin,
                               int64_t nslots, const double *sum_dev, const int64_t *cnt_dev,
                               const int64_t *mm_dev, fsx_table **out);
/* fsx_groupby_dense_accumulate for a batch in pinned HOST memory (float/double values): chunked copies on the
 * copy stream overlap the accumulation (see fsx_groupby_host).  FSX_E_DATA when a key lies outside the range;
 * the state is then to be discarded. */
int32_t fsx_groupby_dense_accumulate_host(fsx_ctx *ctx, int32_t key_dtype, const void *keys_host,
                                          int32_t val_dtype, const void *vals_host, int64_t n,
                                          int64_t chunk_rows, int64_t kmin, int64_t nslots,
                                          double *sum_dev, int64_t *cnt_dev, int64_t *mm_dev);

/* The whole groupBy(key).agg(sum, avg, count(*), min, max) of ComputeFeatures.scala:148-150 for a batch
 * that lives in HOST memory (what the JVM hands over: off-heap column buffers), in one call: the batch is
 * copied in chunks of chunk_rows rows (0 = 8 Mi) on the context's copy stream while the chunks that have
 * landed are accumulated into dense slot arrays on the compute stream, so the call costs about the PCIe
 * transfer plus one chunk's kernel instead of transfer + kernel.  The slot range follows the key range of
 * the chunks seen so far.  Result: the same table as fsx_groupby_agg with those five aggregates
 * (key, sum, avg, count, min, max; rows sorted by key), on the device.
 * FSX_E_UNSUPPORTED (nothing computed; take fsx_column_upload + fsx_groupby_agg instead) for: pageable host
 * memory, integer values (their SUM is a wrapping int64), key ranges wider than max(4 Mi, min(8 n, 128 Mi)).
 * fsx_last_timing: h2d_ms is the copy stream's span and OVERLAPS kernel_ms for this call; main_kernel_ms
 * sums the accumulate kernels of every 4th chunk and the last. */
int32_t fsx_groupby_host(fsx_ctx *ctx, int32_t key_dtype, const void *keys_host, int32_t val_dtype,
                         const void *vals_host, int64_t n, int64_t chunk_rows, fsx_table **out);

int32_t fsx_table_ncols(const fsx_table *t);
int64_t fsx_table_nrows(const fsx_table *t);
/* borrowed handle, valid until fsx_table_free */
const fsx_col *fsx_table_column(const fsx_table *t, int32_t i);
int32_t fsx_table_download(fsx_ctx *ctx, const fsx_table *t, int32_t i, void *data_dst,
                           uint8_t *validity_dst);
int32_t fsx_table_free(fsx_ctx *ctx, fsx_table *t);
/* sort all columns of a table by its column `by` ascending (nulls last); convenience for
 * deterministic output files, never on the timed path.  The table's columns are REPLACED by their sorted copies: handles
 * taken with fsx_table_column before the call are dead afterwards -- take them again. */
int32_t fsx_table_sort_by(fsx_ctx *ctx, fsx_table *t, int32_t by);

/* ---- group-by for callers that must not allocate --------------------------------------------
 * No counterpart in the reference: this entry point exists for hosts of this library that cannot
 * let it allocate device memory -- e.g. an executor plug-in or pipeline stage working out of a
 * pre-sized device arena that already owns the HBM it was given, or a latency-critical loop that
 * cannot tolerate cudaMalloc's device-wide synchronisation.  It computes
 *     SELECT key, sum(val), count(*) GROUP BY key      (int32 keys, float64 values, no nulls)
 * entirely inside caller-provided device buffers plus one fixed table that lives in a static
 * __device__ array of the library (3 MiB + 256 B, held by every GPU that loads libfsx.so until
 * the process exits).  It calls no cudaMalloc / cudaFree / allocating routine, needs no fsx_ctx,
 * and only enqueues work on `stream` (a cudaStream_t; NULL = default stream) of the CURRENT device.
 *  - limit: at most fsx_noalloc_max_groups() (= 65536) distinct keys; max_groups beyond that is
 *    rejected with FSX_E_INVALID.  keys_out/sum_out/count_out must hold max_groups elements.
 *  - result: *ngroups_out_dev (device memory) = number of groups G, rows [0,G) of the outputs are
 *    filled in an unspecified order; a NEGATIVE value means the data held more than max_groups
 *    distinct keys and the outputs are incomplete.  Read it after synchronising `stream`.
 *  - the table is shared by all calls on a GPU and keeps the previous call's contents; every call
 *    clears it first.  Calls that overlap on the same GPU (other streams or threads) race: serialise.
 *  - the first call on each device looks up the table's device address (cudaGetSymbolAddress);
 *    make that first call before starting any stream capture.
 * Errors are reported through fsx_last_error(NULL). */
int32_t fsx_noalloc_max_groups(void);
int32_t fsx_groupby_sum_count_noalloc(const int32_t *keys_dev, const double *vals_dev, int64_t n,
                                      int64_t max_groups, int32_t *keys_out_dev,
                                      double *sum_out_dev, int64_t *count_out_dev,
                                      int64_t *ngroups_out_dev, void *stream);

/* ---- K2 typed-map finalisation ----------------------------------------------------------
 * Replaces rawFeaturesDf.map(row => ...) (CF:151-160,209-218,248-265) in exact JVM float32
 * semantics: out = sum.toFloat / count.toFloat when count != NULL (l2f or d2f, l2f, fdiv --
 * all round-to-nearest-even), out = sum.toFloat otherwise (CF:159,217,262-264; CF:292 for
 * team_budget.toFloat).  A null sum unboxes to 0. sum: i64/f64/i32/f32. */
int32_t fsx_finalize_f32(fsx_ctx *ctx, const fsx_col *sum, const fsx_col *count, fsx_col **out);
/* Spark's Average on merged states: out = (double)sum / (double)count in float64; null where the
 * sum is null or the count is 0 (needed for BASELINE.json config 2's avg after the
 * multi-rank merge of SUM and COUNT states). */
int32_t fsx_finalize_f64(fsx_ctx *ctx, const fsx_col *sum, const fsx_col *count, fsx_col **out);

/* ---- K3 build/probe hash join -----------------------------------------------------------
 * Replaces sum.join(count, "team_id") (CF:150,208,247) and the feature-group joins of
 * createTrainingDataset (CF:317-319): equi-join USING(key); null keys never match.
 * The right side is the build side.  INNER yields all matching (left,right) pairs (M:N);
 * LEFT_SEMI / LEFT_ANTI yield left rows with / without a match (right index = -1). */
int32_t fsx_hash_join(fsx_ctx *ctx, const fsx_col *left_key, const fsx_col *right_key,
                      int32_t join_type, fsx_join **out);
int32_t fsx_join_count(const fsx_join *j, int64_t *out);
/* side 0 = left, 1 = right: gather `col` through the matched row indices into a new column */
int32_t fsx_join_gather(fsx_ctx *ctx, const fsx_join *j, int32_t side, const fsx_col *col,
                        fsx_col **out);
int32_t fsx_join_indices(fsx_ctx *ctx, const fsx_join *j, int64_t *left_idx_dst,
                         int64_t *right_idx_dst);
int32_t fsx_join_free(fsx_ctx *ctx, fsx_join *j);
/* fused probe + aggregate for the star-schema shape of BASELINE.json config 4:
 * SELECT dim_group, sum(fact_val), count(*) FROM fact JOIN dim ON fact_key = dim_key
 * GROUP BY dim_group -- without materialising the join.  Output columns: group, sum, count. */
int32_t fsx_join_groupby_sum(fsx_ctx *ctx, const fsx_col *fact_key, const fsx_col *fact_val,
                             const fsx_col *dim_key, const fsx_col *dim_group,
                             int64_t expected_groups, fsx_table **out);

/* ---- K4/K5/K7/K8 feature-group statistics ------------------------------------------------
 * Replace what hsfs computes on fg.save(df) when the feature group is built with
 * .statisticsConfig(new StatisticsConfig(true, true, true)) (CF:114,129,170,226,273,303,325):
 * descriptive statistics, histograms, correlations.  cols: ncols equally long columns. */
int32_t fsx_stats_describe(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols,
                           fsx_moments *out /* [ncols] */);
/* Spark DoubleRDDFunctions.histogram(nbins) semantics.  lo/hi: per
 * column bounds or NULL / NaN entries = take the column's min/max.  counts_out[ncols*nbins],
 * edges_out[ncols*(nbins+1)] (may be NULL), nbuckets_out[ncols] (may be NULL; 1 when
 * min == max, else nbins). */
int32_t fsx_stats_histogram(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols, int32_t nbins,
                            const double *lo, const double *hi, int64_t *counts_out,
                            double *edges_out, int32_t *nbuckets_out);
/* shifted Gram: G = sum (x-c)(x-c)^T (d*d row-major f64), S = sum (x-c) (d), n rows.
 * shift: d constants or NULL (the library picks the first row).  shift_out may be NULL.
 * No nulls allowed (FSX_E_UNSUPPORTED otherwise). */
int32_t fsx_stats_gram(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols,
                       const double *shift, double *gram_out, double *colsum_out,
                       double *shift_out, int64_t *n_out);
/* MLlib PearsonCorrelation semantics from (G, S, n): runs on the device. */
int32_t fsx_stats_pearson_from_gram(fsx_ctx *ctx, const double *gram, const double *colsum,
                                    int32_t d, int64_t n, double *corr_out);
/* gram + pearson in one call: corr_out[d*d].  Wide float tables (more than 128 columns, at least 2^20 rows) need no describe
 * pass: centre and scale of the tensor-core digits come from a sample of ~16 K rows per column, the conversion kernel
 * verifies every value against them and sums q = rint((x - c) 2^s) exactly (a miss repeats the run with exact bounds). */
int32_t fsx_stats_pearson(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols,
                          double *corr_out);

/* Everything hsfs computes on save with StatisticsConfig(true, true, true) (CF:114) in one call: the
 * describe pass is shared (its min/max are the histogram bounds, its mean/min/max scale the tensor-core
 * Gram).  counts_out/edges_out/nbuckets_out may be NULL (no histograms), corr_out may be NULL (no
 * correlations).  A column without usable bounds (no non-null value, NaN or infinite extremes) gets
 * zero counts and nbuckets 0.  Correlations need null-free columns and more than one row. */
int32_t fsx_stats_profile(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols, int32_t nbins,
                          fsx_moments *moments_out, int64_t *counts_out, double *edges_out,
                          int32_t *nbuckets_out, double *corr_out);

/* ---- frequency- and rank-based profile metrics ----------------
 * The Deequ column profile hsfs stores next to the descriptive statistics (CF:114) and the data-validation
 * rules built on it (notebooks/featurestore/hsfs/data_validation/feature_validation_python.ipynb:218-243:
 * HAS_NUMBER_OF_DISTINCT_VALUES, HAS_DISTINCTNESS, HAS_UNIQUENESS, HAS_UNIQUE_VALUE_RATIO, HAS_ENTROPY,
 * HAS_APPROX_QUANTILE).  Values are compared by bit pattern, as Spark 2.4's hash aggregate does (0.0 and -0.0
 * are two values); nulls are not values.
 *   distinctness = n_distinct / n_nonnull, uniqueness = n_unique / n_nonnull,
 *   unique value ratio = n_unique / n_distinct, entropy = -sum_v (c_v / n_nonnull) ln(c_v / n_nonnull).
 * The distinct count is exact (Deequ's ApproxCountDistinct is a HyperLogLog++ estimate of this number). */
typedef struct fsx_distinct_stats {
    int64_t n_rows;
    int64_t n_nonnull;
    int64_t n_distinct; /* values that occur at least once */
    int64_t n_unique;   /* values that occur exactly once */
    double entropy;     /* natural logarithm */
} fsx_distinct_stats;
int32_t fsx_stats_distinct(fsx_ctx *ctx, const fsx_col *col, fsx_distinct_stats *out);
/* Deequ's approximateNumDistinctValues (CF:114; ApproxCountDistinct = Spark HyperLogLog++): a HyperLogLog sketch of
 * the non-null values (2^12 registers, relative standard error 1.6 %; values hashed by bit pattern) and Ertl's
 * table-free estimator.  registers_out (may be NULL) receives the 4096 registers: sketches of row shards merge by
 * element-wise maximum, fsx_hll_estimate (host code) turns registers into the estimate.  Spark's own hash (xxHash64) and
 * bias tables are not restated, so the estimate agrees with Deequ's within the estimators' errors, not bit for bit. */
int32_t fsx_stats_approx_distinct(fsx_ctx *ctx, const fsx_col *col, int64_t *estimate_out, uint32_t *registers_out);
double fsx_hll_estimate(const uint32_t *registers, int32_t p);
/* String columns (team_name, team_owner, CF:39-44; the `type` key of the reference's group-by golden,
 * notebooks/hive/PyHive.ipynb:686-689) reach the integer-keyed operators as dictionary codes.  Host code, Arrow-style
 * input (UTF-8 bytes, n + 1 offsets, optional validity bitmap): codes_out[i] = index of row i's string among the
 * distinct strings in order of first appearance, -1 for null (upload with a validity bitmap of codes >= 0);
 * first_row_out[k] (may be NULL, room for n entries) = a row holding the string of code k. */
int32_t fsx_dictionary_encode_utf8(const uint8_t *bytes, const int64_t *offsets, const uint8_t *validity, int64_t n,
                                   int32_t *codes_out, int32_t *first_row_out, int32_t *ndistinct_out);
/* Exact order statistics of the non-null values (NaN greatest, as Spark orders doubles): out[i] is the
 * value of rank max(1, ceil(probs[i] * n_nonnull)) in sorted order, i.e. what DataFrame.approxQuantile
 * returns with relativeError = 0 (Deequ's ApproxQuantile is a sketch of the same quantity).  At most 16
 * probabilities per call; NaN when the column has no non-null value.  Eight histogram passes over the
 * column whatever nprobs is; no sorted copy is made. */
int32_t fsx_stats_quantiles(fsx_ctx *ctx, const fsx_col *col, const double *probs, int32_t nprobs,
                            double *out, int64_t *n_nonnull_out);
/* Exact value counts of a column (Deequ's histogram of a low-cardinality column: value -> count, ratio =
 * count / rows): table (value [the column's dtype; null = the null group], count bigint), most frequent value
 * first, ties by ascending value.  Values are compared by bit pattern as in fsx_stats_distinct. */
int32_t fsx_stats_value_counts(fsx_ctx *ctx, const fsx_col *col, fsx_table **out);

/* ---- the tour's own feature engineering -----------------------
 * notebooks/featurestore/hsfs/basics/feature_engineering.ipynb: composite keys (:229 partitionBy(['store',
 * 'dept']); feature_exploration.ipynb:391 joins on (store, date)), groupBy("store").agg(countDistinct("dept"))
 * (:152-159) and the sums over range windows (:229-249). */
/* One dense int64 key for a pair of integer key columns: (a - min a) * (#b values) + (b - min b); null where
 * either input is null.  Equal pairs get equal keys and the key order is the lexicographic (a, b) order, so the
 * result can be handed to fsx_groupby_agg or fsx_partition_by_key or used as a window partition.  (The offsets are
 * those of THIS pair of columns: for a join, combine the concatenation of both sides' key columns so that both sides
 * share them.)  FSX_E_UNSUPPORTED when the two ranges do not fit 63 bits. */
int32_t fsx_column_combine_keys(fsx_ctx *ctx, const fsx_col *a, const fsx_col *b, fsx_col **out);
/* groupBy(key).agg(countDistinct(val)) for integer key and value columns -> table (key, count bigint): the number
 * of distinct non-null values per key (0 for a key whose values are all null); the null-key group is kept, as
 * Spark keeps it.  Two K1 passes: distinct (key, value) pairs, then pairs per key. */
int32_t fsx_groupby_count_distinct(fsx_ctx *ctx, const fsx_col *key, const fsx_col *val, fsx_table **out);
/* sum(val) OVER (PARTITION BY partition ORDER BY order RANGE BETWEEN lo AND hi): for every row, the sum of the
 * non-null values of the rows in its partition whose order value lies in [order + lo, order + hi] (lo <= hi, both
 * may be negative: rangeBetween(days(-30), days(-1))).  partition may be NULL (one partition); the null partition is
 * one partition.  Result: f64 column in the input's row order, null where the frame holds no non-null value
 * (Spark returns null there; the notebook then does fillna(0)).  NaN/Inf follow IEEE as Spark's sum does. */
int32_t fsx_window_range_sum(fsx_ctx *ctx, const fsx_col *partition, const fsx_col *order, const fsx_col *val,
                             int64_t lo, int64_t hi, fsx_col **out);

/* ---- K6 hash partitioner (multi-rank exchange) --------------------------------------------
 * Replaces Spark's `Exchange hashpartitioning(team_id, 200)` between the partial and final
 * HashAggregate / before the join (implicit in CF:148-150).  Rows are routed to
 * rank = mix(key) % nparts.  counts_out[nparts] (host) receives the rows per destination;
 * the reordered columns come back as a table (column 0 = keys) whose rows are grouped by
 * destination, ready to be handed to an all-to-all with counts_out as the split sizes. */
int32_t fsx_partition_by_key(fsx_ctx *ctx, const fsx_col *key, const fsx_col *const *payload,
                             int32_t npayload, int32_t nparts, int64_t *counts_out,
                             fsx_table **out);

/* ---- multi-GPU operators (SURVEY.md 8(b) "Threading", 8(e)) ---------------------------------
 * Replace the exchange steps that are implicit in the reference's operator calls: the `Exchange
 * hashpartitioning` between partial and final HashAggregate and in front of the joins (CF:148-150, CF:317-319) and the
 * tree aggregation behind the statistics on save (CF:114); the driver thread calls them once per operator exactly as
 * MAIN:37-48 calls the single-node ones.  A "group" is the set of devices that exchange data: the n members of a
 * context made by fsx_ctx_create(devices, n > 1) (world = n, all in this process) or the ranks joined with
 * fsx_comm_init_rank (one member per process).  Every per-shard argument is an array with one entry per LOCAL member
 * (length 1 in the one-process-per-GPU launch); column i of member m sits at [m * ncols + i].  All ranks must make
 * the same call with consistent arguments; data-dependent decisions (route, key range) are agreed inside the call, so
 * no rank can take a different sequence of collectives.  NCCL is loaded with dlopen on first use (the libnccl.so.2
 * already in the process, else the loader's; FSX_NCCL_LIB overrides); FSX_E_NCCL if none is found. */
#define FSX_UNIQUE_ID_BYTES 128
/* rank 0 makes the id (ncclGetUniqueId), the launcher hands it to every rank (torch.distributed store, JVM RPC, file) */
int32_t fsx_comm_unique_id(void *id_out /* FSX_UNIQUE_ID_BYTES */);
int32_t fsx_comm_init_rank(fsx_ctx *ctx, const void *unique_id, int32_t rank, int32_t world);
int32_t fsx_ctx_world(const fsx_ctx *ctx, int32_t *world, int32_t *nlocal, int32_t *first_rank);
/* context of local member i (member 0 is ctx itself): uploads, generators and downloads of that device's shard */
fsx_ctx *fsx_ctx_member(fsx_ctx *ctx, int32_t i);

typedef enum {
    FSX_ROUTE_AUTO = 0,    /* dense when every rank qualifies, else partial states                                  */
    FSX_ROUTE_DENSE = 1,   /* one agreed slot layout, merge = reduce-scatter (SUM, SUM, MIN) in one NCCL group        */
    FSX_ROUTE_PARTIAL = 2, /* partial HashAggregate -> K6 -> all-to-all of the states -> final HashAggregate          */
    FSX_ROUTE_ROWS = 3     /* K6 on the raw rows -> all-to-all -> one HashAggregate on the owner (north_star wording) */
} fsx_dist_route;
typedef enum { FSX_EXCHANGE_SHUFFLE = 0, FSX_EXCHANGE_BROADCAST = 1 } fsx_exchange;

/* groupBy(key).agg(sum, avg, count(*), min, max) of one value column over row shards (CF:148-150 with the aggregates
 * of BASELINE.json config 2).  out[m] = the groups member m's rank owns: (key, sum, avg, count, min, max) -- a
 * contiguous slice of the key range, sorted by key, on the dense route; the keys with mix(key) % world == rank
 * otherwise.  The dense route needs null-free keys, null-free float/double values and a global key range of at most
 * 2^26 slots; the range agreed for one batch is remembered and re-verified by the accumulate kernel itself on the
 * next one.  fsx_dist_last_route tells which route the last call took. */
int32_t fsx_dist_groupby_agg(fsx_ctx *ctx, const fsx_col *const *keys, const fsx_col *const *vals, int32_t route,
                             int64_t expected_groups, fsx_table **out);
int32_t fsx_dist_last_route(const fsx_ctx *ctx);
/* the same for batches in HOST memory (one per local member: keys_host[m], vals_host[m], n[m] rows), the multi-GPU form of
 * fsx_groupby_host: with a key range remembered from an earlier batch the rows stream from pinned host memory straight
 * into the dense state (chunked copies overlapped with the accumulation); otherwise, or when a key falls outside the
 * range or a batch does not qualify on some rank (pageable memory, integer values), the batches are uploaded and take
 * fsx_dist_groupby_agg's routes.  Which of the two happens is agreed by all ranks inside the call. */
int32_t fsx_dist_groupby_host(fsx_ctx *ctx, int32_t key_dtype, const void *const *keys_host, int32_t val_dtype,
                              const void *const *vals_host, const int64_t *n, int64_t expected_groups, fsx_table **out);
/* SELECT g, sum(v), count(*) FROM fact JOIN dim ON fk = pk GROUP BY g over row-sharded fact and dimension tables
 * (BASELINE.json config 4; the join shape of CF:317-319).  SHUFFLE: both sides hash-partitioned by the join key with
 * an all-to-all (Spark's shuffled hash join); BROADCAST: the dimension is all-gathered, fact rows never move (Spark's
 * BroadcastHashJoin).  Every member receives the full result (g, sum, count).  fact_rows_joined[m] (may be NULL) = fact
 * rows member m probed (after the shuffle: the skew). */
int32_t fsx_dist_join_groupby_sum(fsx_ctx *ctx, const fsx_col *const *fact_key, const fsx_col *const *fact_val,
                                  const fsx_col *const *dim_key, const fsx_col *const *dim_group, int32_t variant,
                                  fsx_table **out, int64_t *fact_rows_joined);
/* INNER equi-join of two row-sharded tables (CF:150, CF:317-319).  left/right: nleft / nright columns per member (1..7), column 0
 * is the join key.  out[m] = (key, left payload..., right payload...) for the keys member m owns (SHUFFLE) or for its
 * own left rows (BROADCAST: the right side is all-gathered). */
int32_t fsx_dist_hash_join(fsx_ctx *ctx, const fsx_col *const *left, int32_t nleft, const fsx_col *const *right,
                           int32_t nright, int32_t variant, fsx_table **out);
/* fsx_stats_profile / fsx_stats_histogram over row shards (CF:114): moments all-gathered and merged with Chan's
 * formula in rank order (every rank gets the same bits), histogram counts (bounds = global min/max) and Gram blocks
 * all-reduced on the device.  Outputs are the global results, identical on every rank.  moments_out may be NULL when only
 * corr_out is asked for: wide float tables (more than 128 columns) then skip the describe pass -- centre and scale of the
 * tensor-core digits come from an all-reduced sample and every value is verified by the conversion kernel (a value outside
 * the sampled range, on any rank, repeats the call with exact bounds). */
int32_t fsx_dist_stats_profile(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols, int32_t nbins,
                               fsx_moments *moments_out, int64_t *counts_out, double *edges_out,
                               int32_t *nbuckets_out, double *corr_out);
int32_t fsx_dist_stats_histogram(fsx_ctx *ctx, const fsx_col *const *cols, int32_t ncols, int32_t nbins,
                                 const double *lo, const double *hi, int64_t *counts_out, double *edges_out,
                                 int32_t *nbuckets_out);
/* fsx_stats_quantiles / fsx_stats_approx_distinct of ONE column sharded over the ranks (cols[m] = member m's shard; under
 * fsx_comm_init_rank the one local shard).  Quantiles: the radix selection's 256-bin histograms are summed over the ranks
 * every pass, so the results are the exact order statistics of the whole column (rank = max(1, ceil(p * n_nonnull)) over
 * all shards), the same bits on every rank.  Approximate distinct count: the ranks' 4096 HyperLogLog registers merge by
 * maximum in one all-reduce; registers_out (may be NULL) receives the merged sketch. */
int32_t fsx_dist_stats_quantiles(fsx_ctx *ctx, const fsx_col *const *cols, const double *probs, int32_t nprobs,
                                 double *out, int64_t *n_nonnull_out);
int32_t fsx_dist_stats_approx_distinct(fsx_ctx *ctx, const fsx_col *const *cols, int64_t *estimate_out,
                                       uint32_t *registers_out);
/* host-only helpers (no GPU needed): the rank-ordered Chan merge the statistics use, parts[r * ncols + c] -> out[c],
 * and K6's routing function (which rank owns a key). */
int32_t fsx_moments_merge(const fsx_moments *parts, int32_t nparts, int32_t ncols, fsx_moments *out);
int32_t fsx_partition_owner(int32_t key_dtype, const void *keys, int64_t n, int32_t nparts, int32_t *owner_out);

#ifdef __cplusplus
}
#endif
#endif /* FSX_H */
