/* fsx_jni.c -- mechanical JNI shim: one function per @native method of Native.scala, each forwarding
 * to include/fsx.h and turning a non-zero status into a RuntimeException.  Built only where a JDK is
 * present (not in this repository's image; tests/test_jni_shim_cpu.py compiles it against a minimal stand-in jni.h and
 * checks that every @native method of Native.scala has its function here):
 *   gcc -shared -fPIC -I$JAVA_HOME/include -I$JAVA_HOME/include/linux -I../../include fsx_jni.c \
 *       -L../lib -lfsx -o libfsx_jni.so
 */
#include <jni.h>
#include <stdint.h>

#include "fsx.h"

#define NATIVE(name) Java_io_hops_examples_featurestore_1tour_native_Native_00024_##name

static void throw_fsx(JNIEnv *env, fsx_ctx *ctx) {
    jclass cls = (*env)->FindClass(env, "java/lang/RuntimeException");
    (*env)->ThrowNew(env, cls, fsx_last_error(ctx));
}
static void *buf(JNIEnv *env, jobject b) { return b ? (*env)->GetDirectBufferAddress(env, b) : NULL; }
#define CTX(h) ((fsx_ctx *)(intptr_t)(h))
#define COL(h) ((const fsx_col *)(intptr_t)(h))

JNIEXPORT jlong JNICALL NATIVE(ctxCreate)(JNIEnv *env, jobject self, jint device) {
    fsx_ctx *ctx = NULL;
    int32_t dev = device;
    if (fsx_ctx_create(&dev, 1, &ctx) != FSX_OK) { throw_fsx(env, NULL); return 0; }
    return (jlong)(intptr_t)ctx;
}
JNIEXPORT void JNICALL NATIVE(ctxDestroy)(JNIEnv *env, jobject self, jlong ctx) { fsx_ctx_destroy((fsx_ctx *)(intptr_t)ctx); }

JNIEXPORT jlong JNICALL NATIVE(columnUpload)(JNIEnv *env, jobject self, jlong ctx, jint dtype, jlong n, jobject data, jobject validity) {
    fsx_col *out = NULL;
    if (fsx_column_upload((fsx_ctx *)(intptr_t)ctx, dtype, n, buf(env, data), buf(env, validity), &out) != FSX_OK) throw_fsx(env, (fsx_ctx *)(intptr_t)ctx);
    return (jlong)(intptr_t)out;
}
JNIEXPORT void JNICALL NATIVE(columnDownload)(JNIEnv *env, jobject self, jlong ctx, jlong col, jobject data, jobject validity) {
    if (fsx_column_download((fsx_ctx *)(intptr_t)ctx, (const fsx_col *)(intptr_t)col, buf(env, data), buf(env, validity)) != FSX_OK) throw_fsx(env, (fsx_ctx *)(intptr_t)ctx);
}
JNIEXPORT jlong JNICALL NATIVE(groupbyAgg)(JNIEnv *env, jobject self, jlong ctx, jlong key, jlongArray vals, jintArray aggs, jlong expected) {
    jsize n = (*env)->GetArrayLength(env, vals);
    jlong *v = (*env)->GetLongArrayElements(env, vals, NULL);
    jint *a = (*env)->GetIntArrayElements(env, aggs, NULL);
    fsx_table *out = NULL;
    int32_t rc = fsx_groupby_agg((fsx_ctx *)(intptr_t)ctx, (const fsx_col *)(intptr_t)key, (const fsx_col *const *)v, (const int32_t *)a, n, expected, &out);
    (*env)->ReleaseLongArrayElements(env, vals, v, JNI_ABORT);
    (*env)->ReleaseIntArrayElements(env, aggs, a, JNI_ABORT);
    if (rc != FSX_OK) throw_fsx(env, (fsx_ctx *)(intptr_t)ctx);
    return (jlong)(intptr_t)out;
}
JNIEXPORT jlong JNICALL NATIVE(finalizeF32)(JNIEnv *env, jobject self, jlong ctx, jlong sum, jlong count) {
    fsx_col *out = NULL;
    if (fsx_finalize_f32((fsx_ctx *)(intptr_t)ctx, (const fsx_col *)(intptr_t)sum, (const fsx_col *)(intptr_t)count, &out) != FSX_OK) throw_fsx(env, (fsx_ctx *)(intptr_t)ctx);
    return (jlong)(intptr_t)out;
}
JNIEXPORT jlong JNICALL NATIVE(columnLength)(JNIEnv *env, jobject self, jlong col) {
    int64_t n = 0;
    if (fsx_column_info((const fsx_col *)(intptr_t)col, NULL, &n, NULL) != FSX_OK) throw_fsx(env, NULL);
    return (jlong)n;
}
JNIEXPORT void JNICALL NATIVE(columnFree)(JNIEnv *env, jobject self, jlong ctx, jlong col) { fsx_column_free(CTX(ctx), (fsx_col *)(intptr_t)col); }
JNIEXPORT void JNICALL NATIVE(tableSortBy)(JNIEnv *env, jobject self, jlong ctx, jlong table, jint column) {
    if (fsx_table_sort_by(CTX(ctx), (fsx_table *)(intptr_t)table, column) != FSX_OK) throw_fsx(env, CTX(ctx));
}
JNIEXPORT jlong JNICALL NATIVE(tableNrows)(JNIEnv *env, jobject self, jlong table) { return (jlong)fsx_table_nrows((const fsx_table *)(intptr_t)table); }
JNIEXPORT jlong JNICALL NATIVE(tableColumn)(JNIEnv *env, jobject self, jlong table, jint i) {
    return (jlong)(intptr_t)fsx_table_column((const fsx_table *)(intptr_t)table, i);
}
JNIEXPORT void JNICALL NATIVE(tableFree)(JNIEnv *env, jobject self, jlong ctx, jlong table) { fsx_table_free(CTX(ctx), (fsx_table *)(intptr_t)table); }
JNIEXPORT jlong JNICALL NATIVE(hashJoin)(JNIEnv *env, jobject self, jlong ctx, jlong left, jlong right, jint type) {
    fsx_join *out = NULL;
    if (fsx_hash_join(CTX(ctx), COL(left), COL(right), type, &out) != FSX_OK) throw_fsx(env, CTX(ctx));
    return (jlong)(intptr_t)out;
}
JNIEXPORT jlong JNICALL NATIVE(joinGather)(JNIEnv *env, jobject self, jlong ctx, jlong join, jint side, jlong col) {
    fsx_col *out = NULL;
    if (fsx_join_gather(CTX(ctx), (const fsx_join *)(intptr_t)join, side, COL(col), &out) != FSX_OK) throw_fsx(env, CTX(ctx));
    return (jlong)(intptr_t)out;
}
JNIEXPORT void JNICALL NATIVE(joinFree)(JNIEnv *env, jobject self, jlong ctx, jlong join) { fsx_join_free(CTX(ctx), (fsx_join *)(intptr_t)join); }

/* handles travel as jlong arrays; on LP64 a jlong array IS an array of pointers */
static const fsx_col *const *cols_in(JNIEnv *env, jlongArray a, jlong **keep) {
    *keep = (*env)->GetLongArrayElements(env, a, NULL);
    return (const fsx_col *const *)*keep;
}
static void cols_done(JNIEnv *env, jlongArray a, jlong *keep) { (*env)->ReleaseLongArrayElements(env, a, keep, JNI_ABORT); }

JNIEXPORT void JNICALL NATIVE(statsProfile)(JNIEnv *env, jobject self, jlong ctx, jlongArray cols, jint nbins, jobject moments, jobject counts,
                                            jobject edges, jobject nbuckets, jobject corr) {
    jlong *keep;
    const fsx_col *const *c = cols_in(env, cols, &keep);
    int32_t rc = fsx_stats_profile(CTX(ctx), c, (*env)->GetArrayLength(env, cols), nbins, (fsx_moments *)buf(env, moments), (int64_t *)buf(env, counts),
                                   (double *)buf(env, edges), (int32_t *)buf(env, nbuckets), (double *)buf(env, corr));
    cols_done(env, cols, keep);
    if (rc != FSX_OK) throw_fsx(env, CTX(ctx));
}
JNIEXPORT jlong JNICALL NATIVE(groupbyHost)(JNIEnv *env, jobject self, jlong ctx, jint kdt, jobject keys, jint vdt, jobject vals, jlong n, jlong chunk) {
    fsx_table *out = NULL;
    int32_t rc = fsx_groupby_host(CTX(ctx), kdt, buf(env, keys), vdt, buf(env, vals), n, chunk, &out);
    if (rc == FSX_E_UNSUPPORTED) return 0; /* the caller takes columnUpload + groupbyAgg */
    if (rc != FSX_OK) throw_fsx(env, CTX(ctx));
    return (jlong)(intptr_t)out;
}
JNIEXPORT void JNICALL NATIVE(statsDistinct)(JNIEnv *env, jobject self, jlong ctx, jlong col, jobject out5) {
    if (fsx_stats_distinct(CTX(ctx), COL(col), (fsx_distinct_stats *)buf(env, out5)) != FSX_OK) throw_fsx(env, CTX(ctx));
}
JNIEXPORT jdoubleArray JNICALL NATIVE(statsQuantiles)(JNIEnv *env, jobject self, jlong ctx, jlong col, jdoubleArray probs) {
    jsize n = (*env)->GetArrayLength(env, probs);
    jdouble *p = (*env)->GetDoubleArrayElements(env, probs, NULL);
    double out[16];
    int64_t nn = 0;
    int32_t rc = n <= 16 ? fsx_stats_quantiles(CTX(ctx), COL(col), p, n, out, &nn) : FSX_E_INVALID;
    (*env)->ReleaseDoubleArrayElements(env, probs, p, JNI_ABORT);
    if (rc != FSX_OK) { throw_fsx(env, CTX(ctx)); return NULL; }
    jdoubleArray r = (*env)->NewDoubleArray(env, n);
    (*env)->SetDoubleArrayRegion(env, r, 0, n, out);
    return r;
}
JNIEXPORT jlong JNICALL NATIVE(statsValueCounts)(JNIEnv *env, jobject self, jlong ctx, jlong col) {
    fsx_table *out = NULL;
    if (fsx_stats_value_counts(CTX(ctx), COL(col), &out) != FSX_OK) throw_fsx(env, CTX(ctx));
    return (jlong)(intptr_t)out;
}
JNIEXPORT jlong JNICALL NATIVE(columnCombineKeys)(JNIEnv *env, jobject self, jlong ctx, jlong a, jlong b) {
    fsx_col *out = NULL;
    if (fsx_column_combine_keys(CTX(ctx), COL(a), COL(b), &out) != FSX_OK) throw_fsx(env, CTX(ctx));
    return (jlong)(intptr_t)out;
}
JNIEXPORT jlong JNICALL NATIVE(groupbyCountDistinct)(JNIEnv *env, jobject self, jlong ctx, jlong key, jlong value) {
    fsx_table *out = NULL;
    if (fsx_groupby_count_distinct(CTX(ctx), COL(key), COL(value), &out) != FSX_OK) throw_fsx(env, CTX(ctx));
    return (jlong)(intptr_t)out;
}
JNIEXPORT jlong JNICALL NATIVE(windowRangeSum)(JNIEnv *env, jobject self, jlong ctx, jlong part, jlong order, jlong value, jlong lo, jlong hi) {
    fsx_col *out = NULL;
    if (fsx_window_range_sum(CTX(ctx), COL(part), COL(order), COL(value), lo, hi, &out) != FSX_OK) throw_fsx(env, CTX(ctx));
    return (jlong)(intptr_t)out;
}

/* ---- multi-GPU: one driver thread, n devices (or one JVM per GPU joined with commInitRank) ---- */
static jlongArray handles_out(JNIEnv *env, void *const *h, int n) {
    jlong tmp[64];
    for (int i = 0; i < n; i++) tmp[i] = (jlong)(intptr_t)h[i];
    jlongArray r = (*env)->NewLongArray(env, n);
    (*env)->SetLongArrayRegion(env, r, 0, n, tmp);
    return r;
}
static int nlocal(fsx_ctx *ctx) { int32_t n = 1; fsx_ctx_world(ctx, NULL, &n, NULL); return n; }

JNIEXPORT jlong JNICALL NATIVE(ctxCreateGroup)(JNIEnv *env, jobject self, jintArray devices) {
    jsize n = (*env)->GetArrayLength(env, devices);
    jint *d = (*env)->GetIntArrayElements(env, devices, NULL);
    fsx_ctx *ctx = NULL;
    int32_t rc = fsx_ctx_create((const int32_t *)d, n, &ctx);
    (*env)->ReleaseIntArrayElements(env, devices, d, JNI_ABORT);
    if (rc != FSX_OK) { throw_fsx(env, NULL); return 0; }
    return (jlong)(intptr_t)ctx;
}
JNIEXPORT jlong JNICALL NATIVE(ctxMember)(JNIEnv *env, jobject self, jlong ctx, jint i) { return (jlong)(intptr_t)fsx_ctx_member(CTX(ctx), i); }
JNIEXPORT jbyteArray JNICALL NATIVE(commUniqueId)(JNIEnv *env, jobject self) {
    jbyte id[FSX_UNIQUE_ID_BYTES];
    if (fsx_comm_unique_id(id) != FSX_OK) { throw_fsx(env, NULL); return NULL; }
    jbyteArray r = (*env)->NewByteArray(env, FSX_UNIQUE_ID_BYTES);
    (*env)->SetByteArrayRegion(env, r, 0, FSX_UNIQUE_ID_BYTES, id);
    return r;
}
JNIEXPORT void JNICALL NATIVE(commInitRank)(JNIEnv *env, jobject self, jlong ctx, jbyteArray id, jint rank, jint world) {
    jbyte *b = (*env)->GetByteArrayElements(env, id, NULL);
    int32_t rc = fsx_comm_init_rank(CTX(ctx), b, rank, world);
    (*env)->ReleaseByteArrayElements(env, id, b, JNI_ABORT);
    if (rc != FSX_OK) throw_fsx(env, CTX(ctx));
}
JNIEXPORT jlongArray JNICALL NATIVE(distGroupbyAgg)(JNIEnv *env, jobject self, jlong ctx, jlongArray keys, jlongArray vals, jint route, jlong expected) {
    jlong *kk, *vk;
    const fsx_col *const *k = cols_in(env, keys, &kk), *const *v = cols_in(env, vals, &vk);
    fsx_table *out[64] = {0};
    int32_t rc = fsx_dist_groupby_agg(CTX(ctx), k, v, route, expected, out);
    cols_done(env, keys, kk); cols_done(env, vals, vk);
    if (rc != FSX_OK) { throw_fsx(env, CTX(ctx)); return NULL; }
    return handles_out(env, (void *const *)out, nlocal(CTX(ctx)));
}
JNIEXPORT jlongArray JNICALL NATIVE(distGroupbyHost)(JNIEnv *env, jobject self, jlong ctx, jint kdt, jobjectArray keys, jint vdt, jobjectArray vals,
                                                     jlongArray n, jlong expected) {
    const void *k[64], *v[64];
    int m = nlocal(CTX(ctx));
    for (int i = 0; i < m; i++) {
        k[i] = buf(env, (*env)->GetObjectArrayElement(env, keys, i));
        v[i] = buf(env, (*env)->GetObjectArrayElement(env, vals, i));
    }
    jlong *rows = (*env)->GetLongArrayElements(env, n, NULL);
    fsx_table *out[64] = {0};
    int32_t rc = fsx_dist_groupby_host(CTX(ctx), kdt, k, vdt, v, (const int64_t *)rows, expected, out);
    (*env)->ReleaseLongArrayElements(env, n, rows, JNI_ABORT);
    if (rc != FSX_OK) { throw_fsx(env, CTX(ctx)); return NULL; }
    return handles_out(env, (void *const *)out, m);
}
JNIEXPORT jlongArray JNICALL NATIVE(distJoinGroupbySum)(JNIEnv *env, jobject self, jlong ctx, jlongArray fk, jlongArray fv, jlongArray pk, jlongArray g, jint variant) {
    jlong *k1, *k2, *k3, *k4;
    const fsx_col *const *a = cols_in(env, fk, &k1), *const *b = cols_in(env, fv, &k2), *const *c = cols_in(env, pk, &k3), *const *d = cols_in(env, g, &k4);
    fsx_table *out[64] = {0};
    int32_t rc = fsx_dist_join_groupby_sum(CTX(ctx), a, b, c, d, variant, out, NULL);
    cols_done(env, fk, k1); cols_done(env, fv, k2); cols_done(env, pk, k3); cols_done(env, g, k4);
    if (rc != FSX_OK) { throw_fsx(env, CTX(ctx)); return NULL; }
    return handles_out(env, (void *const *)out, nlocal(CTX(ctx)));
}
JNIEXPORT jlongArray JNICALL NATIVE(distHashJoin)(JNIEnv *env, jobject self, jlong ctx, jlongArray left, jint nleft, jlongArray right, jint nright, jint variant) {
    jlong *k1, *k2;
    const fsx_col *const *l = cols_in(env, left, &k1), *const *r = cols_in(env, right, &k2);
    fsx_table *out[64] = {0};
    int32_t rc = fsx_dist_hash_join(CTX(ctx), l, nleft, r, nright, variant, out);
    cols_done(env, left, k1); cols_done(env, right, k2);
    if (rc != FSX_OK) { throw_fsx(env, CTX(ctx)); return NULL; }
    return handles_out(env, (void *const *)out, nlocal(CTX(ctx)));
}
JNIEXPORT void JNICALL NATIVE(distStatsProfile)(JNIEnv *env, jobject self, jlong ctx, jlongArray cols, jint ncols, jint nbins, jobject moments, jobject counts,
                                                jobject edges, jobject nbuckets, jobject corr) {
    jlong *keep;
    const fsx_col *const *c = cols_in(env, cols, &keep);
    int32_t rc = fsx_dist_stats_profile(CTX(ctx), c, ncols, nbins, (fsx_moments *)buf(env, moments), (int64_t *)buf(env, counts), (double *)buf(env, edges),
                                        (int32_t *)buf(env, nbuckets), (double *)buf(env, corr));
    cols_done(env, cols, keep);
    if (rc != FSX_OK) throw_fsx(env, CTX(ctx));
}

/* ---- scans and string columns ---------------------------------------------------------------------------------------- */
JNIEXPORT jlong JNICALL NATIVE(statsApproxDistinct)(JNIEnv *env, jobject self, jlong ctx, jlong col) {
    int64_t est = 0;
    if (fsx_stats_approx_distinct(CTX(ctx), COL(col), &est, NULL) != FSX_OK) throw_fsx(env, CTX(ctx));
    return (jlong)est;
}
JNIEXPORT jlong JNICALL NATIVE(csvDecode)(JNIEnv *env, jobject self, jlong ctx, jobject body, jlong nbytes, jint ncols, jintArray types) {
    jint *t = (*env)->GetIntArrayElements(env, types, NULL);
    fsx_table *out = NULL;
    int64_t nrows = 0;
    int32_t rc = fsx_csv_decode(CTX(ctx), (const char *)buf(env, body), nbytes, ncols, (int32_t *)t, &out, &nrows);
    (*env)->ReleaseIntArrayElements(env, types, t, 0);
    if (rc != FSX_OK) { throw_fsx(env, CTX(ctx)); return 0; }
    return (jlong)(intptr_t)out;
}
JNIEXPORT jstring JNICALL NATIVE(parquetSchema)(JNIEnv *env, jobject self, jobject file, jlong nbytes) {
    static __thread char text[1 << 16];
    int64_t nrows = 0;
    if (fsx_parquet_schema(buf(env, file), nbytes, text, sizeof text, &nrows) != FSX_OK) { throw_fsx(env, NULL); return NULL; }
    return (*env)->NewStringUTF(env, text);
}
JNIEXPORT jlong JNICALL NATIVE(parquetDecodeColumn)(JNIEnv *env, jobject self, jlong ctx, jobject file, jlong nbytes, jstring column) {
    const char *name = (*env)->GetStringUTFChars(env, column, NULL);
    fsx_col *out = NULL;
    int32_t rc = fsx_parquet_decode_column(CTX(ctx), buf(env, file), nbytes, name, &out);
    (*env)->ReleaseStringUTFChars(env, column, name);
    if (rc != FSX_OK) { throw_fsx(env, CTX(ctx)); return 0; }
    return (jlong)(intptr_t)out;
}
JNIEXPORT jlong JNICALL NATIVE(parquetDecodeColumnFiles)(JNIEnv *env, jobject self, jlong ctx, jobjectArray files, jlongArray nbytes, jstring column) {
    const void *f[256];
    jsize n = (*env)->GetArrayLength(env, files);
    if (n > 256) n = 256;
    for (jsize i = 0; i < n; i++) f[i] = buf(env, (*env)->GetObjectArrayElement(env, files, i));
    jlong *sizes = (*env)->GetLongArrayElements(env, nbytes, NULL);
    const char *name = (*env)->GetStringUTFChars(env, column, NULL);
    fsx_col *out = NULL;
    int32_t rc = fsx_parquet_decode_column_files(CTX(ctx), f, (const int64_t *)sizes, n, name, &out);
    (*env)->ReleaseStringUTFChars(env, column, name);
    (*env)->ReleaseLongArrayElements(env, nbytes, sizes, JNI_ABORT);
    if (rc != FSX_OK) { throw_fsx(env, CTX(ctx)); return 0; }
    return (jlong)(intptr_t)out;
}
JNIEXPORT jlong JNICALL NATIVE(parquetDecodeStrings)(JNIEnv *env, jobject self, jlong ctx, jobject file, jlong nbytes, jstring column, jobject dict_bytes,
                                                     jobject dict_offsets, jlongArray sizes) {
    const char *name = (*env)->GetStringUTFChars(env, column, NULL);
    fsx_col *out = NULL;
    int32_t nd = 0;
    int64_t need = 0;
    const jlong bytes_cap = dict_bytes ? (*env)->GetDirectBufferCapacity(env, dict_bytes) : 0;
    const jlong offsets_cap = dict_offsets ? (*env)->GetDirectBufferCapacity(env, dict_offsets) / 8 : 0;
    int32_t rc = fsx_parquet_decode_strings(CTX(ctx), buf(env, file), nbytes, name, &out, (char *)buf(env, dict_bytes), bytes_cap, (int64_t *)buf(env, dict_offsets),
                                            (int32_t)offsets_cap, &nd, &need);
    (*env)->ReleaseStringUTFChars(env, column, name);
    jlong s[2] = {nd, need};
    (*env)->SetLongArrayRegion(env, sizes, 0, 2, s);
    if (rc == FSX_E_INVALID && (nd >= offsets_cap || need > bytes_cap)) return 0;      /* buffers too small: sizes says how large */
    if (rc != FSX_OK) { throw_fsx(env, CTX(ctx)); return 0; }
    return (jlong)(intptr_t)out;
}

/* ---- statistics of one column sharded over the members ---------------------------------------------------------------- */
JNIEXPORT jdoubleArray JNICALL NATIVE(distStatsQuantiles)(JNIEnv *env, jobject self, jlong ctx, jlongArray cols, jdoubleArray probs) {
    jsize n = (*env)->GetArrayLength(env, probs);
    jdouble *p = (*env)->GetDoubleArrayElements(env, probs, NULL);
    jlong *keep;
    const fsx_col *const *c = cols_in(env, cols, &keep);
    double out[16];
    int64_t nn = 0;
    int32_t rc = n <= 16 ? fsx_dist_stats_quantiles(CTX(ctx), c, p, n, out, &nn) : FSX_E_INVALID;
    cols_done(env, cols, keep);
    (*env)->ReleaseDoubleArrayElements(env, probs, p, JNI_ABORT);
    if (rc != FSX_OK) { throw_fsx(env, CTX(ctx)); return NULL; }
    jdoubleArray r = (*env)->NewDoubleArray(env, n);
    (*env)->SetDoubleArrayRegion(env, r, 0, n, out);
    return r;
}
JNIEXPORT jlong JNICALL NATIVE(distStatsApproxDistinct)(JNIEnv *env, jobject self, jlong ctx, jlongArray cols) {
    jlong *keep;
    const fsx_col *const *c = cols_in(env, cols, &keep);
    int64_t est = 0;
    int32_t rc = fsx_dist_stats_approx_distinct(CTX(ctx), c, &est, NULL);
    cols_done(env, cols, keep);
    if (rc != FSX_OK) throw_fsx(env, CTX(ctx));
    return (jlong)est;
}

/* ---- the remaining single-GPU operators ----------------------------------------------------------------------------- */
JNIEXPORT jlong JNICALL NATIVE(joinGroupbySum)(JNIEnv *env, jobject self, jlong ctx, jlong fact_key, jlong fact_val, jlong dim_key, jlong dim_group,
                                               jlong expected) {
    fsx_table *out = NULL;
    if (fsx_join_groupby_sum(CTX(ctx), COL(fact_key), COL(fact_val), COL(dim_key), COL(dim_group), expected, &out) != FSX_OK) throw_fsx(env, CTX(ctx));
    return (jlong)(intptr_t)out;
}
JNIEXPORT jint JNICALL NATIVE(dictionaryEncodeUtf8)(JNIEnv *env, jobject self, jobject bytes, jobject offsets, jobject validity, jlong n, jobject codes_out,
                                                    jobject first_row_out) {
    int32_t nd = 0;
    if (fsx_dictionary_encode_utf8((const uint8_t *)buf(env, bytes), (const int64_t *)buf(env, offsets), (const uint8_t *)buf(env, validity), n,
                                   (int32_t *)buf(env, codes_out), (int32_t *)buf(env, first_row_out), &nd) != FSX_OK) throw_fsx(env, NULL);
    return nd;
}
JNIEXPORT jdoubleArray JNICALL NATIVE(lastTiming)(JNIEnv *env, jobject self, jlong ctx) {
    fsx_timing t;
    if (fsx_last_timing(CTX(ctx), &t) != FSX_OK) { throw_fsx(env, CTX(ctx)); return NULL; }
    jdouble v[8] = {t.total_ms, t.h2d_ms, t.kernel_ms, t.d2h_ms, t.main_kernel_ms, (jdouble)t.h2d_bytes, (jdouble)t.d2h_bytes, (jdouble)t.kernel_launches};
    jdoubleArray r = (*env)->NewDoubleArray(env, 8);
    (*env)->SetDoubleArrayRegion(env, r, 0, 8, v);
    return r;
}
/* every address is a device address inside the caller's own arena; nothing is allocated on either side of this call */
JNIEXPORT void JNICALL NATIVE(groupbySumCountNoalloc)(JNIEnv *env, jobject self, jlong keys_dev, jlong vals_dev, jlong n, jlong max_groups, jlong keys_out_dev,
                                                      jlong sum_out_dev, jlong count_out_dev, jlong ngroups_out_dev, jlong stream) {
    if (fsx_groupby_sum_count_noalloc((const int32_t *)(intptr_t)keys_dev, (const double *)(intptr_t)vals_dev, n, max_groups, (int32_t *)(intptr_t)keys_out_dev,
                                      (double *)(intptr_t)sum_out_dev, (int64_t *)(intptr_t)count_out_dev, (int64_t *)(intptr_t)ngroups_out_dev,
                                      (void *)(intptr_t)stream) != FSX_OK) throw_fsx(env, NULL);
}
JNIEXPORT jint JNICALL NATIVE(noallocMaxGroups)(JNIEnv *env, jobject self) { return fsx_noalloc_max_groups(); }
