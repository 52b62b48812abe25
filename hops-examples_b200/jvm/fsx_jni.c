/* fsx_jni.c -- mechanical JNI shim: one function per @native method of Native.scala, each forwarding
 * to include/fsx.h and turning a non-zero status into a RuntimeException.  Built only where a JDK is
 * present (not in this repository's image):
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
/* ... the remaining methods (tableSortBy, tableNrows, tableColumn, tableFree, columnLength, columnFree,
 * hashJoin, joinGather, joinFree, statsProfile) follow the same pattern: unwrap handles and direct
 * buffers, call the fsx_* function of the same name, throw on a non-zero status. */
