/* Minimal stand-in for <jni.h>: just enough declarations for hops-examples_b200/jvm/fsx_jni.c to be syntax- and
 * type-checked in an image without a JDK (tests/test_jni_shim_cpu.py).  Not a JNI implementation. */
#ifndef MOCK_JNI_H
#define MOCK_JNI_H
#include <stdint.h>
#define JNIEXPORT __attribute__((visibility("default")))
#define JNICALL
#define JNI_ABORT 2
typedef int32_t jint; typedef int64_t jlong; typedef int8_t jbyte; typedef double jdouble; typedef jint jsize;
typedef void *jobject; typedef jobject jclass; typedef jobject jarray; typedef jarray jlongArray; typedef jarray jintArray;
typedef jarray jdoubleArray; typedef jarray jbyteArray; typedef jarray jobjectArray; typedef jobject jstring;
struct JNINativeInterface_;
typedef const struct JNINativeInterface_ *JNIEnv;
struct JNINativeInterface_ {
    jclass (*FindClass)(JNIEnv *, const char *);
    jint (*ThrowNew)(JNIEnv *, jclass, const char *);
    void *(*GetDirectBufferAddress)(JNIEnv *, jobject);
    jsize (*GetArrayLength)(JNIEnv *, jarray);
    jlong *(*GetLongArrayElements)(JNIEnv *, jlongArray, void *);
    void (*ReleaseLongArrayElements)(JNIEnv *, jlongArray, jlong *, jint);
    jint *(*GetIntArrayElements)(JNIEnv *, jintArray, void *);
    void (*ReleaseIntArrayElements)(JNIEnv *, jintArray, jint *, jint);
    jdouble *(*GetDoubleArrayElements)(JNIEnv *, jdoubleArray, void *);
    void (*ReleaseDoubleArrayElements)(JNIEnv *, jdoubleArray, jdouble *, jint);
    jbyte *(*GetByteArrayElements)(JNIEnv *, jbyteArray, void *);
    void (*ReleaseByteArrayElements)(JNIEnv *, jbyteArray, jbyte *, jint);
    jdoubleArray (*NewDoubleArray)(JNIEnv *, jsize);
    void (*SetDoubleArrayRegion)(JNIEnv *, jdoubleArray, jsize, jsize, const jdouble *);
    jlongArray (*NewLongArray)(JNIEnv *, jsize);
    void (*SetLongArrayRegion)(JNIEnv *, jlongArray, jsize, jsize, const jlong *);
    jbyteArray (*NewByteArray)(JNIEnv *, jsize);
    void (*SetByteArrayRegion)(JNIEnv *, jbyteArray, jsize, jsize, const jbyte *);
    jobject (*GetObjectArrayElement)(JNIEnv *, jobjectArray, jsize);
    jstring (*NewStringUTF)(JNIEnv *, const char *);
    const char *(*GetStringUTFChars)(JNIEnv *, jstring, void *);
    void (*ReleaseStringUTFChars)(JNIEnv *, jstring, const char *);
    jlong (*GetDirectBufferCapacity)(JNIEnv *, jobject);
};
#endif
