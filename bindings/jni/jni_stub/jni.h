/* Minimal stand-in for the JDK's <jni.h>, ONLY so that `gcc -fsyntax-only` can check bindings/jni/qegpu_jni.c
 * against include/qegpu.h in an image without a JDK.  Never shipped, never linked. */
#ifndef QE_JNI_STUB_H
#define QE_JNI_STUB_H
#include <stdint.h>
typedef int32_t jint; typedef int64_t jlong; typedef uint8_t jboolean; typedef double jdouble; typedef jint jsize;
typedef void *jobject; typedef jobject jclass; typedef jobject jarray;
typedef jarray jintArray; typedef jarray jlongArray; typedef jarray jbooleanArray;
#define JNIEXPORT
#define JNICALL
#define JNI_ABORT 2
#define JNI_TRUE 1
#define JNI_FALSE 0
struct JNINativeInterface_;
typedef const struct JNINativeInterface_ *JNIEnv;
struct JNINativeInterface_ {
    jclass (*FindClass)(JNIEnv *, const char *);
    jint (*ThrowNew)(JNIEnv *, jclass, const char *);
    void *(*GetDirectBufferAddress)(JNIEnv *, jobject);
    jsize (*GetArrayLength)(JNIEnv *, jarray);
    jint *(*GetIntArrayElements)(JNIEnv *, jintArray, jboolean *);
    void (*ReleaseIntArrayElements)(JNIEnv *, jintArray, jint *, jint);
    jlong *(*GetLongArrayElements)(JNIEnv *, jlongArray, jboolean *);
    void (*ReleaseLongArrayElements)(JNIEnv *, jlongArray, jlong *, jint);
    jboolean *(*GetBooleanArrayElements)(JNIEnv *, jbooleanArray, jboolean *);
    void (*ReleaseBooleanArrayElements)(JNIEnv *, jbooleanArray, jboolean *, jint);
    void (*SetBooleanArrayRegion)(JNIEnv *, jbooleanArray, jsize, jsize, const jboolean *);
};
#endif
