/*
 * jni_mock.c -- a JNIEnv for executing bindings/jni/qegpu_jni.c without a JVM (there is none in this image).
 * TEST INFRASTRUCTURE ONLY.
 *
 * The function table is the stand-in bindings/jni/jni_stub/jni.h (JNI specification layout); the dozen functions
 * the shim calls are implemented over plain C objects: primitive arrays (length + elements), direct byte buffers
 * (address + capacity), classes (a name) and one pending exception (class name + message) -- the JNI rule that a
 * native method returns normally after ThrowNew and the exception surfaces in the caller is what
 * tests/test_gpu_jni.py checks.  Get / Release<Prim>ArrayElements calls are counted so that a missing release in
 * the shim shows up.
 */
#include <jni.h>
#include <stdlib.h>
#include <string.h>

enum { MOCK_INT_ARRAY = 1, MOCK_LONG_ARRAY, MOCK_BOOLEAN_ARRAY, MOCK_BUFFER, MOCK_CLASS };

typedef struct mock_object {
    int kind;
    jsize length;          /* arrays: elements */
    void *data;            /* arrays: elements (owned); buffers: address (borrowed) */
    jlong capacity;        /* buffers */
    char name[96];         /* classes */
} mock_object;

static __thread char pending_class[96];
static __thread char pending_message[1024];
static __thread int pending;
static __thread int outstanding;      /* Get...ArrayElements not yet released */

static jclass m_FindClass(JNIEnv *env, const char *name)
{
    mock_object *o = (mock_object *)calloc(1, sizeof *o);
    o->kind = MOCK_CLASS;
    strncpy(o->name, name, sizeof o->name - 1);
    return o;                         /* a local reference: leaked on purpose, as a JVM frees it with the frame */
}

static jint m_ThrowNew(JNIEnv *env, jclass cls, const char *msg)
{
    const mock_object *c = (const mock_object *)cls;
    if (!c || c->kind != MOCK_CLASS) return -1;
    strncpy(pending_class, c->name, sizeof pending_class - 1);
    strncpy(pending_message, msg ? msg : "", sizeof pending_message - 1);
    pending = 1;
    return 0;
}

static jobject m_ExceptionOccurred(JNIEnv *env) { return pending ? (jobject)pending_class : NULL; }
static void m_ExceptionClear(JNIEnv *env) { pending = 0; pending_class[0] = 0; pending_message[0] = 0; }
static jboolean m_ExceptionCheck(JNIEnv *env) { return pending ? JNI_TRUE : JNI_FALSE; }

static jsize m_GetArrayLength(JNIEnv *env, jarray a) { return ((const mock_object *)a)->length; }

/* like HotSpot for large arrays: the caller gets a COPY, Release copies back unless JNI_ABORT */
static void *get_elements(jarray a, int kind, size_t width, jboolean *is_copy)
{
    mock_object *o = (mock_object *)a;
    if (!o || o->kind != kind) return NULL;
    void *copy = malloc(width * (size_t)(o->length > 0 ? o->length : 1));
    memcpy(copy, o->data, width * (size_t)o->length);
    if (is_copy) *is_copy = JNI_TRUE;
    outstanding++;
    return copy;
}

static void release_elements(jarray a, void *elems, size_t width, jint mode)
{
    mock_object *o = (mock_object *)a;
    if (mode != JNI_ABORT) memcpy(o->data, elems, width * (size_t)o->length);
    if (mode != JNI_COMMIT) { free(elems); outstanding--; }
}

static jint *m_GetIntArrayElements(JNIEnv *env, jintArray a, jboolean *c) { return (jint *)get_elements(a, MOCK_INT_ARRAY, 4, c); }
static jlong *m_GetLongArrayElements(JNIEnv *env, jlongArray a, jboolean *c) { return (jlong *)get_elements(a, MOCK_LONG_ARRAY, 8, c); }
static jboolean *m_GetBooleanArrayElements(JNIEnv *env, jbooleanArray a, jboolean *c) { return (jboolean *)get_elements(a, MOCK_BOOLEAN_ARRAY, 1, c); }
static void m_ReleaseIntArrayElements(JNIEnv *env, jintArray a, jint *e, jint mode) { release_elements(a, e, 4, mode); }
static void m_ReleaseLongArrayElements(JNIEnv *env, jlongArray a, jlong *e, jint mode) { release_elements(a, e, 8, mode); }
static void m_ReleaseBooleanArrayElements(JNIEnv *env, jbooleanArray a, jboolean *e, jint mode) { release_elements(a, e, 1, mode); }

static void m_SetBooleanArrayRegion(JNIEnv *env, jbooleanArray a, jsize start, jsize len, const jboolean *buf)
{
    mock_object *o = (mock_object *)a;
    if (!o || o->kind != MOCK_BOOLEAN_ARRAY || start < 0 || len < 0 || start + len > o->length) {
        mock_object cls = {MOCK_CLASS, 0, NULL, 0, "java/lang/ArrayIndexOutOfBoundsException"};
        m_ThrowNew(env, &cls, "SetBooleanArrayRegion");
        return;
    }
    memcpy((jboolean *)o->data + start, buf, (size_t)len);
}

static void *m_GetDirectBufferAddress(JNIEnv *env, jobject b)
{
    const mock_object *o = (const mock_object *)b;
    return o && o->kind == MOCK_BUFFER ? o->data : NULL;
}

static jlong m_GetDirectBufferCapacity(JNIEnv *env, jobject b)
{
    const mock_object *o = (const mock_object *)b;
    return o && o->kind == MOCK_BUFFER ? o->capacity : -1;
}

static struct JNINativeInterface_ table;
static JNIEnv the_env;

/* ---- what the test driver (ctypes) calls ---------------------------------- */
#define MOCK_API __attribute__((visibility("default")))

MOCK_API JNIEnv *mock_env(void)
{
    if (!the_env) {
        memset(&table, 0, sizeof table);
        table.FindClass = m_FindClass;
        table.ThrowNew = m_ThrowNew;
        table.ExceptionOccurred = m_ExceptionOccurred;
        table.ExceptionClear = m_ExceptionClear;
        table.ExceptionCheck = m_ExceptionCheck;
        table.GetArrayLength = m_GetArrayLength;
        table.GetIntArrayElements = m_GetIntArrayElements;
        table.GetLongArrayElements = m_GetLongArrayElements;
        table.GetBooleanArrayElements = m_GetBooleanArrayElements;
        table.ReleaseIntArrayElements = m_ReleaseIntArrayElements;
        table.ReleaseLongArrayElements = m_ReleaseLongArrayElements;
        table.ReleaseBooleanArrayElements = m_ReleaseBooleanArrayElements;
        table.SetBooleanArrayRegion = m_SetBooleanArrayRegion;
        table.GetDirectBufferAddress = m_GetDirectBufferAddress;
        table.GetDirectBufferCapacity = m_GetDirectBufferCapacity;
        the_env = &table;
    }
    return &the_env;
}

static jobject new_array(int kind, size_t width, const void *src, jsize n)
{
    mock_object *o = (mock_object *)calloc(1, sizeof *o);
    o->kind = kind;
    o->length = n;
    o->data = malloc(width * (size_t)(n > 0 ? n : 1));
    if (src && n > 0) memcpy(o->data, src, width * (size_t)n);
    return o;
}

MOCK_API jobject mock_new_int_array(const jint *src, jsize n) { return new_array(MOCK_INT_ARRAY, 4, src, n); }
MOCK_API jobject mock_new_long_array(const jlong *src, jsize n) { return new_array(MOCK_LONG_ARRAY, 8, src, n); }
MOCK_API jobject mock_new_boolean_array(const jboolean *src, jsize n) { return new_array(MOCK_BOOLEAN_ARRAY, 1, src, n); }
MOCK_API int mock_boolean_array_get(jobject a, jsize i) { return ((jboolean *)((mock_object *)a)->data)[i]; }

MOCK_API jobject mock_new_direct_buffer(void *address, jlong capacity)
{
    mock_object *o = (mock_object *)calloc(1, sizeof *o);
    o->kind = MOCK_BUFFER;
    o->data = address;
    o->capacity = capacity;
    return o;
}

MOCK_API void mock_free_object(jobject obj)
{
    mock_object *o = (mock_object *)obj;
    if (!o) return;
    if (o->kind != MOCK_BUFFER && o->kind != MOCK_CLASS) free(o->data);
    free(o);
}

MOCK_API int mock_exception_pending(void) { return pending; }
MOCK_API const char *mock_exception_class(void) { return pending_class; }
MOCK_API const char *mock_exception_message(void) { return pending_message; }
MOCK_API void mock_exception_clear(void) { m_ExceptionClear(NULL); }
MOCK_API int mock_outstanding_array_elements(void) { return outstanding; }
MOCK_API int mock_function_table_slots(void) { return (int)(sizeof(struct JNINativeInterface_) / sizeof(void *)); }
