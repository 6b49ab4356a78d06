/*
 * qegpu_jni.c -- JNI glue between the reference's JVM (Java 11, pom.xml:13-15) and the C ABI of
 * libqegpu.so (include/qegpu.h).  One native method per entry point the Kotlin adapter
 * (bindings/kotlin/.../gpu/QeGpu.kt) declares; handles travel as jlong, host buffers as direct
 * ByteBuffers (never copied, never retained), small argument lists as primitive arrays.  A non-zero
 * qe_status becomes an IllegalStateException carrying qe_last_error() -- the reference's own error style
 * (unchecked exceptions, e.g. M/operator/OrderByOperator.kt:23).
 *
 * This image has no JDK: here the file is compiled against bindings/jni/jni_stub/jni.h (the JNI specification's function
 * table, same slot order) and EXECUTED through a mock JNIEnv (tests/jni_mock.c, tests/test_jni_cpu.py,
 * tests/test_gpu_jni.py).  Build where a JDK exists:
 *   gcc -shared -fPIC -I$JAVA_HOME/include -I$JAVA_HOME/include/linux -Iinclude \
 *       bindings/jni/qegpu_jni.c -Lqueryengine_b200 -lqegpu -o libqegpu_jni.so
 */
#include <jni.h>
#include <stdint.h>
#include <stdlib.h>

#include "qegpu.h"

#define QE_JNI(name) Java_net_jhorstmann_queryengine_gpu_QeGpu_##name
#define PTR(T, h) ((T *)(intptr_t)(h))

static int check(JNIEnv *env, int status)
{
    if (status == QE_OK) return 0;
    jclass cls = (*env)->FindClass(env, "java/lang/IllegalStateException");
    if (cls) (*env)->ThrowNew(env, cls, qe_last_error());
    return status;
}

static void *buffer(JNIEnv *env, jobject buf)
{
    return buf ? (*env)->GetDirectBufferAddress(env, buf) : NULL;
}

JNIEXPORT jlong JNICALL QE_JNI(ctxCreate)(JNIEnv *env, jclass c, jint device)
{
    qe_ctx *ctx = NULL;
    check(env, qe_ctx_create(device, &ctx));
    return (jlong)(intptr_t)ctx;
}

JNIEXPORT void JNICALL QE_JNI(ctxDestroy)(JNIEnv *env, jclass c, jlong ctx)
{
    check(env, qe_ctx_destroy(PTR(qe_ctx, ctx)));
}

/* compileExpression(expr, Mode.GPU): the flattened tree, M/evaluator/Compiler.kt:20-26 */
JNIEXPORT jlong JNICALL QE_JNI(programCreate)(JNIEnv *env, jclass c, jlong ctx, jintArray words)
{
    qe_program *p = NULL;
    const jsize n = (*env)->GetArrayLength(env, words);
    jint *w = (*env)->GetIntArrayElements(env, words, NULL);
    check(env, qe_program_create(PTR(qe_ctx, ctx), (const uint32_t *)w, (size_t)n, &p));
    (*env)->ReleaseIntArrayElements(env, words, w, JNI_ABORT);
    return (jlong)(intptr_t)p;
}

JNIEXPORT void JNICALL QE_JNI(programDestroy)(JNIEnv *env, jclass c, jlong program)
{
    check(env, qe_program_destroy(PTR(qe_program, program)));
}

/* MemoryTable(schema, values), M/data/MemoryTable.kt:7 -- columnar, resident in HBM */
JNIEXPORT jlong JNICALL QE_JNI(tableCreate)(JNIEnv *env, jclass c, jlong ctx, jintArray types, jbooleanArray nullable, jlong rows)
{
    qe_table *t = NULL;
    const jsize n = (*env)->GetArrayLength(env, types);
    jint *ty = (*env)->GetIntArrayElements(env, types, NULL);
    jboolean *nu = (*env)->GetBooleanArrayElements(env, nullable, NULL);
    qe_col_desc *cols = (qe_col_desc *)malloc(sizeof(qe_col_desc) * (size_t)(n > 0 ? n : 1));
    for (jsize i = 0; i < n; i++) { cols[i].type = ty[i]; cols[i].nullable = nu[i] ? 1 : 0; }
    check(env, qe_table_create(PTR(qe_ctx, ctx), (int)n, cols, (int64_t)rows, &t));
    free(cols);
    (*env)->ReleaseIntArrayElements(env, types, ty, JNI_ABORT);
    (*env)->ReleaseBooleanArrayElements(env, nullable, nu, JNI_ABORT);
    return (jlong)(intptr_t)t;
}

JNIEXPORT void JNICALL QE_JNI(tableUpload)(JNIEnv *env, jclass c, jlong table, jint col, jobject values, jobject validity,
                                           jlong row0, jlong n)
{
    check(env, qe_table_upload(PTR(qe_table, table), col, buffer(env, values), (const uint8_t *)buffer(env, validity),
                               (int64_t)row0, (int64_t)n));
}

JNIEXPORT void JNICALL QE_JNI(tableDestroy)(JNIEnv *env, jclass c, jlong table)
{
    check(env, qe_table_destroy(PTR(qe_table, table)));
}

/* the operator subtree of buildPhysicalPlan, M/evaluator/Planner.kt:30-63, pulled dry */
JNIEXPORT jlong JNICALL QE_JNI(pipelineRun)(JNIEnv *env, jclass c, jlong ctx, jlong table, jlong filter, jlongArray proj,
                                            jintArray aggFns, jlong previous)
{
    qe_result *r = PTR(qe_result, previous);
    const jsize np = (*env)->GetArrayLength(env, proj);
    const jsize na = (*env)->GetArrayLength(env, aggFns);
    jlong *ph = (*env)->GetLongArrayElements(env, proj, NULL);
    jint *af = (*env)->GetIntArrayElements(env, aggFns, NULL);
    const qe_program *progs[QE_MAX_PROJECTIONS];
    for (jsize i = 0; i < np && i < QE_MAX_PROJECTIONS; i++) progs[i] = PTR(qe_program, ph[i]);
    check(env, qe_pipeline_run(PTR(qe_ctx, ctx), PTR(qe_table, table), PTR(qe_program, filter), progs, (int)np,
                               (const int32_t *)af, (int)na, &r));
    (*env)->ReleaseLongArrayElements(env, proj, ph, JNI_ABORT);
    (*env)->ReleaseIntArrayElements(env, aggFns, af, JNI_ABORT);
    return (jlong)(intptr_t)r;
}

/* GroupByAggregationOperator, M/operator/GroupByAggregationOperator.kt:21-75 */
JNIEXPORT jlong JNICALL QE_JNI(groupByRun)(JNIEnv *env, jclass c, jlong ctx, jlong table, jlong filter, jlongArray keys,
                                           jlongArray inputs, jintArray aggFns, jlong maxGroups, jlong previous)
{
    qe_result *r = PTR(qe_result, previous);
    const jsize nk = (*env)->GetArrayLength(env, keys);
    const jsize na = (*env)->GetArrayLength(env, inputs);
    jlong *kh = (*env)->GetLongArrayElements(env, keys, NULL);
    jlong *ih = (*env)->GetLongArrayElements(env, inputs, NULL);
    jint *af = (*env)->GetIntArrayElements(env, aggFns, NULL);
    const qe_program *kp[8], *ip[16];
    for (jsize i = 0; i < nk && i < 8; i++) kp[i] = PTR(qe_program, kh[i]);
    for (jsize i = 0; i < na && i < 16; i++) ip[i] = PTR(qe_program, ih[i]);
    qe_groupby_desc d;
    d.filter = PTR(qe_program, filter);
    d.keys = kp;
    d.nkeys = (int32_t)nk;
    d.inputs = ip;
    d.agg_fns = (const int32_t *)af;
    d.nagg = (int32_t)na;
    d.row_begin = 0;
    d.row_end = -1;
    d.max_groups = (int64_t)maxGroups;
    d.flags = 0;
    check(env, qe_groupby_run(PTR(qe_ctx, ctx), PTR(qe_table, table), &d, &r));
    (*env)->ReleaseLongArrayElements(env, keys, kh, JNI_ABORT);
    (*env)->ReleaseLongArrayElements(env, inputs, ih, JNI_ABORT);
    (*env)->ReleaseIntArrayElements(env, aggFns, af, JNI_ABORT);
    return (jlong)(intptr_t)r;
}

JNIEXPORT jlong JNICALL QE_JNI(resultNumRows)(JNIEnv *env, jclass c, jlong result)
{
    int64_t n = 0;
    check(env, qe_result_num_rows(PTR(qe_result, result), &n));
    return (jlong)n;
}

JNIEXPORT jint JNICALL QE_JNI(resultNumColumns)(JNIEnv *env, jclass c, jlong result)
{
    int n = 0;
    check(env, qe_result_num_columns(PTR(qe_result, result), &n));
    return n;
}

/* type | hasValidity << 8 */
JNIEXPORT jint JNICALL QE_JNI(resultColumnType)(JNIEnv *env, jclass c, jlong result, jint k)
{
    int type = 0, has_validity = 0;
    check(env, qe_result_column_type(PTR(qe_result, result), k, &type, &has_validity));
    return type | (has_validity << 8);
}

JNIEXPORT void JNICALL QE_JNI(resultColumn)(JNIEnv *env, jclass c, jlong result, jint k, jobject values, jobject validity)
{
    check(env, qe_result_column(PTR(qe_result, result), k, buffer(env, values), (uint8_t *)buffer(env, validity)));
}

/* Accumulator.finish(), M/evaluator/Accumulators.kt:21-107; NaN payload 0x7ff8dead00000000 = SQL NULL */
JNIEXPORT jdouble JNICALL QE_JNI(resultAgg)(JNIEnv *env, jclass c, jlong result, jint k, jbooleanArray isNull)
{
    double v = 0.0;
    int is_null = 0;
    check(env, qe_result_agg(PTR(qe_result, result), k, &v, &is_null));
    jboolean b = is_null ? JNI_TRUE : JNI_FALSE;
    (*env)->SetBooleanArrayRegion(env, isNull, 0, 1, &b);
    return v;
}

JNIEXPORT void JNICALL QE_JNI(resultDestroy)(JNIEnv *env, jclass c, jlong result)
{
    check(env, qe_result_destroy(PTR(qe_result, result)));
}

/* the same with qe_pipeline_desc.flags (QE_RUN_NO_VALUES: FilterOperator hands on the row objects the JVM already holds,
 * M/operator/FilterOperator.kt:20-22 -- the selection vector is all it needs back) */
JNIEXPORT jlong JNICALL QE_JNI(pipelineRunFlags)(JNIEnv *env, jclass c, jlong ctx, jlong table, jlong filter, jlongArray proj,
                                                 jintArray aggFns, jint flags, jlong previous)
{
    qe_result *r = PTR(qe_result, previous);
    const jsize np = (*env)->GetArrayLength(env, proj);
    const jsize na = (*env)->GetArrayLength(env, aggFns);
    jlong *ph = (*env)->GetLongArrayElements(env, proj, NULL);
    jint *af = (*env)->GetIntArrayElements(env, aggFns, NULL);
    const qe_program *progs[QE_MAX_PROJECTIONS];
    for (jsize i = 0; i < np && i < QE_MAX_PROJECTIONS; i++) progs[i] = PTR(qe_program, ph[i]);
    qe_pipeline_desc d;
    d.filter = PTR(qe_program, filter);
    d.proj = progs;
    d.nproj = (int32_t)np;
    d.agg_fns = (const int32_t *)af;
    d.nagg = (int32_t)na;
    d.row_begin = 0;
    d.row_end = -1;
    d.out_capacity_rows = 0;
    d.flags = (uint32_t)flags;
    d.sel_offset = 0;
    check(env, qe_pipeline_run_ex(PTR(qe_ctx, ctx), PTR(qe_table, table), &d, &r));
    (*env)->ReleaseLongArrayElements(env, proj, ph, JNI_ABORT);
    (*env)->ReleaseIntArrayElements(env, aggFns, af, JNI_ABORT);
    return (jlong)(intptr_t)r;
}

/* the selection vector: uint32 table row indices of the rows the filter kept, ascending */
JNIEXPORT void JNICALL QE_JNI(resultSelection)(JNIEnv *env, jclass c, jlong result, jobject rows)
{
    check(env, qe_result_selection(PTR(qe_result, result), (uint32_t *)buffer(env, rows)));
}

/* ---- every GPU of the box from the one process a JVM is (qe_mgpu_*, include/qegpu.h) ---- */

JNIEXPORT jlong JNICALL QE_JNI(mgpuCreate)(JNIEnv *env, jclass c, jintArray devices)
{
    qe_mgpu *g = NULL;
    const jsize n = (*env)->GetArrayLength(env, devices);
    jint *d = (*env)->GetIntArrayElements(env, devices, NULL);
    check(env, qe_mgpu_create((const int *)d, (int)n, &g));
    (*env)->ReleaseIntArrayElements(env, devices, d, JNI_ABORT);
    return (jlong)(intptr_t)g;
}

JNIEXPORT void JNICALL QE_JNI(mgpuDestroy)(JNIEnv *env, jclass c, jlong group)
{
    check(env, qe_mgpu_destroy(PTR(qe_mgpu, group)));
}

/* the context of partition `part`: programs created with any of them run on all */
JNIEXPORT jlong JNICALL QE_JNI(mgpuCtx)(JNIEnv *env, jclass c, jlong group, jint part)
{
    return (jlong)(intptr_t)qe_mgpu_ctx(PTR(qe_mgpu, group), part);
}

JNIEXPORT jlong JNICALL QE_JNI(mtableCreate)(JNIEnv *env, jclass c, jlong group, jintArray types, jbooleanArray nullable, jlong rows)
{
    qe_mtable *t = NULL;
    const jsize n = (*env)->GetArrayLength(env, types);
    jint *ty = (*env)->GetIntArrayElements(env, types, NULL);
    jboolean *nu = (*env)->GetBooleanArrayElements(env, nullable, NULL);
    qe_col_desc *cols = (qe_col_desc *)malloc(sizeof(qe_col_desc) * (size_t)(n > 0 ? n : 1));
    for (jsize i = 0; i < n; i++) { cols[i].type = ty[i]; cols[i].nullable = nu[i] ? 1 : 0; }
    check(env, qe_mtable_create(PTR(qe_mgpu, group), (int)n, cols, (int64_t)rows, &t));
    free(cols);
    (*env)->ReleaseIntArrayElements(env, types, ty, JNI_ABORT);
    (*env)->ReleaseBooleanArrayElements(env, nullable, nu, JNI_ABORT);
    return (jlong)(intptr_t)t;
}

JNIEXPORT void JNICALL QE_JNI(mtableUpload)(JNIEnv *env, jclass c, jlong table, jint col, jobject values, jobject validity)
{
    check(env, qe_mtable_upload(PTR(qe_mtable, table), col, buffer(env, values), (const uint8_t *)buffer(env, validity)));
}

JNIEXPORT void JNICALL QE_JNI(mtableDestroy)(JNIEnv *env, jclass c, jlong table)
{
    check(env, qe_mtable_destroy(PTR(qe_mtable, table)));
}

JNIEXPORT jlong JNICALL QE_JNI(mgpuPipelineRun)(JNIEnv *env, jclass c, jlong group, jlong table, jlong filter, jlongArray proj,
                                                jintArray aggFns, jint flags)
{
    qe_mresult *r = NULL;
    const jsize np = (*env)->GetArrayLength(env, proj);
    const jsize na = (*env)->GetArrayLength(env, aggFns);
    jlong *ph = (*env)->GetLongArrayElements(env, proj, NULL);
    jint *af = (*env)->GetIntArrayElements(env, aggFns, NULL);
    const qe_program *progs[QE_MAX_PROJECTIONS];
    for (jsize i = 0; i < np && i < QE_MAX_PROJECTIONS; i++) progs[i] = PTR(qe_program, ph[i]);
    qe_pipeline_desc d;
    d.filter = PTR(qe_program, filter);
    d.proj = progs;
    d.nproj = (int32_t)np;
    d.agg_fns = (const int32_t *)af;
    d.nagg = (int32_t)na;
    d.row_begin = 0;
    d.row_end = -1;
    d.out_capacity_rows = 0;
    d.flags = (uint32_t)flags;
    d.sel_offset = 0;
    check(env, qe_mgpu_pipeline_run(PTR(qe_mgpu, group), PTR(qe_mtable, table), &d, &r));
    (*env)->ReleaseLongArrayElements(env, proj, ph, JNI_ABORT);
    (*env)->ReleaseIntArrayElements(env, aggFns, af, JNI_ABORT);
    return (jlong)(intptr_t)r;
}

JNIEXPORT jlong JNICALL QE_JNI(mgpuGroupByRun)(JNIEnv *env, jclass c, jlong group, jlong table, jlong filter, jlongArray keys,
                                               jlongArray inputs, jintArray aggFns, jlong maxGroups)
{
    qe_mresult *r = NULL;
    const jsize nk = (*env)->GetArrayLength(env, keys);
    const jsize na = (*env)->GetArrayLength(env, inputs);
    jlong *kh = (*env)->GetLongArrayElements(env, keys, NULL);
    jlong *ih = (*env)->GetLongArrayElements(env, inputs, NULL);
    jint *af = (*env)->GetIntArrayElements(env, aggFns, NULL);
    const qe_program *kp[8], *ip[16];
    for (jsize i = 0; i < nk && i < 8; i++) kp[i] = PTR(qe_program, kh[i]);
    for (jsize i = 0; i < na && i < 16; i++) ip[i] = PTR(qe_program, ih[i]);
    qe_groupby_desc d;
    d.filter = PTR(qe_program, filter);
    d.keys = kp;
    d.nkeys = (int32_t)nk;
    d.inputs = ip;
    d.agg_fns = (const int32_t *)af;
    d.nagg = (int32_t)na;
    d.row_begin = 0;
    d.row_end = -1;
    d.max_groups = (int64_t)maxGroups;
    d.flags = 0;
    check(env, qe_mgpu_groupby_run(PTR(qe_mgpu, group), PTR(qe_mtable, table), &d, &r));
    (*env)->ReleaseLongArrayElements(env, keys, kh, JNI_ABORT);
    (*env)->ReleaseLongArrayElements(env, inputs, ih, JNI_ABORT);
    (*env)->ReleaseIntArrayElements(env, aggFns, af, JNI_ABORT);
    return (jlong)(intptr_t)r;
}

JNIEXPORT jlong JNICALL QE_JNI(mresultNumRows)(JNIEnv *env, jclass c, jlong result)
{
    int64_t n = 0;
    check(env, qe_mresult_num_rows(PTR(qe_mresult, result), &n));
    return (jlong)n;
}

JNIEXPORT jint JNICALL QE_JNI(mresultNumColumns)(JNIEnv *env, jclass c, jlong result)
{
    int n = 0;
    check(env, qe_mresult_num_columns(PTR(qe_mresult, result), &n));
    return n;
}

/* all partitions in order = the reference's scan order (M/operator/FilterOperator.kt:17-22) */
JNIEXPORT void JNICALL QE_JNI(mresultColumn)(JNIEnv *env, jclass c, jlong result, jint k, jobject values, jobject validity)
{
    check(env, qe_mresult_column(PTR(qe_mresult, result), k, buffer(env, values), (uint8_t *)buffer(env, validity)));
}

/* uint64 row indices into the whole table */
JNIEXPORT void JNICALL QE_JNI(mresultSelection)(JNIEnv *env, jclass c, jlong result, jobject rows)
{
    check(env, qe_mresult_selection(PTR(qe_mresult, result), (uint64_t *)buffer(env, rows)));
}

JNIEXPORT jdouble JNICALL QE_JNI(mresultAgg)(JNIEnv *env, jclass c, jlong result, jint k, jbooleanArray isNull)
{
    double v = 0.0;
    int is_null = 0;
    check(env, qe_mresult_agg(PTR(qe_mresult, result), k, &v, &is_null));
    jboolean b = is_null ? JNI_TRUE : JNI_FALSE;
    (*env)->SetBooleanArrayRegion(env, isNull, 0, 1, &b);
    return v;
}

JNIEXPORT void JNICALL QE_JNI(mresultDestroy)(JNIEnv *env, jclass c, jlong result)
{
    check(env, qe_mresult_destroy(PTR(qe_mresult, result)));
}
