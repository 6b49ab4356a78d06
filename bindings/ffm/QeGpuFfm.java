package net.jhorstmann.queryengine.gpu;

import java.lang.foreign.Arena;
import java.lang.foreign.FunctionDescriptor;
import java.lang.foreign.Linker;
import java.lang.foreign.MemorySegment;
import java.lang.foreign.SymbolLookup;
import java.lang.invoke.MethodHandle;

import static java.lang.foreign.ValueLayout.ADDRESS;
import static java.lang.foreign.ValueLayout.JAVA_INT;
import static java.lang.foreign.ValueLayout.JAVA_LONG;

/**
 * The same C ABI (include/qegpu.h) through the Foreign Function & Memory API of JDK 22+ instead of the JNI shim of
 * bindings/jni/qegpu_jni.c: every entry point takes plain pointers and integers, so a downcall handle per function is
 * all the binding there is.  The reference targets Java 11 (pom.xml:13-15), which is why the JNI shim is the primary
 * binding; this file is what replaces it once the JDK is raised.  Source only: no JDK in the build image.
 *
 * Column buffers are off-heap MemorySegments (Arena.ofShared() or pinned memory from qe_host_alloc wrapped with
 * MemorySegment.ofAddress(..).reinterpret(bytes)); nothing is retained by the library past a call.  A non-zero status
 * becomes an IllegalStateException carrying qe_last_error(), like the unchecked exceptions of the reference
 * (evaluator/TypeCheck.kt:8, operator/OrderByOperator.kt:23).
 */
public final class QeGpuFfm {
    private static final Linker LINKER = Linker.nativeLinker();
    private static final SymbolLookup LIB = SymbolLookup.libraryLookup(System.mapLibraryName("qegpu"), Arena.global());

    private static MethodHandle h(String name, FunctionDescriptor fd) {
        return LINKER.downcallHandle(LIB.find(name).orElseThrow(() -> new UnsatisfiedLinkError(name)), fd);
    }

    // int qe_ctx_create(int device, qe_ctx **out);  int qe_ctx_destroy(qe_ctx *ctx);
    private static final MethodHandle CTX_CREATE = h("qe_ctx_create", FunctionDescriptor.of(JAVA_INT, JAVA_INT, ADDRESS));
    private static final MethodHandle CTX_DESTROY = h("qe_ctx_destroy", FunctionDescriptor.of(JAVA_INT, ADDRESS));
    // int qe_program_create(qe_ctx *ctx, const uint32_t *words, size_t nwords, qe_program **out);
    private static final MethodHandle PROGRAM_CREATE = h("qe_program_create", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS));
    private static final MethodHandle PROGRAM_DESTROY = h("qe_program_destroy", FunctionDescriptor.of(JAVA_INT, ADDRESS));
    // int qe_table_create(qe_ctx *ctx, int ncols, const qe_col_desc *cols, int64_t nrows, qe_table **out);
    private static final MethodHandle TABLE_CREATE = h("qe_table_create", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, ADDRESS, JAVA_LONG, ADDRESS));
    // int qe_table_upload(qe_table *table, int col, const void *values, const uint8_t *validity, int64_t row0, int64_t nrows);
    private static final MethodHandle TABLE_UPLOAD = h("qe_table_upload", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, JAVA_LONG));
    private static final MethodHandle TABLE_DESTROY = h("qe_table_destroy", FunctionDescriptor.of(JAVA_INT, ADDRESS));
    // int qe_pipeline_run(qe_ctx *ctx, const qe_table *table, const qe_program *filter, const qe_program *const *proj, int nproj,
    //                     const int32_t *agg_fns, int nagg, qe_result **result);
    private static final MethodHandle PIPELINE_RUN = h("qe_pipeline_run",
            FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS, ADDRESS, JAVA_INT, ADDRESS, JAVA_INT, ADDRESS));
    // int qe_result_num_rows(qe_result *result, int64_t *nrows);
    private static final MethodHandle RESULT_NUM_ROWS = h("qe_result_num_rows", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
    // int qe_result_column(qe_result *result, int k, void *dst_values, uint8_t *dst_validity);
    private static final MethodHandle RESULT_COLUMN = h("qe_result_column", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, ADDRESS, ADDRESS));
    // int qe_result_selection(qe_result *result, uint32_t *dst);
    private static final MethodHandle RESULT_SELECTION = h("qe_result_selection", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
    // int qe_result_agg(qe_result *result, int k, double *value, int *is_null);
    private static final MethodHandle RESULT_AGG = h("qe_result_agg", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, ADDRESS, ADDRESS));
    private static final MethodHandle RESULT_DESTROY = h("qe_result_destroy", FunctionDescriptor.of(JAVA_INT, ADDRESS));
    // const char *qe_last_error(void);
    private static final MethodHandle LAST_ERROR = h("qe_last_error", FunctionDescriptor.of(ADDRESS));

    private QeGpuFfm() {
    }

    private static void check(int status) throws Throwable {
        if (status != 0) {
            MemorySegment msg = ((MemorySegment) LAST_ERROR.invokeExact()).reinterpret(4096);
            throw new IllegalStateException(msg.getString(0));
        }
    }

    /** qe_ctx_create: one context per GPU (a handle is used by one thread at a time). */
    public static MemorySegment ctxCreate(int device) throws Throwable {
        try (Arena a = Arena.ofConfined()) {
            MemorySegment out = a.allocate(ADDRESS);
            check((int) CTX_CREATE.invokeExact(device, out));
            return out.get(ADDRESS, 0);
        }
    }

    public static void ctxDestroy(MemorySegment ctx) throws Throwable {
        check((int) CTX_DESTROY.invokeExact(ctx));
    }

    /** compileExpression(expr, mode) (evaluator/Compiler.kt:20-26): the postfix words of ProgramSerializer. */
    public static MemorySegment programCreate(MemorySegment ctx, int[] words) throws Throwable {
        try (Arena a = Arena.ofConfined()) {
            MemorySegment w = a.allocateFrom(JAVA_INT, words);
            MemorySegment out = a.allocate(ADDRESS);
            check((int) PROGRAM_CREATE.invokeExact(ctx, w, (long) words.length, out));
            return out.get(ADDRESS, 0);
        }
    }

    public static void programDestroy(MemorySegment program) throws Throwable {
        check((int) PROGRAM_DESTROY.invokeExact(program));
    }

    /** types[i] = qe_type (STRING, DOUBLE, BOOLEAN = DataType.ordinal, data/Schema.kt:3-5); qe_col_desc = { int32 type, int32 nullable }. */
    public static MemorySegment tableCreate(MemorySegment ctx, int[] types, boolean[] nullable, long rows) throws Throwable {
        try (Arena a = Arena.ofConfined()) {
            MemorySegment cols = a.allocate(JAVA_INT, 2L * types.length);
            for (int i = 0; i < types.length; i++) {
                cols.setAtIndex(JAVA_INT, 2L * i, types[i]);
                cols.setAtIndex(JAVA_INT, 2L * i + 1, nullable[i] ? 1 : 0);
            }
            MemorySegment out = a.allocate(ADDRESS);
            check((int) TABLE_CREATE.invokeExact(ctx, types.length, cols, rows, out));
            return out.get(ADDRESS, 0);
        }
    }

    /** values / validity: off-heap segments (validity may be MemorySegment.NULL: no nulls); copied before the call returns. */
    public static void tableUpload(MemorySegment table, int col, MemorySegment values, MemorySegment validity, long row0, long rows) throws Throwable {
        check((int) TABLE_UPLOAD.invokeExact(table, col, values, validity, row0, rows));
    }

    public static void tableDestroy(MemorySegment table) throws Throwable {
        check((int) TABLE_DESTROY.invokeExact(table));
    }

    /** The operator subtree buildPhysicalPlan assembles (evaluator/Planner.kt:30-63), pulled dry; result = previous result to reuse or NULL. */
    public static MemorySegment pipelineRun(MemorySegment ctx, MemorySegment table, MemorySegment filter, MemorySegment[] projections,
                                            int[] aggregates, MemorySegment result) throws Throwable {
        try (Arena a = Arena.ofConfined()) {
            MemorySegment proj = a.allocate(ADDRESS, Math.max(1, projections.length));
            for (int i = 0; i < projections.length; i++) proj.setAtIndex(ADDRESS, i, projections[i]);
            MemorySegment aggs = a.allocateFrom(JAVA_INT, aggregates.length == 0 ? new int[1] : aggregates);
            MemorySegment out = a.allocate(ADDRESS);
            out.set(ADDRESS, 0, result);
            check((int) PIPELINE_RUN.invokeExact(ctx, table, filter, proj, projections.length, aggs, aggregates.length, out));
            return out.get(ADDRESS, 0);
        }
    }

    public static long resultNumRows(MemorySegment result) throws Throwable {
        try (Arena a = Arena.ofConfined()) {
            MemorySegment n = a.allocate(JAVA_LONG);
            check((int) RESULT_NUM_ROWS.invokeExact(result, n));
            return n.get(JAVA_LONG, 0);
        }
    }

    public static void resultColumn(MemorySegment result, int k, MemorySegment values, MemorySegment validity) throws Throwable {
        check((int) RESULT_COLUMN.invokeExact(result, k, values, validity));
    }

    public static void resultSelection(MemorySegment result, MemorySegment dst) throws Throwable {
        check((int) RESULT_SELECTION.invokeExact(result, dst));
    }

    /** Accumulator.finish() (evaluator/Accumulators.kt:21-107): null when no non-null input was seen (COUNT: 0). */
    public static Double resultAgg(MemorySegment result, int k) throws Throwable {
        try (Arena a = Arena.ofConfined()) {
            MemorySegment v = a.allocate(java.lang.foreign.ValueLayout.JAVA_DOUBLE);
            MemorySegment isNull = a.allocate(JAVA_INT);
            check((int) RESULT_AGG.invokeExact(result, k, v, isNull));
            return isNull.get(JAVA_INT, 0) != 0 ? null : v.get(java.lang.foreign.ValueLayout.JAVA_DOUBLE, 0);
        }
    }

    public static void resultDestroy(MemorySegment result) throws Throwable {
        check((int) RESULT_DESTROY.invokeExact(result));
    }
}
