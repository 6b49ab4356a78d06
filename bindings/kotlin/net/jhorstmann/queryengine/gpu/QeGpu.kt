package net.jhorstmann.queryengine.gpu

import java.nio.ByteBuffer

/**
 * JNI surface over the C ABI of libqegpu.so (include/qegpu.h); the native side is bindings/jni/qegpu_jni.c.
 * Handles are jlong; host buffers are direct ByteBuffers (little endian), owned by the caller.
 * A failing call throws IllegalStateException with qe_last_error().
 *
 * NOT COMPILED in the B200 build image (no JVM / Kotlin compiler there): source for the maintainer of
 * jhorstmann/queryengine (Kotlin 1.3.61, Java 11, pom.xml:13-17).  The natives behind every `external fun` below ARE
 * compiled and executed there, through a mock JNIEnv (tests/test_jni_cpu.py checks names and arities against this file,
 * tests/test_gpu_jni.py runs them on the GPU).
 */
object QeGpu {
    init {
        System.loadLibrary("qegpu_jni")
    }

    // qe_type (include/qegpu.h); STRING, DOUBLE, BOOLEAN are DataType.ordinal (data/Schema.kt:3-5)
    const val TYPE_STRING = 0
    const val TYPE_DOUBLE = 1
    const val TYPE_BOOLEAN = 2

    @JvmStatic external fun ctxCreate(device: Int): Long
    @JvmStatic external fun ctxDestroy(ctx: Long)
    @JvmStatic external fun programCreate(ctx: Long, words: IntArray): Long
    @JvmStatic external fun programDestroy(program: Long)
    @JvmStatic external fun tableCreate(ctx: Long, types: IntArray, nullable: BooleanArray, rows: Long): Long
    @JvmStatic external fun tableUpload(table: Long, col: Int, values: ByteBuffer, validity: ByteBuffer?, row0: Long, n: Long)
    @JvmStatic external fun tableDestroy(table: Long)
    @JvmStatic external fun pipelineRun(ctx: Long, table: Long, filter: Long, proj: LongArray, aggFns: IntArray, previous: Long): Long
    @JvmStatic external fun groupByRun(ctx: Long, table: Long, filter: Long, keys: LongArray, inputs: LongArray, aggFns: IntArray,
                                       maxGroups: Long, previous: Long): Long
    @JvmStatic external fun resultNumRows(result: Long): Long
    @JvmStatic external fun resultNumColumns(result: Long): Int
    @JvmStatic external fun resultColumnType(result: Long, k: Int): Int      // type | hasValidity shl 8
    @JvmStatic external fun resultColumn(result: Long, k: Int, values: ByteBuffer, validity: ByteBuffer)
    @JvmStatic external fun resultAgg(result: Long, k: Int, isNull: BooleanArray): Double
    @JvmStatic external fun resultDestroy(result: Long)

    // qe_pipeline_desc.flags
    const val RUN_NO_SELECTION = 1
    const val RUN_INTERPRETER = 2
    const val RUN_NO_VALUES = 8      // the selection vector is the whole output: FilterOperator over rows the JVM already holds

    @JvmStatic external fun pipelineRunFlags(ctx: Long, table: Long, filter: Long, proj: LongArray, aggFns: IntArray, flags: Int,
                                             previous: Long): Long
    @JvmStatic external fun resultSelection(result: Long, rows: ByteBuffer)            // uint32 per kept row

    // every GPU of the box from this one process (qe_mgpu_*): contiguous row ranges, results concatenated in scan order,
    // SUM / AVG bit-identical to one GPU, group-by partitions merged by key
    @JvmStatic external fun mgpuCreate(devices: IntArray): Long
    @JvmStatic external fun mgpuDestroy(group: Long)
    @JvmStatic external fun mgpuCtx(group: Long, part: Int): Long                      // programs of any of them run on all
    @JvmStatic external fun mtableCreate(group: Long, types: IntArray, nullable: BooleanArray, rows: Long): Long
    @JvmStatic external fun mtableUpload(table: Long, col: Int, values: ByteBuffer, validity: ByteBuffer?)
    @JvmStatic external fun mtableDestroy(table: Long)
    @JvmStatic external fun mgpuPipelineRun(group: Long, table: Long, filter: Long, proj: LongArray, aggFns: IntArray, flags: Int): Long
    @JvmStatic external fun mgpuGroupByRun(group: Long, table: Long, filter: Long, keys: LongArray, inputs: LongArray, aggFns: IntArray,
                                           maxGroups: Long): Long
    @JvmStatic external fun mresultNumRows(result: Long): Long
    @JvmStatic external fun mresultNumColumns(result: Long): Int
    @JvmStatic external fun mresultColumn(result: Long, k: Int, values: ByteBuffer, validity: ByteBuffer)
    @JvmStatic external fun mresultSelection(result: Long, rows: ByteBuffer)           // uint64 per kept row, whole-table indices
    @JvmStatic external fun mresultAgg(result: Long, k: Int, isNull: BooleanArray): Double
    @JvmStatic external fun mresultDestroy(result: Long)
}

/** One dictionary per context: STRING columns and literals are int32 codes of it (equal codes <=> equal strings). */
class StringDictionary {
    private val codes = HashMap<String, Int>()
    private val strings = ArrayList<String>()
    fun encode(s: String): Int = codes.getOrPut(s) { strings.add(s); strings.size - 1 }
    fun decode(code: Int): String = strings[code]
}

class GpuContext(device: Int = 0) : AutoCloseable {
    val handle: Long = QeGpu.ctxCreate(device)
    val dictionary = StringDictionary()
    override fun close() = QeGpu.ctxDestroy(handle)
}
