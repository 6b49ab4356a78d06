package net.jhorstmann.queryengine.gpu

import net.jhorstmann.queryengine.data.DataType
import net.jhorstmann.queryengine.data.Schema
import net.jhorstmann.queryengine.data.Table
import net.jhorstmann.queryengine.operator.Operator
import java.nio.ByteBuffer
import java.nio.ByteOrder

/**
 * Same contract as MemoryTable (data/MemoryTable.kt:7-19), data resident in HBM as columns:
 * DOUBLE -> double[n], BOOLEAN -> LSB-first bitmap, STRING -> int32 dictionary codes; every column
 * nullable (validity bitmap), like the boxed List<List<Any?>> it replaces.
 */
class ColumnarGpuTable(val ctx: GpuContext, override val schema: Schema, values: List<List<Any?>>) : Table(), AutoCloseable {
    val rows: Long = values.size.toLong()
    val handle: Long

    init {
        val fields = schema.fields
        handle = QeGpu.tableCreate(ctx.handle, IntArray(fields.size) { fields[it].type.ordinal }, BooleanArray(fields.size) { true }, rows)
        val n = values.size
        for ((c, field) in fields.withIndex()) {
            val validity = direct((n + 7) / 8)
            val data = when (field.type) {
                DataType.DOUBLE -> direct(8 * n)
                DataType.STRING -> direct(4 * n)
                DataType.BOOLEAN -> direct((n + 7) / 8)
            }
            for (i in 0 until n) {
                val v = values[i][c] ?: continue
                validity.put(i / 8, (validity.get(i / 8).toInt() or (1 shl (i % 8))).toByte())
                when (field.type) {
                    DataType.DOUBLE -> data.putDouble(8 * i, v as Double)
                    DataType.STRING -> data.putInt(4 * i, ctx.dictionary.encode(v as String))
                    DataType.BOOLEAN -> if (v as Boolean) data.put(i / 8, (data.get(i / 8).toInt() or (1 shl (i % 8))).toByte())
                }
            }
            QeGpu.tableUpload(handle, c, data, validity, 0, rows)
        }
    }

    /** Table.getScanOperator(projection) (data/Table.kt:8): a scan of the named columns. */
    override fun getScanOperator(projection: List<String>): Operator {
        val indices = projection.map { name ->
            val idx = schema.fields.indexOfFirst { it.name == name }
            if (idx < 0) throw IllegalArgumentException("Unknown field $name")      // as MemoryTable.kt:12
            idx
        }
        return GpuPipelineOperator.scan(this, indices)
    }

    override fun close() = QeGpu.tableDestroy(handle)

    private fun direct(bytes: Int): ByteBuffer = ByteBuffer.allocateDirect(maxOf(bytes, 1)).order(ByteOrder.LITTLE_ENDIAN)
}
