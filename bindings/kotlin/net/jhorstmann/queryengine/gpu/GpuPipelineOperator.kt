package net.jhorstmann.queryengine.gpu

import net.jhorstmann.queryengine.ast.AggregationFunction
import net.jhorstmann.queryengine.ast.ColumnExpression
import net.jhorstmann.queryengine.ast.Expression
import net.jhorstmann.queryengine.operator.Operator
import java.nio.ByteBuffer
import java.nio.ByteOrder

/**
 * The subtree [GlobalAggregation | GroupByAggregation] <- [Projection] <- [Filter] <- Scan that buildPhysicalPlan
 * assembles (evaluator/Planner.kt:30-63), as ONE Operator (operator/Operators.kt:5-11).  open() runs the pipeline
 * on the GPU eagerly -- as the aggregation operators already do (operator/GlobalAggregationOperator.kt:10) --,
 * next() hands out boxed rows one at a time, close() releases the device buffers.
 *
 * `groupCount` < 0: no aggregation, or a global one (`aggregates` non-empty, one row out);
 * `groupCount` >= 0: `projections` = group keys, then the aggregate inputs (operator/GroupByAggregationOperator.kt:21-48).
 */
class GpuPipelineOperator(private val table: ColumnarGpuTable, private val filter: Expression?,
                          private val projections: List<Expression>, private val aggregates: List<AggregationFunction>,
                          private val groupCount: Int = -1, private val scanColumns: IntArray? = null,
                          private val ieeeOrder: Boolean = false) : Operator() {
    private var result = 0L
    private var cursor = 0L
    private var rows = 0L
    private var columns: Array<Column>? = null

    private class Column(val type: Int, val values: ByteBuffer, val validity: ByteBuffer)

    override fun open() {
        val ctx = table.ctx
        val ser = ProgramSerializer(ctx.dictionary, scanColumns, ieeeOrder)     // true: the semantics of Mode.CLOSURE_COMPILER
        val f = filter?.let { QeGpu.programCreate(ctx.handle, ser.serialize(it)) } ?: 0L
        val progs = LongArray(projections.size) { QeGpu.programCreate(ctx.handle, ser.serialize(projections[it])) }
        val fns = IntArray(aggregates.size) { aggregates[it].ordinal }
        try {
            result = if (groupCount >= 0) {
                QeGpu.groupByRun(ctx.handle, table.handle, f, progs.copyOfRange(0, groupCount),
                        progs.copyOfRange(groupCount, progs.size), fns, 0, result)
            } else {
                QeGpu.pipelineRun(ctx.handle, table.handle, f, progs, fns, result)
            }
        } finally {
            progs.forEach { QeGpu.programDestroy(it) }
            if (f != 0L) QeGpu.programDestroy(f)
        }
        rows = QeGpu.resultNumRows(result)
        cursor = 0
        val n = rows.toInt()
        columns = Array(QeGpu.resultNumColumns(result)) { k ->
            val type = QeGpu.resultColumnType(result, k) and 0xff
            val width = when (type) { QeGpu.TYPE_BOOLEAN -> 0; QeGpu.TYPE_STRING -> 4; else -> 8 }
            val values = direct(if (width == 0) (n + 7) / 8 else width * n)
            val validity = direct((n + 7) / 8)
            QeGpu.resultColumn(result, k, values, validity)
            Column(type, values, validity)
        }
    }

    override fun next(): Array<Any?>? {
        val cols = columns ?: throw IllegalStateException("Operator not opened")
        if (cursor >= rows) return null
        val i = (cursor++).toInt()
        return Array(cols.size) { k ->
            val c = cols[k]
            if ((c.validity.get(i / 8).toInt() shr (i % 8)) and 1 == 0) {
                null
            } else when (c.type) {
                QeGpu.TYPE_BOOLEAN -> (c.values.get(i / 8).toInt() shr (i % 8)) and 1 == 1
                QeGpu.TYPE_STRING -> table.ctx.dictionary.decode(c.values.getInt(4 * i))
                else -> {
                    val v = c.values.getDouble(8 * i)
                    val agg = k - maxOf(groupCount, 0)
                    // CountAccumulator.finish() is a Kotlin Int (evaluator/Accumulators.kt:33-35)
                    if (aggregates.isNotEmpty() && agg >= 0 && aggregates[agg] == AggregationFunction.COUNT) v.toInt() else v
                }
            }
        }
    }

    override fun close() {
        if (result != 0L) QeGpu.resultDestroy(result)
        result = 0
        columns = null
    }

    private fun direct(bytes: Int): ByteBuffer = ByteBuffer.allocateDirect(maxOf(bytes, 1)).order(ByteOrder.LITTLE_ENDIAN)

    companion object {
        fun scan(table: ColumnarGpuTable, indices: List<Int>): GpuPipelineOperator {
            val fields = table.schema.fields
            return GpuPipelineOperator(table, null, indices.map { ColumnExpression(fields[it].name, it, fields[it].type) }, emptyList())
        }
    }
}
