package net.jhorstmann.queryengine.gpu

import net.jhorstmann.queryengine.evaluator.*
import net.jhorstmann.queryengine.operator.Operator

/**
 * The branch buildPhysicalPlan (evaluator/Planner.kt:30-63) gains for Mode.GPU: fold the logical nodes
 *     [Aggregation] <- [Projection] <- [Filter] <- Scan(ColumnarGpuTable)
 * into one GpuPipelineOperator and return null for anything else, so that the caller falls through to the
 * existing interpreter / closure / bytecode operators (and builds OrderBy, or the final projection over the
 * aggregation's output row, on top as before).
 *
 * In Planner.kt, first line of buildPhysicalPlan:
 *     if (mode == Mode.GPU) buildGpuSubtree(tableRegistry, plan)?.let { return it }
 * (nodes the GPU does not take -- OrderBy, the final projection over an aggregation's row -- are built by the
 * existing branches, whose recursive buildPhysicalPlan(source, mode) call then reaches the GPU subtree)
 * and `enum class Mode { INTERPRETER, CLOSURE_COMPILER, BYTECODE_COMPILER, GPU }` (evaluator/Compiler.kt:5-7).
 */
fun buildGpuSubtree(tableRegistry: net.jhorstmann.queryengine.data.TableRegistry, plan: LogicalNode): Operator? {
    var node = plan
    var groupCount = -1
    var aggregates = emptyList<net.jhorstmann.queryengine.ast.AggregationFunction>()
    if (node is LogicalAggregationNode) {
        groupCount = if (node.groupCount > 0) node.groupCount else -1
        aggregates = node.aggregateFunctions
        node = node.source
    }
    val projection = node as? LogicalProjectionNode
    if (projection != null) node = projection.source
    val filter = node as? LogicalFilterNode
    if (filter != null) node = filter.source
    val scan = node as? LogicalScanNode ?: return null
    val table = tableRegistry.getTable(scan.table) as? ColumnarGpuTable ?: return null
    // expressions index the pruned scan schema (ResolveSchema.kt:53-68): map to the table's columns
    val names = scan.schema.fields.map { it.name }
    if (projection == null) {
        return if (filter == null && aggregates.isEmpty()) table.getScanOperator(names) else null
    }
    val scanColumns = IntArray(names.size) { i -> table.schema.fields.indexOfFirst { it.name == names[i] } }
    return GpuPipelineOperator(table, filter?.filter, projection.expressions, aggregates, groupCount, scanColumns)
}
