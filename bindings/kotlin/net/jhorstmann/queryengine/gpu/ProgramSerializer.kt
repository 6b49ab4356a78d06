package net.jhorstmann.queryengine.gpu

import net.jhorstmann.queryengine.ast.*

/**
 * Resolved expression tree -> the postfix 32-bit words qe_program_create takes (include/qegpu.h):
 * word = opcode | argument shl 8; opcodes 0..16 are Function.ordinal (ast/Functions.kt:7-22).
 * ColumnExpression.index is the position in the PRUNED scan schema (evaluator/ResolveSchema.kt:53-68); `scanColumns`
 * maps it to the column of the HBM-resident table (scanColumns[i] = table column of pruned field i).
 * FunctionExpression.dataType is not consulted: the library infers and checks types with the rules of
 * evaluator/TypeCheck.kt:38-96 (hand-built trees leave it unset, test/.../ByteCodeCompilerTest.kt:28).
 * `ieeeOrder` = the semantics of Mode.CLOSURE_COMPILER, whose ordering comparisons unbox to primitive doubles
 * (evaluator/ClosureCompiler.kt:127-130): QE_CMP_IEEE in the argument of CMP_LT / LE / GE / GT.  The default is
 * Double.compare's total order (evaluator/Interpreter.kt:104-107, BytecodeCompiler.kt:300).
 */
class ProgramSerializer(private val dictionary: StringDictionary, private val scanColumns: IntArray? = null,
                        private val ieeeOrder: Boolean = false) : ExpressionVisitor<Unit> {
    private val words = ArrayList<Int>()

    fun serialize(expr: Expression): IntArray {
        words.clear()
        expr.accept(this)
        return words.toIntArray()
    }

    override fun visitColumn(expr: ColumnExpression) {
        words.add(OP_LOAD_COL or ((scanColumns?.get(expr.index) ?: expr.index) shl 8))
    }

    override fun visitNumericLiteral(expr: NumericLiteralExpression) {
        val bits = java.lang.Double.doubleToRawLongBits(expr.value)
        words.add(OP_CONST_F64)
        words.add(bits.toInt())
        words.add((bits ushr 32).toInt())
    }

    override fun visitBooleanLiteral(expr: BooleanLiteralExpression) {
        words.add(OP_CONST_BOOL or ((if (expr.value) 1 else 0) shl 8))
    }

    override fun visitStringLiteral(expr: StringLiteralExpression) {
        words.add(OP_CONST_STR)
        words.add(dictionary.encode(expr.value))
    }

    override fun visitFunction(expr: FunctionExpression) {
        expr.operands.forEach { it.accept(this) }
        val f = expr.function
        val ordering = f == Function.CMP_LT || f == Function.CMP_LE || f == Function.CMP_GE || f == Function.CMP_GT
        words.add(if (ieeeOrder && ordering) f.ordinal or (QE_CMP_IEEE shl 8) else f.ordinal)
    }

    // identifiers never reach the evaluator (evaluator/Interpreter.kt:9-11); aggregates are split off by
    // rewriteAggregates (evaluator/RewriteAggregates.kt:18-49) before a physical plan is built
    override fun visitIdentifier(expr: IdentifierExpression) = throw IllegalStateException("Unresolved identifier ${expr.name}")

    override fun visitAggregationFunction(expr: AggregationFunctionExpression) =
            throw IllegalStateException("Aggregation functions are planned, not evaluated")

    companion object {
        const val OP_LOAD_COL = 32
        const val OP_CONST_F64 = 33
        const val OP_CONST_BOOL = 34
        const val OP_CONST_STR = 35
        const val QE_CMP_IEEE = 1
    }
}
