"""Pins the CPU oracle against every golden vector the reference's own tests
hold for the hot path.  Each test cites the reference
test it transcribes; T/ = src/test/java/net/jhorstmann/queryengine/.

These read like T/evaluator/CompilerTest.kt on purpose: same trees, same rows,
same expectations, parameterised over the same three modes.
"""
import math

import numpy as np
import pytest

from oracle_binding import EvalError, OraclePlan, OracleTable, compile_expression
from queryengine_b200.ast import (
    AggregationFunction, BooleanLiteralExpression, ColumnExpression, DataType, Function,
    FunctionExpression, Mode, NumericLiteralExpression, StringLiteralExpression,
)

MODES = [Mode.INTERPRETER, Mode.CLOSURE_COMPILER, Mode.BYTECODE_COMPILER]


@pytest.mark.parametrize("mode", MODES)
def test_should_load_column(mode):
    # T/evaluator/CompilerTest.kt:15-23
    expr = ColumnExpression("foo", 0, DataType.STRING)
    callable_ = compile_expression(expr, mode)
    assert callable_(["foobar"]) == "foobar"


@pytest.mark.parametrize("mode", MODES)
def test_should_add_numeric_literals(mode):
    # T/evaluator/CompilerTest.kt:27-35
    expr = FunctionExpression(Function.ADD, [NumericLiteralExpression(1.0), NumericLiteralExpression(2.0)],
                              DataType.DOUBLE)
    assert compile_expression(expr, mode)([]) == 3.0


@pytest.mark.parametrize("mode", MODES)
def test_should_handle_null_values(mode):
    # T/evaluator/CompilerTest.kt:39-47
    expr = FunctionExpression(Function.MUL, [ColumnExpression("foo", 0, DataType.DOUBLE),
                                             ColumnExpression("bar", 1, DataType.DOUBLE)], DataType.DOUBLE)
    assert compile_expression(expr, mode)([10.0, None]) is None


AND_TRUTH = [  # T/evaluator/CompilerTest.kt:55-65
    ((True, True), True), ((True, False), False), ((True, None), None),
    ((False, True), False), ((False, False), False), ((False, None), False),
    ((None, True), None), ((None, False), False), ((None, None), None),
]
OR_TRUTH = [  # T/evaluator/CompilerTest.kt:81-91
    ((True, True), True), ((True, False), True), ((True, None), True),
    ((False, True), True), ((False, False), False), ((False, None), None),
    ((None, True), True), ((None, False), None), ((None, None), None),
]


@pytest.mark.parametrize("mode", MODES)
def test_should_handle_null_in_boolean_and_expressions(mode):
    expr = FunctionExpression(Function.AND, [ColumnExpression("p", 0, DataType.BOOLEAN),
                                             ColumnExpression("q", 1, DataType.BOOLEAN)], DataType.BOOLEAN)
    callable_ = compile_expression(expr, mode)
    for row, expected in AND_TRUTH:
        assert callable_(list(row)) == expected and type(callable_(list(row))) is type(expected), row


@pytest.mark.parametrize("mode", MODES)
def test_should_handle_null_in_boolean_or_expressions(mode):
    expr = FunctionExpression(Function.OR, [ColumnExpression("p", 0, DataType.BOOLEAN),
                                            ColumnExpression("q", 1, DataType.BOOLEAN)], DataType.BOOLEAN)
    callable_ = compile_expression(expr, mode)
    for row, expected in OR_TRUTH:
        assert callable_(list(row)) == expected and type(callable_(list(row))) is type(expected), row


@pytest.mark.parametrize("mode", MODES)
def test_should_handle_null_in_if_expressions(mode):
    # T/evaluator/CompilerTest.kt:104-120
    expr = FunctionExpression(Function.IF, [ColumnExpression("cond", 0, DataType.BOOLEAN),
                                            StringLiteralExpression("t"), StringLiteralExpression("f")],
                              DataType.STRING)
    callable_ = compile_expression(expr, mode)
    for row, expected in [((True,), "t"), ((False,), "f"), ((None,), None)]:
        assert callable_(list(row)) == expected


def test_should_group_by_multiple_columns():
    # T/evaluator/QueryTest.kt:14-32.  The SQL string goes through the parser and
    # planner (out of scope); the plan they produce for
    #   SELECT bar, SUM(num), foo FROM table
    # is Projection[_0, SUM#0 -> col 2, _1] <- GroupBy(2 keys, [SUM]) <- Projection[bar, foo, num]
    # (M/evaluator/RewriteAggregates.kt:18-49), built here by hand.
    S, D = DataType.STRING, DataType.DOUBLE
    table = OracleTable.from_rows([S, S, D], [
        ["a", "A", 1.0], ["a", "B", 2.0], ["a", "B", 3.0],
        ["b", "B", 4.0], ["b", "B", None], ["c", None, None]])
    plan = OraclePlan(table, [0, 1, 2],
                      projections=[ColumnExpression("bar", 1, S), ColumnExpression("foo", 0, S),
                                   ColumnExpression("num", 2, D)],
                      aggregates=[AggregationFunction.SUM], group_count=2, mode=Mode.INTERPRETER)
    grouped = plan.rows()
    actual = [[r[0], r[2], r[1]] for r in grouped]          # final projection: bar, SUM(num), foo
    assert actual == [["A", 1.0, "a"], ["B", 5.0, "a"], ["B", 4.0, "b"], [None, None, "c"]]


@pytest.mark.parametrize("mode", MODES)
def test_bytecode_compiler_projection_example(mode):
    # T/ByteCodeCompilerTest.kt:12-40 prints (asserts nothing); values by inspection
    D = DataType.DOUBLE
    table = OracleTable.from_rows([D, D, D], [[10.0, 11.0, 12.0], [20.0, 21.0, 22.0], [30.0, 31.0, 32.0]])
    plan = OraclePlan(table, [0, 1, 2], projections=[
        FunctionExpression(Function.ADD, [ColumnExpression("foo", 0, D), ColumnExpression("bar", 1, D)]),
        FunctionExpression(Function.ADD, [ColumnExpression("bar", 1, D), ColumnExpression("baz", 2, D)])],
        mode=mode)
    assert plan.rows() == [[21.0, 23.0], [41.0, 43.0], [61.0, 63.0]]


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("size", [1000, 1_000_000])
def test_simple_sum_benchmark_known_answer(mode, size):
    # T/SimpleSumBenchmark.java:26,41-45,53: SELECT SUM(foo + 10*bar), foo = bar = (double)(i / 1000).
    # Every partial sum is an integer below 2^53, so any order gives exactly this.
    D = DataType.DOUBLE
    i = np.arange(size, dtype=np.int64)
    v = (i // 1000).astype(np.float64)
    table = OracleTable([D, D], size)
    table.set_column(0, v)
    table.set_column(1, v)
    expr = FunctionExpression(Function.ADD, [
        ColumnExpression("foo", 0, D),
        FunctionExpression(Function.MUL, [NumericLiteralExpression(10.0), ColumnExpression("bar", 1, D)])])
    plan = OraclePlan(table, [0, 1], projections=[expr], aggregates=[AggregationFunction.SUM], mode=mode)
    expected = 11.0 * 1000.0 * sum(range(size // 1000)) if size >= 1000 else 0.0
    assert plan.rows() == [[expected]]
    if size == 1_000_000:
        assert expected == 5_494_500_000.0


# ---- behaviour the reference's tests do not pin: restated from the Java SE spec ----

def _cmp(fn_, a, b, mode=Mode.BYTECODE_COMPILER):
    D = DataType.DOUBLE
    e = FunctionExpression(fn_, [ColumnExpression("a", 0, D), ColumnExpression("b", 1, D)])
    return compile_expression(e, mode)([a, b])


def test_total_order_comparisons_spec():
    nan, inf = float("nan"), float("inf")
    for mode in (Mode.INTERPRETER, Mode.BYTECODE_COMPILER):
        # Double.compare: -0.0 < 0.0, NaN == NaN, NaN > +Inf  (M/evaluator/Interpreter.kt:104-107)
        assert _cmp(Function.CMP_LT, -0.0, 0.0, mode) is True
        assert _cmp(Function.CMP_GE, -0.0, 0.0, mode) is False
        assert _cmp(Function.CMP_GT, nan, inf, mode) is True
        assert _cmp(Function.CMP_LE, nan, nan, mode) is True
        assert _cmp(Function.CMP_LT, 1.0, nan, mode) is True
    # ClosureCompiler.kt:127-130 is IEEE
    assert _cmp(Function.CMP_LT, -0.0, 0.0, Mode.CLOSURE_COMPILER) is False
    assert _cmp(Function.CMP_GT, nan, inf, Mode.CLOSURE_COMPILER) is False
    # equality: all three modes use boxed equals (Interpreter.kt:102-103)
    for mode in MODES:
        assert _cmp(Function.CMP_EQ, nan, nan, mode) is True
        assert _cmp(Function.CMP_EQ, 0.0, -0.0, mode) is False
        assert _cmp(Function.CMP_NE, 0.0, -0.0, mode) is True
        assert _cmp(Function.CMP_EQ, 1.5, None, mode) is None


def test_arithmetic_spec():
    D = DataType.DOUBLE
    a, b = ColumnExpression("a", 0, D), ColumnExpression("b", 1, D)

    def ev(f, x, y):
        return compile_expression(FunctionExpression(f, [a, b]))([x, y])
    assert ev(Function.DIV, 1.0, 0.0) == float("inf")
    assert ev(Function.DIV, -1.0, 0.0) == float("-inf")
    assert math.isnan(ev(Function.DIV, 0.0, 0.0))
    assert ev(Function.MOD, 7.5, 2.0) == 1.5
    assert ev(Function.MOD, -7.5, 2.0) == -1.5          # sign of the dividend (JVM DREM)
    assert math.isnan(ev(Function.MOD, 1.0, 0.0))
    assert ev(Function.SUB, 0.1, 0.3) == 0.1 - 0.3
    neg = compile_expression(FunctionExpression(Function.UNARY_MINUS, [a]))
    assert math.copysign(1.0, neg([0.0])) == -1.0
    assert neg([None]) is None
    plus = compile_expression(FunctionExpression(Function.UNARY_PLUS, [a]))
    assert plus([2.5]) == 2.5


def test_not_null_disagreement():
    B = DataType.BOOLEAN
    e = FunctionExpression(Function.NOT, [ColumnExpression("p", 0, B)])
    assert compile_expression(e, Mode.CLOSURE_COMPILER)([None]) is None     # ClosureCompiler.kt:115
    assert compile_expression(e, Mode.BYTECODE_COMPILER)([None]) is None    # BytecodeCompiler.kt:346-351
    with pytest.raises(EvalError):                                          # Interpreter.kt:92
        compile_expression(e, Mode.INTERPRETER)([None])
    for mode in MODES:
        assert compile_expression(e, mode)([True]) is False
        assert compile_expression(e, mode)([False]) is True


def test_filter_drops_null_predicate_and_keeps_order():
    # M/operator/FilterOperator.kt:20
    D, B = DataType.DOUBLE, DataType.BOOLEAN
    table = OracleTable.from_rows([D, B], [[0.0, True], [1.0, None], [2.0, False], [3.0, True]])
    plan = OraclePlan(table, [0, 1], filter=ColumnExpression("p", 1, B))
    assert plan.rows() == [[0.0, True], [3.0, True]]


def test_accumulators_spec():
    # M/evaluator/Accumulators.kt:26-107
    D = DataType.DOUBLE
    A = AggregationFunction
    table = OracleTable.from_rows([D], [[3.0], [None], [-0.0], [0.0], [7.0]])
    c = ColumnExpression("x", 0, D)
    plan = OraclePlan(table, [0], projections=[c, c, c, c, c], aggregates=[A.MIN, A.MAX, A.SUM, A.COUNT, A.AVG])
    (mn, mx, sm, cnt, avg), = plan.rows()
    assert mn == 0.0 and math.copysign(1.0, mn) == -1.0 and mx == 7.0
    assert sm == 10.0 and cnt == 4 and isinstance(cnt, int) and avg == 2.5
    empty = OracleTable.from_rows([D], [[None], [None]])
    plan = OraclePlan(empty, [0], projections=[c, c, c, c, c], aggregates=[A.MIN, A.MAX, A.SUM, A.COUNT, A.AVG])
    assert plan.rows() == [[None, None, None, 0, None]]
    nan_t = OracleTable.from_rows([D], [[1.0], [float("nan")], [2.0]])
    plan = OraclePlan(nan_t, [0], projections=[c, c], aggregates=[A.MIN, A.MAX])
    (mn, mx), = plan.rows()
    assert math.isnan(mn) and math.isnan(mx)          # Math.min/max propagate NaN


def test_jdk_published_known_answers():
    """The JLS / Javadoc examples for the primitives the reference delegates to the JDK (tests/jdk_vectors.py)."""
    import jdk_vectors as J
    D = DataType.DOUBLE
    a, b = ColumnExpression("a", 0, D), ColumnExpression("b", 1, D)
    for name, x, y, want in J.ARITHMETIC:
        for mode in MODES:
            got = compile_expression(FunctionExpression(Function[name], [a, b]), mode)([x, y])
            assert J.same(got, want), (name, x, y, got, want, mode)
    for name, x, y, want in J.COMPARISONS:
        modes = MODES if name in ("CMP_EQ", "CMP_NE") else (Mode.INTERPRETER, Mode.BYTECODE_COMPILER)   # ClosureCompiler is IEEE on <,<=,>=,>
        for mode in modes:
            got = compile_expression(FunctionExpression(Function[name], [a, b]), mode)([x, y])
            assert got is want, (name, x, y, got, want, mode)
    A = AggregationFunction
    c = ColumnExpression("x", 0, D)
    for values, mn, mx in J.MIN_MAX:
        table = OracleTable.from_rows([D], [[v] for v in values])
        (gmn, gmx), = OraclePlan(table, [0], projections=[c, c], aggregates=[A.MIN, A.MAX]).rows()
        assert J.same(gmn, mn) and J.same(gmx, mx), (values, gmn, gmx)
