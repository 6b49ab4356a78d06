"""Parity of the CUDA path (through the C ABI) against the CPU oracle.  Needs a B200.

Small cases transcribe the reference's own tests
(T/ = src/test/java/net/jhorstmann/queryengine/); the rest are seeded random
trees and tables at sizes the oracle finishes in seconds, plus size-independent
properties at larger sizes.
"""
import numpy as np
import pytest

from exprgen import ExprGen
from parity import TableSpec, compare_aggregation, compare_pipeline, f64_equal, random_table
from queryengine_b200 import capi
from queryengine_b200.ast import (
    AggregationFunction, BooleanLiteralExpression, ColumnExpression, DataType, Function, FunctionExpression, Mode,
    NumericLiteralExpression, StringLiteralExpression, col, fn, lit,
)
from queryengine_b200.engine import ColumnarTable, Context, Field, GpuPipelineOperator, Pipeline

pytestmark = pytest.mark.gpu
D, B, I, T = DataType.DOUBLE, DataType.BOOLEAN, DataType.INT64, DataType.DATE32
A = AggregationFunction


@pytest.fixture(scope="module")
def ctx():
    c = Context(0)
    yield c
    c.close()


def _rows_table(ctx, fields, rows):
    cols, valid = [], []
    for c, f in enumerate(fields):
        vals = [r[c] for r in rows]
        valid.append(np.array([v is not None for v in vals]))
        if f.type == B:
            cols.append(np.array([bool(v) for v in vals]))
        elif f.type == DataType.STRING:
            cols.append(vals)                       # str / None: dictionary-encoded by ColumnarTable.upload
        else:
            cols.append(np.array([0.0 if v is None else v for v in vals], dtype=np.float64))
    return ColumnarTable.from_arrays(ctx, fields, cols, valid)


def _eval_rows(ctx, fields, rows, expr, interpreter):
    """compileExpression(expr, Mode.GPU) applied to each row (RowCallable.invoke)."""
    t = _rows_table(ctx, fields, rows)
    op = GpuPipelineOperator(t, None, [expr], None, interpreter=interpreter)
    out = [r[0] for r in op.map()]
    t.close()
    return out


@pytest.mark.parametrize("interpreter", [False, True])
def test_compilertest_golden_vectors(ctx, interpreter):
    # T/evaluator/CompilerTest.kt:27-35  1.0 + 2.0 == 3.0 (evaluated once per row of a 3-row table)
    e = FunctionExpression(Function.ADD, [NumericLiteralExpression(1.0), NumericLiteralExpression(2.0)], D)
    assert _eval_rows(ctx, [Field("x", D, False)], [[0.0], [1.0], [2.0]], e, interpreter) == [3.0, 3.0, 3.0]
    # :39-47  10.0 * null == null
    e = FunctionExpression(Function.MUL, [ColumnExpression("foo", 0, D), ColumnExpression("bar", 1, D)], D)
    assert _eval_rows(ctx, [Field("foo", D, True), Field("bar", D, True)], [[10.0, None], [10.0, 2.0]], e,
                      interpreter) == [None, 20.0]
    # :52-74 AND truth table, :78-100 OR truth table
    rows = [[p, q] for p in (True, False, None) for q in (True, False, None)]
    fields = [Field("p", B, True), Field("q", B, True)]
    e = FunctionExpression(Function.AND, [ColumnExpression("p", 0, B), ColumnExpression("q", 1, B)], B)
    assert _eval_rows(ctx, fields, rows, e, interpreter) == [True, False, None, False, False, False, None, False, None]
    e = FunctionExpression(Function.OR, [ColumnExpression("p", 0, B), ColumnExpression("q", 1, B)], B)
    assert _eval_rows(ctx, fields, rows, e, interpreter) == [True, True, True, True, False, None, True, None, None]
    # :104-120 IF with a null condition (string literals replaced by doubles: strings are not on the GPU path)
    e = FunctionExpression(Function.IF, [ColumnExpression("cond", 0, B), NumericLiteralExpression(1.0),
                                         NumericLiteralExpression(2.0)], D)
    assert _eval_rows(ctx, [Field("cond", B, True)], [[True], [False], [None]], e, interpreter) == [1.0, 2.0, None]


@pytest.mark.parametrize("interpreter", [False, True])
def test_bytecode_compiler_projection_example(ctx, interpreter):
    # T/ByteCodeCompilerTest.kt:12-40
    fields = [Field("foo", D, False), Field("bar", D, False), Field("baz", D, False)]
    t = _rows_table(ctx, fields, [[10.0, 11.0, 12.0], [20.0, 21.0, 22.0], [30.0, 31.0, 32.0]])
    op = GpuPipelineOperator(t, None, [fn(Function.ADD, col(0), col(1)), fn(Function.ADD, col(1), col(2))], None,
                             interpreter=interpreter)
    assert op.map() == [[21.0, 23.0], [41.0, 43.0], [61.0, 63.0]]


@pytest.mark.parametrize("size", [1000, 1_000_000])
def test_simple_sum_benchmark_known_answer(ctx, size):
    # T/SimpleSumBenchmark.java:41-53  SELECT SUM(foo + 10*bar); exact in any summation order
    v = (np.arange(size) // 1000).astype(np.float64)
    t = ColumnarTable.from_arrays(ctx, [Field("foo", D, False), Field("bar", D, False)], [v, v])
    e = fn(Function.ADD, col(0), fn(Function.MUL, 10.0, col(1)))
    op = GpuPipelineOperator(t, None, [e], [A.SUM])
    assert op.map() == [[11.0 * 1000.0 * sum(range(size // 1000))]]
    if size == 1_000_000:
        assert op.map() == [[5_494_500_000.0]]
    t.close()


@pytest.mark.parametrize("interpreter", [False, True])
def test_jdk_published_known_answers(ctx, interpreter):
    """The same JLS / Javadoc examples (tests/jdk_vectors.py) through the GPU path: generated code and interpreter kernels,
    operands from columns and (for the second operand) as literals, MIN / MAX through the aggregation kernel."""
    import jdk_vectors as J
    fields = [Field("a", D, True), Field("b", D, True)]
    for table_, literal_rhs in ((J.ARITHMETIC + J.COMPARISONS, False), (J.ARITHMETIC + J.COMPARISONS, True)):
        for name, x, y, want in table_:
            if literal_rhs and y != y:
                continue                                  # a NaN literal is not something the grammar can write
            rhs = NumericLiteralExpression(y) if literal_rhs else col(1)
            got, = _eval_rows(ctx, fields, [[x, y]], FunctionExpression(Function[name], [col(0), rhs]), interpreter)
            assert J.same(got, want), (name, x, y, got, want, literal_rhs)
    for values, mn, mx in J.MIN_MAX:
        t = _rows_table(ctx, [Field("x", D, True)], [[v] for v in values])
        op = GpuPipelineOperator(t, None, [col(0), col(0)], [A.MIN, A.MAX], interpreter=interpreter)
        (gmn, gmx), = op.map()
        assert J.same(gmn, mn) and J.same(gmx, mx), (values, gmn, gmx)
        t.close()


def test_filter_returns_row_unchanged_and_drops_null_predicate(ctx):
    # M/operator/FilterOperator.kt:14-25
    fields = [Field("x", D, False), Field("p", B, True)]
    t = _rows_table(ctx, fields, [[0.0, True], [1.0, None], [2.0, False], [3.0, True]])
    op = GpuPipelineOperator(t, ColumnExpression("p", 1, B), None, None)
    assert op.map() == [[0.0, True], [3.0, True]]
    assert op.next() is None          # exhausted and closed
    t.close()


@pytest.mark.parametrize("nrows", [0, 1, 63, 64, 65, 511, 4096, 4097, 8191, 20000])
@pytest.mark.parametrize("interpreter", [False, True])
def test_single_column_filter_all_sizes(ctx, nrows, interpreter):
    rng = np.random.default_rng(nrows + 7)
    spec = random_table(rng, nrows, "d")
    compare_pipeline(ctx, spec, fn(Function.CMP_LT, col(0), 0.0), [col(0)], interpreter=interpreter)


@pytest.mark.parametrize("k", [-np.inf, -1000.0, -0.0, 0.0, 5e-324, np.inf, np.nan])
def test_compare_against_extreme_literals(ctx, k):
    rng = np.random.default_rng(11)
    spec = random_table(rng, 5000, "dD", special_rate=0.3)
    for f in (Function.CMP_LT, Function.CMP_LE, Function.CMP_GE, Function.CMP_GT, Function.CMP_EQ, Function.CMP_NE):
        compare_pipeline(ctx, spec, fn(f, col(0), float(k)), [fn(f, float(k), col(1))])


@pytest.mark.parametrize("k", [-3.5, -3.0, -0.0, 0.0, 2.0, 2.5, 9000.0, 1e18, 1e19, -1e19, np.inf, -np.inf, np.nan])
def test_integer_columns_against_literals(ctx, k):
    # INT64 / DATE32 columns compare in the integer domain when the literal allows it; same answers
    # as widening to DOUBLE first (the oracle sees the widened values)
    rng = np.random.default_rng(17)
    spec = random_table(rng, 3000, "iIt")
    for f in (Function.CMP_LT, Function.CMP_LE, Function.CMP_GE, Function.CMP_GT, Function.CMP_EQ, Function.CMP_NE):
        compare_pipeline(ctx, spec, fn(f, col(0, I), float(k)),
                         [fn(f, float(k), col(1, I)), fn(f, col(2, T), float(k) + 9000.0), col(0, I)])


@pytest.mark.parametrize("seed", range(12))
def test_random_trees_jit(ctx, seed):
    rng = np.random.default_rng(1000 + seed)
    layout = ["dDbB", "DDDD", "dibT", "DBIt", "dd", "BB d".replace(" ", "")][seed % 6]
    spec = random_table(rng, int(rng.integers(1, 30000)), layout)
    g = ExprGen(rng, spec.fields)
    filt = g.boolean(4) if seed % 4 != 3 else None
    projs = [g.double(5), g.boolean(3), g.double(2)]
    if seed % 3 == 0:
        projs.append(ColumnExpression(spec.fields[0].name, 0, spec.fields[0].type))
    compare_pipeline(ctx, spec, filt, projs, interpreter=False)


@pytest.mark.parametrize("seed", range(6))
def test_random_trees_interpreter(ctx, seed):
    rng = np.random.default_rng(2000 + seed)
    spec = random_table(rng, int(rng.integers(1, 5000)), ["dDbB", "DIbt", "DDBB"][seed % 3])
    g = ExprGen(rng, spec.fields)
    filt = g.boolean(4) if seed % 3 != 2 else None
    compare_pipeline(ctx, spec, filt, [g.double(5), g.boolean(3)], interpreter=True)


# ---- Mode.CLOSURE_COMPILER: IEEE ordering comparisons (QE_CMP_IEEE), M/evaluator/ClosureCompiler.kt:127-130 ------------------

@pytest.mark.parametrize("interpreter", [False, True])
def test_closure_compiler_mode_orders_like_ieee(ctx, interpreter):
    """The one place the reference's CLOSURE_COMPILER mode differs from its other two on well-typed trees:
    `<, <=, >=, >` unbox to primitive doubles, so a NaN compares false and -0.0 equals 0.0, while `=` / `<>` stay boxed
    equals (ClosureCompiler.kt:125-130).  Known answers of IEEE 754 / JLS 15.20.1, and the same rows in the default mode."""
    nan, inf = float("nan"), float("inf")
    fields = [Field("a", D, True), Field("b", D, True)]
    rows = [[nan, 1.0], [1.0, nan], [nan, nan], [-0.0, 0.0], [0.0, -0.0], [nan, inf], [inf, nan], [1.0, 2.0], [None, 1.0]]
    F = Function
    ieee = {F.CMP_LT: [False, False, False, False, False, False, False, True, None],
            F.CMP_LE: [False, False, False, True, True, False, False, True, None],
            F.CMP_GE: [False, False, False, True, True, False, False, False, None],
            F.CMP_GT: [False, False, False, False, False, False, False, False, None],
            F.CMP_EQ: [False, False, True, False, False, False, False, False, None],     # boxed equals in every mode
            F.CMP_NE: [True, True, False, True, True, True, True, True, None]}
    total = {F.CMP_LT: [False, True, False, True, False, False, True, True, None],
             F.CMP_GE: [True, False, True, False, True, True, False, False, None]}
    t = _rows_table(ctx, fields, rows)
    for f, want in ieee.items():
        op = GpuPipelineOperator(t, None, [fn(f, col(0), col(1))], None, interpreter=interpreter, mode=Mode.CLOSURE_COMPILER)
        assert [r[0] for r in op.map()] == want, f
    for f, want in total.items():
        op = GpuPipelineOperator(t, None, [fn(f, col(0), col(1))], None, interpreter=interpreter)
        assert [r[0] for r in op.map()] == want, f
    # both orderings of one comparison in one pipeline stay two values (the code generator's common-subexpression key)
    from queryengine_b200.program import QE_CMP_IEEE, serialize
    w_ieee = serialize(fn(F.CMP_GE, col(0), col(1)), ieee_order=True)
    assert int(w_ieee[-1]) == int(F.CMP_GE) | (QE_CMP_IEEE << 8)
    t.close()


@pytest.mark.parametrize("k", [-np.inf, -0.0, 0.0, 2.5, np.inf, np.nan])
def test_closure_compiler_mode_against_literals(ctx, k):
    rng = np.random.default_rng(12)
    spec = random_table(rng, 5000, "dDiIt", special_rate=0.3)
    for f in (Function.CMP_LT, Function.CMP_LE, Function.CMP_GE, Function.CMP_GT, Function.CMP_EQ, Function.CMP_NE):
        for interpreter in (False, True):
            compare_pipeline(ctx, spec, fn(f, col(0), float(k)),
                             [fn(f, float(k), col(1)), fn(f, col(2, I), float(k)), fn(f, float(k), col(3, I)),
                              fn(f, col(4, T), float(k) + 9000.0)], interpreter=interpreter, mode=Mode.CLOSURE_COMPILER)


@pytest.mark.parametrize("seed", range(8))
def test_random_trees_closure_compiler_mode(ctx, seed):
    rng = np.random.default_rng(3000 + seed)
    layout = ["dDbB", "DDDD", "dibT", "DBIt"][seed % 4]
    spec = random_table(rng, int(rng.integers(1, 20000)), layout, special_rate=0.25)
    g = ExprGen(rng, spec.fields)
    filt = g.boolean(4) if seed % 4 != 3 else None
    compare_pipeline(ctx, spec, filt, [g.double(5), g.boolean(4), g.boolean(3)], interpreter=seed % 2 == 1,
                     mode=Mode.CLOSURE_COMPILER)


@pytest.mark.parametrize("steps,vec,staging", [(1, 2, 0), (2, 2, 0), (8, 2, 0), (1, 4, 0), (4, 4, 0), (8, 4, 0),
                                               (1, 2, 1), (2, 2, 1), (4, 2, 1), (8, 2, 1)])
def test_every_tile_shape(ctx, steps, vec, staging):
    c = Context(0)
    c.set_option("steps", steps)
    c.set_option("vec", vec)
    c.set_option("staging", staging)
    c.set_option("pipeline", 0 if (steps == 2 and vec == 2) or (steps == 4 and vec == 4) else 1)
    if staging and steps == 2:
        c.set_option("stages", 4)
    rng = np.random.default_rng(steps * 10 + vec)
    spec = random_table(rng, 30011, "DbiT")
    g = ExprGen(rng, spec.fields)
    compare_pipeline(c, spec, g.boolean(3), [g.double(4), g.boolean(2), col(2, I), col(3, T)])
    compare_pipeline(c, spec, None, [g.double(4), g.boolean(2), col(2, I)])
    finite = random_table(rng, 30011, "DbiT", special_rate=0.0)        # sums of Inf/NaN depend on the order
    compare_aggregation(c, finite, g.boolean(2), [fn(Function.MUL, col(0), col(2, I)), col(0)], [A.SUM, A.MAX])
    c.close()


@pytest.mark.parametrize("ncw,stage_kb,nrows", [(0, 0, 30011), (4, 4, 1), (8, 16, 4096), (20, 0, 100003), (12, 48, 65536)])
def test_dense_projection_through_the_ring_of_stages(ctx, ncw, stage_kb, nrows):
    """qe_project_ring_kernel ("staging" = 2 without a filter): bulk-copied stages, one row per lane and step, dense
    value stores and validity / BOOLEAN words by ballot; every column type, nullable inputs, ragged tails."""
    c = Context(0)
    c.set_option("staging", 2)
    c.set_option("ncw", ncw)
    c.set_option("stage_kb", stage_kb)
    rng = np.random.default_rng(77 + ncw)
    spec = random_table(rng, nrows, "DbiTdB")
    g = ExprGen(rng, spec.fields)
    compare_pipeline(c, spec, None, [g.double(5), g.boolean(3), col(2, I), col(3, T), g.double(3), col(1, B)])
    c.close()


@pytest.mark.parametrize("stage_kb,ncw,defer,stages,nlb", [(0, 0, 0, 0, 0), (4, 4, 1, 2, 1), (8, 8, 1, 3, 3), (16, 16, 3, 5, 2),
                                                           (48, 8, 2, 4, 4), (2, 12, 9, 16, 8)])
def test_every_ring_shape(ctx, stage_kb, ncw, defer, stages, nlb):
    """The warp-specialised ring variant of the filter kernel (the default) over its option space."""
    c = Context(0)
    c.set_option("staging", 2)
    for k, v in (("stage_kb", stage_kb), ("ncw", ncw), ("defer", defer), ("stages", stages), ("nlb", nlb)):
        c.set_option(k, v)
    rng = np.random.default_rng(stage_kb * 100 + ncw)
    for n in (0, 1, 127, 128, 129, 4095, 4096, 4097, 100003, 1_000_003):
        v = rng.integers(0, 1000, n).astype(np.int64)
        t = ColumnarTable(c, [Field("col", I, False)], n)
        t.upload(0, v)
        for k in (0.0, 10.0, 500.0, 990.0, 1000.0):
            res = Pipeline(t, fn(Function.CMP_LT, col(0, I), k), [col(0, I)]).run()
            want = np.nonzero(v < k)[0].astype(np.uint32)
            assert res.num_rows == want.size
            assert np.array_equal(res.selection(), want)
            assert np.array_equal(res.column(0)[0], v[want])
            res.close()
        t.close()
    spec = random_table(rng, 70001, "DbiT")
    g = ExprGen(rng, spec.fields)
    compare_pipeline(c, spec, g.boolean(3), [g.double(4), g.boolean(2), col(2, I), col(3, T)])
    compare_pipeline(c, spec, g.boolean(4), [col(0), col(1, spec.fields[1].type)])
    c.close()


@pytest.mark.parametrize("list_cap,nsw", [(0, 0), (-1, 0), (32, -1), (32, 1), (32, 4), (64, 2), (128, 3)])
def test_sparse_tiles_row_lists_and_sweeper_warps(ctx, list_cap, nsw):
    """Ring filter, sparse tiles: a consumer warp that kept at most `list_cap` rows of a tile leaves their row numbers as a
    packed list in shared memory; `nsw` sweeper warps (or, with nsw = -1, the warp itself) store them once the tile's
    prefix is known, the other warps keep the mask paths.  Same rows, same order as the masks-only kernel (list_cap = -1)
    and the oracle, over selectivities that mix both kinds of warps in one tile, with and without selection vector /
    values, appended batches, and an output capacity that ends inside a swept list."""
    c = Context(0)
    c.set_option("list_cap", list_cap)
    c.set_option("nsw", nsw)
    rng = np.random.default_rng(77 + list_cap + nsw)
    n = 300_007
    x = rng.integers(0, 100_000, n).astype(np.int64)
    y = rng.uniform(-1.0, 1.0, n)
    yv = rng.random(n) < 0.9
    b = rng.random(n) < 0.7
    d = rng.integers(8000, 9000, n).astype(np.int32)
    fields = [Field("x", I, False), Field("y", D, True), Field("b", B, False), Field("d", T, False)]
    t = ColumnarTable.from_arrays(c, fields, [x, y, b, d], [None, yv, None, None])
    for frac in (0.0, 0.002, 0.01, 0.03, 0.1, 0.15, 0.3, 1.0):
        k = float(int(frac * 100_000))
        # one 8-byte column: the benchmark's shape (20 consumer warps)
        want = np.nonzero(x < k)[0].astype(np.uint32)
        res = Pipeline(t, fn(Function.CMP_LT, col(0, I), k), [col(0, I)]).run()
        assert res.num_rows == want.size and np.array_equal(res.selection(), want) and np.array_equal(res.column(0)[0], x[want])
        res = Pipeline(t, fn(Function.CMP_LT, col(0, I), k), [col(0, I)]).run(res)        # reused: shape by the measured selectivity
        assert res.num_rows == want.size and np.array_equal(res.selection(), want) and np.array_equal(res.column(0)[0], x[want])
        res.close()
        only = Pipeline(t, fn(Function.CMP_LT, col(0, I), k), [col(0, I)], values=False).run()
        assert only.num_rows == want.size and np.array_equal(only.selection(), want)
        only.close()
        nosel = Pipeline(t, fn(Function.CMP_LT, col(0, I), k), [col(0, I)], selection=False).run()
        assert nosel.num_rows == want.size and np.array_equal(nosel.column(0)[0], x[want])
        nosel.close()
        # nullable and bit inputs (read at arbitrary rows by the list paths), computed non-null outputs, 4-byte column
        filt = fn(Function.AND, fn(Function.AND, fn(Function.CMP_LT, col(0, I), k * 2.0), fn(Function.CMP_GT, col(1), -0.5)), col(2, B))
        keep = (x < k * 2.0) & yv & (y > -0.5) & b
        want = np.nonzero(keep)[0].astype(np.uint32)
        res = Pipeline(t, filt, [col(0, I), col(3, T), fn(Function.ADD, fn(Function.MUL, col(0, I), 0.5), col(3, T))]).run()
        assert res.num_rows == want.size and np.array_equal(res.selection(), want)
        assert np.array_equal(res.column(0)[0], x[want]) and np.array_equal(res.column(1)[0], d[want])
        assert np.array_equal(res.column(2)[0], x[want] * 0.5 + d[want])
        res.close()
        # outputs the filter has made non-null (y is read by the predicate): validity = runs of ones, written per tile
        res = Pipeline(t, filt, [col(1), fn(Function.MUL, col(1), 2.0), col(0, I), fn(Function.CMP_GT, col(1), 0.0)]).run()
        assert res.num_rows == want.size and np.array_equal(res.selection(), want)
        for kk, ref in ((0, y[want]), (1, y[want] * 2.0), (2, x[want]), (3, y[want] > 0.0)):
            vv, nn = res.column(kk)
            assert nn.all() and np.array_equal(vv, ref), kk
        res.close()
        # nullable and BOOLEAN outputs: compacted bit columns (listed rows: one OR per set bit by the sweepers)
        filt2 = fn(Function.AND, fn(Function.CMP_LT, col(0, I), k * 1.5), col(2, B))
        keep2 = (x < k * 1.5) & b
        want2 = np.nonzero(keep2)[0].astype(np.uint32)
        res = Pipeline(t, filt2, [col(1), col(2, B), fn(Function.MUL, col(1), 2.0), fn(Function.CMP_GT, col(1), 0.0), col(0, I)]).run()
        assert res.num_rows == want2.size and np.array_equal(res.selection(), want2)
        v0, n0 = res.column(0)
        assert np.array_equal(n0, yv[want2]) and np.array_equal(v0[n0], y[want2][n0])
        v1, n1 = res.column(1)
        assert n1.all() and np.array_equal(v1, b[want2])
        v2, n2 = res.column(2)
        assert np.array_equal(n2, yv[want2]) and np.array_equal(v2[n2], (y[want2] * 2.0)[n2])
        v3, n3 = res.column(3)
        assert np.array_equal(n3, yv[want2]) and np.array_equal(v3[n3], (y[want2] > 0.0)[n3])
        assert np.array_equal(res.column(4)[0], x[want2])
        res.close()
        # capacity: exactly enough, and five rows short (the last list is cut, the overflow is reported)
        if want.size > 5:
            res = Pipeline(t, filt, [col(0, I)], out_capacity_rows=int(want.size)).run()
            assert res.num_rows == want.size and np.array_equal(res.column(0)[0], x[want])
            res.close()
            res = Pipeline(t, filt, [col(0, I)], out_capacity_rows=int(want.size) - 5).run()
            with pytest.raises(capi.QeError) as ei:
                _ = res.num_rows
            assert ei.value.status == capi.QE_ERR_CAPACITY
            res.close()
    # appended batches (the running count is the base of the next batch's lists)
    k = 1500.0
    want = np.nonzero(x < k)[0].astype(np.uint32)
    pipe = Pipeline(t, fn(Function.CMP_LT, col(0, I), k), [col(0, I)])
    res = pipe.run(row_begin=0, row_end=100_096)
    res = pipe.run_append(res, 100_096, 200_064)
    res = pipe.run_append(res, 200_064, n)
    assert res.num_rows == want.size and np.array_equal(res.selection(), want) and np.array_equal(res.column(0)[0], x[want])
    res.close()
    t.close()
    # and against the oracle, with the specials
    spec = random_table(rng, 50_021, "dDbi", special_rate=0.1)
    compare_pipeline(c, spec, fn(Function.AND, fn(Function.CMP_LT, col(0), -950.0), col(2, B)), [col(0), col(3, I), fn(Function.SUB, col(0), 1.0)])
    compare_pipeline(c, spec, fn(Function.CMP_GT, col(1), 900.0), [col(0), col(3, I)])
    compare_pipeline(c, spec, fn(Function.CMP_GT, col(1), 900.0), [col(1), col(2, B), fn(Function.DIV, col(1), col(0))])
    c.close()


def test_single_narrow_columns_under_every_remembered_selectivity(ctx):
    """One 4-byte column (DATE32), one bit column (BOOLEAN, nullable) and one 8-byte column, each filtered three times into
    one reused result at selectivities on both sides of every shape threshold (known-sparse up to 8 %, lists up to 13.5 %,
    20 / 16 consumer warps at 40 %): the tile sizes these shapes derive for narrow rows (7 680-row tiles of a 4-byte column,
    u16 list entries) are run at least once each."""
    rng = np.random.default_rng(99)
    n = 3_000_017
    d = rng.integers(0, 100_000, n).astype(np.int32)
    b = rng.random(n) < 0.03
    bv = rng.random(n) < 0.9
    x = rng.uniform(0.0, 1.0, n)
    t = ColumnarTable.from_arrays(ctx, [Field("d", T, False), Field("b", B, True), Field("x", D, False)], [d, b, x], [None, bv, None])
    for frac in (0.002, 0.05, 0.1, 0.3, 0.9):
        k = float(int(frac * 100_000))
        want = np.nonzero(d < k)[0].astype(np.uint32)
        res = None
        for rep in range(3):
            res = Pipeline(t, fn(Function.CMP_LT, col(0, T), k), [col(0, T)]).run(res)
            assert res.num_rows == want.size and np.array_equal(res.selection(), want) and np.array_equal(res.column(0)[0], d[want]), (frac, rep)
        res.close()
        want = np.nonzero(x < frac)[0].astype(np.uint32)
        res = None
        for rep in range(3):
            res = Pipeline(t, fn(Function.CMP_LT, col(2), float(frac)), [col(2)]).run(res)
            assert res.num_rows == want.size and np.array_equal(res.selection(), want) and np.array_equal(res.column(0)[0], x[want]), (frac, rep)
        res.close()
    want = np.nonzero(b & bv)[0].astype(np.uint32)                  # the column is its own predicate: null drops the row
    res = None
    for rep in range(3):
        res = Pipeline(t, col(1, B), [col(1, B)]).run(res)
        vals, valid = res.column(0)
        assert res.num_rows == want.size and np.array_equal(res.selection(), want) and vals.all() and valid.all(), rep
    res.close()
    res = None
    for rep in range(3):                                            # ... and its complement (most rows kept)
        res = Pipeline(t, fn(Function.NOT, col(1, B)), [col(1, B), col(0, T)]).run(res)
        want = np.nonzero(~b & bv)[0].astype(np.uint32)
        vals, valid = res.column(0)
        assert res.num_rows == want.size and np.array_equal(res.selection(), want) and not vals.any() and valid.all(), rep
        assert np.array_equal(res.column(1)[0], d[want])
    res.close()
    t.close()


def test_sparse_path_stress_random_shapes_and_selectivities(ctx):
    """compute-sanitizer is closed on this pool, so the hand-offs of the ring kernel (producer / publisher / look-back /
    sweepers / consumers over five mbarrier arrays) are exercised the blunt way: many tiles per CTA, random shapes and
    selectivities, repeated runs into one reused result (every run picks its shape from the selectivity the last one
    measured), each compared with numpy -- a lost or doubled row, a stale list or a stage released too early shows."""
    rng = np.random.default_rng(2024)
    n = 6_000_011                                       # ~1 300 tiles of 4 480 rows: nine per CTA
    x = rng.integers(0, 1 << 20, n).astype(np.int64)
    y = rng.uniform(-1.0, 1.0, n)
    yv = rng.random(n) < 0.85
    for trial in range(6):
        c = Context(0)
        opts = {"list_cap": int(rng.choice([0, 0, 32, 64, -1])), "nsw": int(rng.choice([0, 0, 1, 3, 4, -1])),
                "stages": int(rng.choice([0, 0, 3, 4, 5])), "ncw": int(rng.choice([0, 0, 8, 16, 24])), "nlb": int(rng.choice([0, 0, 2]))}
        for k_, v_ in opts.items():
            c.set_option(k_, v_)
        t = ColumnarTable.from_arrays(c, [Field("x", I, False), Field("y", D, True)], [x, y], [None, yv])
        res = None
        res2 = None
        for rep in range(8):
            frac = float(rng.choice([0.0005, 0.004, 0.01, 0.03, 0.07, 0.12, 0.2, 0.6, 1.0]))
            k = float(int(frac * (1 << 20)))
            want = np.nonzero(x < k)[0].astype(np.uint32)
            if rep % 3 == 2 and res is not None:           # a fresh result: its shape hint comes from the context's memo of the query
                res.close()
                res = None
            res = Pipeline(t, fn(Function.CMP_LT, col(0, I), k), [col(0, I)]).run(res)
            assert res.num_rows == want.size, (opts, frac)
            assert np.array_equal(res.selection(), want) and np.array_equal(res.column(0)[0], x[want]), (opts, frac)
            # nullable payload + a nullable column the predicate reads: bit columns by runs, by the sweepers, by the masks
            keep = (x < k) & yv & (y < 0.5)
            want2 = np.nonzero(keep)[0].astype(np.uint32)
            res2 = Pipeline(t, fn(Function.AND, fn(Function.CMP_LT, col(0, I), k), fn(Function.CMP_LT, col(1), 0.5)),
                            [col(1), fn(Function.CMP_GT, col(1), 0.0)]).run(res2)
            assert res2.num_rows == want2.size, (opts, frac)
            v0, n0 = res2.column(0)
            v1, n1 = res2.column(1)
            assert np.array_equal(res2.selection(), want2) and n0.all() and n1.all(), (opts, frac)
            assert np.array_equal(v0, y[want2]) and np.array_equal(v1, y[want2] > 0.0), (opts, frac)
        res.close()
        res2.close()
        t.close()
        c.close()


@pytest.mark.parametrize("interpreter", [False, True])
def test_aggregates(ctx, interpreter):
    rng = np.random.default_rng(5)
    spec = random_table(rng, 50000, "DdB", special_rate=0.0)
    e = fn(Function.MUL, col(0), col(1))
    compare_aggregation(ctx, spec, fn(Function.CMP_GT, col(1), -500.0), [e, e, e, e, e, col(2, B)],
                        [A.MIN, A.MAX, A.SUM, A.COUNT, A.AVG, A.COUNT], interpreter=interpreter)
    # no row passes: SUM/AVG/MIN/MAX null, COUNT 0 (M/evaluator/Accumulators.kt:33-35,47-53)
    out = compare_aggregation(ctx, spec, fn(Function.CMP_GT, col(1), 1e9), [e, e, e, e, e],
                              [A.MIN, A.MAX, A.SUM, A.COUNT, A.AVG], interpreter=interpreter)
    assert out == [None, None, None, 0.0, None]
    # NaN and signed zeros through MIN / MAX (Math.min / Math.max)
    spec = random_table(rng, 3000, "D", special_rate=0.5)
    compare_aggregation(ctx, spec, None, [col(0), col(0)], [A.MIN, A.MAX], interpreter=interpreter)
    zeros = TableSpec([Field("z", D, False)], [np.array([0.0, -0.0, 0.0])], [None])
    mn, mx = compare_aggregation(ctx, zeros, None, [col(0), col(0)], [A.MIN, A.MAX], interpreter=interpreter)
    assert np.signbit(mn) and not np.signbit(mx)


def test_q6_shaped_query(ctx):
    # BASELINE.json configs[2] at a size the oracle handles: filter + SUM(extendedprice * discount)
    n = 200_000
    rng = np.random.default_rng(42)
    qty = rng.integers(1, 51, n).astype(np.float64)
    disc = rng.integers(0, 11, n) / 100.0
    ep = np.round(qty * rng.uniform(900, 2100, n) * 100) / 100
    sd = rng.integers(8036, 10562, n).astype(np.float64)
    spec = TableSpec([Field("l_quantity", D, False), Field("l_extendedprice", D, False), Field("l_discount", D, False),
                      Field("l_shipdate", D, False)], [qty, ep, disc, sd], [None] * 4)
    filt = fn(Function.AND, fn(Function.AND, fn(Function.AND, fn(Function.AND,
              fn(Function.CMP_GE, col(3), 8766.0), fn(Function.CMP_LT, col(3), 9131.0)),
              fn(Function.CMP_GE, col(2), 0.05)), fn(Function.CMP_LE, col(2), 0.07)), fn(Function.CMP_LT, col(0), 24.0))
    compare_aggregation(ctx, spec, filt, [fn(Function.MUL, col(1), col(2))], [A.SUM])
    compare_pipeline(ctx, spec, filt, [fn(Function.MUL, col(1), col(2))])
    # date32 variant of the ship date: same rows
    spec32 = TableSpec(spec.fields[:3] + [Field("l_shipdate", T, False)], [qty, ep, disc, sd.astype(np.int32)], [None] * 4)
    compare_pipeline(ctx, spec32, filt, [fn(Function.MUL, col(1), col(2)), col(3, T)])


def test_generators_match_their_published_formulas(ctx):
    # qe_table_generate: value(i) = f(splitmix64(seed ^ i)); reproduced here in numpy
    def splitmix64(x):
        x = (x + np.uint64(0x9E3779B97F4A7C15))
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return x ^ (x >> np.uint64(31))
    n, seed, off = 10007, 42, 1 << 33
    t = ColumnarTable(ctx, [Field("a", I, False), Field("b", D, True), Field("c", D, False), Field("d", T, False)], n)
    t.generate(0, capi.QE_GEN_UNIFORM_U40, seed, row_offset=off)
    t.generate(1, capi.QE_GEN_UNIFORM_REAL, seed + 1, -1000.0, 1000.0, null_rate=0.1, row_offset=off)
    t.generate(2, capi.QE_GEN_CENTS, seed + 2, 0, 10, row_offset=off)
    t.generate(3, capi.QE_GEN_UNIFORM_INT, seed + 3, 8036, 10561, row_offset=off)
    res = Pipeline(t, None, [col(0, I), col(1), col(2), col(3, T)]).run()
    with np.errstate(over="ignore"):
        i = np.arange(n, dtype=np.uint64) + np.uint64(off)
        h = splitmix64(np.uint64(seed) ^ i)
        assert np.array_equal(res.column(0)[0], (h >> np.uint64(24)).astype(np.int64))
        h1 = splitmix64(np.uint64(seed + 1) ^ i)
        exp1 = -1000.0 + ((h1 >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)) * 2000.0
        vals, valid = res.column(1)
        vh = splitmix64(~np.uint64(seed + 1) ^ (i * np.uint64(0x9E3779B97F4A7C15)))
        exp_valid = vh >= np.uint64(int(0.1 * 18446744073709551616.0))
        assert np.array_equal(valid, exp_valid) and 0.05 < 1 - valid.mean() < 0.15
        assert f64_equal(vals[valid], exp1[valid]).all()
        h2 = splitmix64(np.uint64(seed + 2) ^ i)
        assert np.array_equal(res.column(2)[0], (h2 % np.uint64(11)).astype(np.float64) / 100.0)
        h3 = splitmix64(np.uint64(seed + 3) ^ i)
        assert np.array_equal(res.column(3)[0], (8036 + (h3 % np.uint64(2526)).astype(np.int64)).astype(np.int32))
    res.close()
    t.close()


def test_capacity_overflow_is_reported_not_written(ctx):
    v = np.arange(100000, dtype=np.float64)
    t = ColumnarTable.from_arrays(ctx, [Field("x", D, False)], [v])
    res = Pipeline(t, fn(Function.CMP_GE, col(0), 0.0), [col(0)], out_capacity_rows=1000).run()
    with pytest.raises(capi.QeError) as ei:
        _ = res.num_rows
    assert ei.value.status == capi.QE_ERR_CAPACITY
    res2 = Pipeline(t, fn(Function.CMP_LT, col(0), 1000.0), [col(0)], out_capacity_rows=1000).run()
    assert res2.num_rows == 1000 and np.array_equal(res2.column(0)[0], v[:1000])
    t.close()


def test_row_range_batches_concatenate_to_the_whole(ctx):
    # MemorySourceOperator scans in order; batches are row ranges of the same scan
    rng = np.random.default_rng(3)
    n = 100_000
    v = rng.uniform(-1, 1, n)
    t = ColumnarTable.from_arrays(ctx, [Field("x", D, False)], [v])
    pipe = Pipeline(t, fn(Function.CMP_LT, col(0), 0.25), [col(0)])
    whole = pipe.run()
    sel_whole = whole.selection()
    parts = []
    for b in range(0, n, 32768):
        r = pipe.run(row_begin=b, row_end=min(n, b + 32768))
        parts.append(r.selection())
        r.close()
    assert np.array_equal(np.concatenate(parts), sel_whole)
    assert np.array_equal(sel_whole, np.nonzero(v < 0.25)[0].astype(np.uint32))
    t.close()


def test_large_scan_properties(ctx):
    """BASELINE.json configs[1] shape at 64 Mi rows: properties that hold at any size."""
    n = 1 << 26
    t = ColumnarTable(ctx, [Field("col", I, False)], n)
    t.generate(0, capi.QE_GEN_UNIFORM_U40, 42)
    expect_total = None
    for s in (0.01, 0.5, 0.99):
        k = float(int(s * 2 ** 40))
        res = Pipeline(t, fn(Function.CMP_LT, col(0, I), k), [col(0, I)]).run()
        assert res.used_jit
        m = res.num_rows
        assert abs(m / n - s) < 1e-3
        sel = res.selection()
        vals, _ = res.column(0)
        assert (np.diff(sel.astype(np.int64)) > 0).all()                  # sorted, unique
        assert (vals < k).all()                                           # every kept row passes
        # idempotence: complement filter keeps exactly the rest
        res2 = Pipeline(t, fn(Function.CMP_GE, col(0, I), k), [col(0, I)]).run()
        assert res2.num_rows == n - m
        # checksum of checksums: kept + dropped values add up to the column total (exact in int64 wraparound)
        total = int(vals.sum(dtype=np.int64)) + int(res2.column(0)[0].sum(dtype=np.int64))
        expect_total = total if expect_total is None else expect_total
        assert total == expect_total
        res.close()
        res2.close()
    t.close()


@pytest.mark.parametrize("interpreter", [False, True])
def test_append_batches_equal_one_run(ctx, interpreter):
    """QE_RUN_APPEND: row-range batches appended into one result = the whole table in one run (values, selection,
    compacted validity / boolean bitmaps; dense projections too)."""
    rng = np.random.default_rng(77)
    spec = random_table(rng, 50_000, "DbiT")
    g = ExprGen(rng, spec.fields)
    t = spec.gpu(ctx)
    filt = g.boolean(3)
    projs = [g.double(4), g.boolean(2), col(2, I), col(3, T)]
    for f in (filt, None):
        whole = Pipeline(t, f, projs, interpreter=interpreter).run()
        pipe = Pipeline(t, f, projs, interpreter=interpreter, out_capacity_rows=spec.nrows)
        bounds = [0, 128, 128, 4224, 20096, 20096 + 8192, spec.nrows]
        res = pipe.run(row_begin=0, row_end=bounds[1])
        for lo, hi in zip(bounds[1:], bounds[2:]):
            res = pipe.run_append(res, lo, hi)
        assert res.num_rows == whole.num_rows
        if f is not None:
            assert np.array_equal(res.selection(), whole.selection())
        for k in range(len(projs)):
            a, av = res.column(k)
            b, bv = whole.column(k)
            assert np.array_equal(av, bv)
            a, b = np.asarray(a)[bv], np.asarray(b)[bv]
            assert f64_equal(a, b).all() if a.dtype == np.float64 else np.array_equal(a, b)
        res.close()
        whole.close()
    t.close()


@pytest.mark.parametrize("chunk", [128, 4096, 1 << 16])
def test_stream_host_in_host_out(ctx, chunk):
    """qe_pipeline_stream: chunked upload / run / download equals upload + run + read back."""
    from queryengine_b200.engine import stream_pipelines
    rng = np.random.default_rng(chunk)
    n = 200_003
    v = rng.integers(0, 1000, n).astype(np.int64)
    d = rng.normal(size=n)
    dv = rng.random(n) > 0.2
    fields = [Field("i", I, False), Field("d", D, True)]
    ref_t = ColumnarTable.from_arrays(ctx, fields, [v, d], [None, dv])
    filt = fn(Function.CMP_LT, col(0, I), 400.0)
    projs = [col(0, I), fn(Function.MUL, col(1), 2.0)]
    whole = Pipeline(ref_t, filt, projs).run()
    dense = Pipeline(ref_t, None, [fn(Function.ADD, col(1), 1.0)]).run()
    want_n = whole.num_rows
    # streamed: the table starts empty, host buffers are plain numpy arrays (pageable is allowed, pinned is faster)
    t = ColumnarTable(ctx, fields, n)
    vbits = np.packbits(dv.astype(np.uint8), bitorder="little")
    sel = np.zeros(n, np.uint32)
    o0 = np.zeros(n, np.int64)
    o1 = np.zeros(n, np.float64)
    o1v = np.zeros((n + 7) // 8, np.uint8)
    o2 = np.zeros(n, np.float64)
    o2v = np.zeros((n + 7) // 8, np.uint8)
    pipes = [Pipeline(t, filt, projs), Pipeline(t, None, [fn(Function.ADD, col(1), 1.0)])]
    res = stream_pipelines(t, pipes, [v.ctypes.data, d.ctypes.data], [None, vbits.ctypes.data], chunk_rows=chunk,
                           host_selection=[sel.ctypes.data, None],
                           host_out_values=[[o0.ctypes.data, o1.ctypes.data], [o2.ctypes.data]],
                           host_out_validity=[[None, o1v.ctypes.data], [o2v.ctypes.data]])
    assert res[0].num_rows == want_n and res[1].num_rows == n
    assert np.array_equal(sel[:want_n], whole.selection())
    w0, _ = whole.column(0)
    w1, w1v = whole.column(1)
    assert np.array_equal(o0[:want_n], w0)
    got1v = np.unpackbits(o1v, bitorder="little")[:want_n].astype(bool)
    assert np.array_equal(got1v, w1v)
    assert np.array_equal(o1[:want_n][w1v], np.asarray(w1)[w1v])
    dd, ddv = dense.column(0)
    got2v = np.unpackbits(o2v, bitorder="little")[:n].astype(bool)
    assert np.array_equal(got2v, ddv) and np.array_equal(o2[ddv], np.asarray(dd)[ddv])
    # the device-side results are complete as well
    assert np.array_equal(res[0].selection(), whole.selection())
    for r in res + [whole, dense]:
        r.close()
    t.close()
    ref_t.close()


@pytest.mark.parametrize("interpreter", [False, True])
def test_mod_fast_path_is_bit_exact(ctx, interpreter):
    """MOD = JVM DREM = C fmod (M/evaluator/Interpreter.kt:100).  The device code takes an exact
    three-instruction path when the quotient fits 53 bits and the library loop otherwise: both sides of
    that switch, signed zeros, infinities, NaN, denormals and quotients next to integers."""
    rng = np.random.default_rng(123)
    n = 200_000
    mant = rng.random(n) + 1.0
    a = mant * 2.0 ** rng.integers(-40, 80, n) * rng.choice([-1.0, 1.0], n)
    b = (rng.random(n) + 1.0) * 2.0 ** rng.integers(-40, 40, n) * rng.choice([-1.0, 1.0], n)
    # exact multiples and near-multiples: quotient an integer, or one ulp off either side
    k = rng.integers(1, 1 << 30, n // 4).astype(np.float64)
    bb = np.round(rng.random(n // 4) * 1000 + 1, 3)
    a[: n // 4] = k * bb
    b[: n // 4] = bb
    a[n // 4: n // 4 + 1000] = np.nextafter(a[:1000], np.inf)
    b[n // 4: n // 4 + 1000] = b[:1000]
    special = np.array([0.0, -0.0, np.inf, -np.inf, np.nan, 5e-324, -5e-324, 2.2250738585072014e-308, 1.7976931348623157e308,
                        1.0, -1.0, 7.0, 2.0 ** 53, 2.0 ** 53 + 2, 2.0 ** 52 + 1, 3.0, 0.1, 1e-300, 1e300])
    sa, sb = np.meshgrid(special, special)
    m = sa.size
    a[-m:] = sa.ravel()
    b[-m:] = sb.ravel()
    # quotients around 2^53: the switch between the two paths
    a[-m - 2000:-m] = (2.0 ** 53 + rng.integers(-1000, 1000, 2000)) * 3.0
    b[-m - 2000:-m] = 3.0
    spec = TableSpec([Field("a", D, False), Field("b", D, False)], [a, b], [None, None])
    compare_pipeline(ctx, spec, None, [fn(Function.MOD, col(0), col(1)), fn(Function.MOD, col(0), 7.0),
                                       fn(Function.MOD, col(1), col(0))], interpreter=interpreter)


@pytest.mark.parametrize("k", [7.0, -7.0, 0.1, 3.0, 1e-3, 1e10, -2.5, 3.0 * 2.0 ** -20, 1.0, 2.0 ** -600, 2.0 ** 600, 0.7,
                               1e-320, 2.0 ** 1010, 0.0, float("inf"), float("nan")])
def test_mod_by_a_literal_is_bit_exact(ctx, k):
    """`MOD(x, literal)`: the generated code multiplies by the literal's reciprocal and corrects the quotient
    (qe_mod_k); dividends on and next to multiples of k, quotients up to and beyond the 2^51 switch to the general
    path, signed zeros, infinities, NaN, denormals; divisors the fast path must refuse (denormal, huge, 0, Inf, NaN)."""
    rng = np.random.default_rng(321)
    n = 300_000
    kk = abs(k) if np.isfinite(k) and k != 0.0 else 1.0
    a = (rng.random(n) + 1.0) * 2.0 ** rng.integers(-60, 80, n) * rng.choice([-1.0, 1.0], n)
    with np.errstate(over="ignore", invalid="ignore"):
        q = rng.integers(0, 1 << 40, n // 3).astype(np.float64)
        mult = q * kk                                           # (nearly) exact multiples, then one ulp either side
        a[: n // 3] = mult * rng.choice([-1.0, 1.0], n // 3)
        a[n // 3: n // 3 + 30000] = np.nextafter(mult[:30000], np.inf)
        a[n // 3 + 30000: n // 3 + 60000] = np.nextafter(mult[:30000], -np.inf)
        big = (2.0 ** 51 + rng.integers(-2000, 2000, 4000)) * kk  # quotients around the 2^51 switch
        a[-5000:-1000] = big * rng.choice([-1.0, 1.0], 4000)
    special = np.array([0.0, -0.0, np.inf, -np.inf, np.nan, 5e-324, -5e-324, 2.2250738585072014e-308, 1.7976931348623157e308,
                        kk, -kk, np.nextafter(kk, 0.0), np.nextafter(kk, np.inf), 2 * kk, 0.5 * kk, 2.0 ** 53, 2.0 ** 53 + 2, 1e300])
    a[-special.size:] = special
    a = np.where(np.isfinite(a) | np.isnan(a) | np.isinf(a), a, 0.0)
    spec = TableSpec([Field("a", D, False)], [a], [None])
    compare_pipeline(ctx, spec, None, [fn(Function.MOD, col(0), k), fn(Function.ADD, fn(Function.MOD, fn(Function.MUL, col(0), 3.0), k), 1.0)])


# ---------------------------------------------------------------- STRING columns as dictionary codes

S = DataType.STRING


@pytest.mark.parametrize("interpreter", [False, True])
def test_compilertest_if_with_string_literals_golden(ctx, interpreter):
    # T/evaluator/CompilerTest.kt:104-120, as written there: IF(cond, "t", "f") over true / false / null
    e = FunctionExpression(Function.IF, [ColumnExpression("cond", 0, B), StringLiteralExpression("t"),
                                         StringLiteralExpression("f")], S)
    assert _eval_rows(ctx, [Field("cond", B, True)], [[True], [False], [None]], e, interpreter) == ["t", "f", None]


def _string_table(rng, n):
    vocab = ["a", "b", "B", "", "lineitem", "R", "N", "\u00e4\u00f6", "a "]
    s0 = [None if rng.random() < 0.2 else vocab[int(rng.integers(len(vocab)))] for _ in range(n)]
    s1 = [vocab[int(rng.integers(4))] for _ in range(n)]
    x = np.round(rng.uniform(-10, 10, n), 1)
    b = rng.random(n) < 0.5
    bv = rng.random(n) < 0.9
    rows = [[s0[i], s1[i], float(x[i]), (bool(b[i]) if bv[i] else None)] for i in range(n)]
    fields = [Field("s0", S, True), Field("s1", S, False), Field("x", D, False), Field("b", B, True)]
    return fields, rows


@pytest.mark.parametrize("interpreter", [False, True])
@pytest.mark.parametrize("nrows", [0, 1, 777, 20000])
def test_string_equality_if_and_filter_match_the_oracle(ctx, nrows, interpreter):
    from oracle_binding import OraclePlan, OracleTable
    rng = np.random.default_rng(100 + nrows)
    fields, rows = _string_table(rng, nrows)
    s0, s1, x, b = col(0, S), col(1, S), col(2), col(3, B)
    lit_s = StringLiteralExpression
    projs = [
        fn(Function.CMP_EQ, s0, s1),
        fn(Function.CMP_NE, s0, lit_s("b")),
        fn(Function.IF, fn(Function.CMP_LT, x, 0.0), s0, lit_s("neg")),
        fn(Function.IF, fn(Function.CMP_EQ, s0, s1), s1, fn(Function.IF, b, lit_s("never seen in a column"), s0)),
        s0, s1,
        fn(Function.CMP_EQ, lit_s("a"), lit_s("a")),
    ]
    filters = [None,
               fn(Function.OR, fn(Function.CMP_EQ, s0, lit_s("a")), fn(Function.CMP_NE, s1, s0)),
               fn(Function.AND, fn(Function.CMP_NE, s0, lit_s("")), b)]
    otable = OracleTable.from_rows([f.type for f in fields], rows) if nrows else None
    t = _rows_table(ctx, fields, rows) if nrows else ColumnarTable(ctx, fields, 0)
    for f in filters:
        got = GpuPipelineOperator(t, f, projs, None, interpreter=interpreter).map()
        want = OraclePlan(otable, [0, 1, 2, 3], filter=f, projections=projs).rows() if nrows else []
        assert got == want, (f, next((g, w) for g, w in zip(got, want) if g != w) if len(got) == len(want) else (len(got), len(want)))
    # COUNT over a STRING expression skips nulls (GlobalAggregationOperator.kt:16-21, TypeCheck.kt:118-120)
    if nrows:
        got = GpuPipelineOperator(t, filters[2], [projs[2], s0], [A.COUNT, A.COUNT], interpreter=interpreter).map()
        want = OraclePlan(otable, [0, 1, 2, 3], filter=filters[2], projections=[projs[2], s0], aggregates=[A.COUNT, A.COUNT]).rows()
        assert got == want
    t.close()


def _row_values_equal(a, b):
    import math
    if isinstance(a, float) and isinstance(b, float):
        return (math.isnan(a) and math.isnan(b)) or (a == b and math.copysign(1.0, a) == math.copysign(1.0, b))
    return a == b and type(a) == type(b)


@pytest.mark.parametrize("interpreter", [False, True])
@pytest.mark.parametrize("seed", range(8))
def test_random_trees_with_string_columns(ctx, seed, interpreter):
    """Random well-typed trees over DOUBLE / BOOLEAN / STRING columns (string equality, IF over strings, strings
    inside filters): rows out of the GPU operator == rows out of the oracle, which evaluates real strings."""
    from exprgen import STRINGS
    from oracle_binding import OraclePlan, OracleTable
    rng = np.random.default_rng(4000 + seed)
    n = int(rng.integers(1, 4000))
    fields = [Field("d0", D, False), Field("d1", D, True), Field("b0", B, True), Field("s0", S, True), Field("s1", S, False)]
    vocab = STRINGS[:5] + ["x"]
    rows = []
    for _ in range(n):
        rows.append([float(np.round(rng.uniform(-5, 5), 1)) if rng.random() < 0.9 else float(rng.choice([np.nan, -0.0, np.inf])),
                     None if rng.random() < 0.2 else float(int(rng.integers(-3, 4))),
                     None if rng.random() < 0.2 else bool(rng.random() < 0.5),
                     None if rng.random() < 0.2 else vocab[int(rng.integers(len(vocab)))],
                     vocab[int(rng.integers(3))]])
    g = ExprGen(rng, fields)
    filt = g.boolean(4) if seed % 3 else None
    projs = [g.string(3), g.boolean(4), g.double(4), g.string(2)]
    t = _rows_table(ctx, fields, rows)
    got = GpuPipelineOperator(t, filt, projs, None, interpreter=interpreter).map()
    want = OraclePlan(OracleTable.from_rows([f.type for f in fields], rows), list(range(len(fields))), filter=filt, projections=projs).rows()
    assert len(got) == len(want)
    for i, (g_, w_) in enumerate(zip(got, want)):
        assert all(_row_values_equal(a, b) for a, b in zip(g_, w_)), (i, g_, w_)
    t.close()


@pytest.mark.parametrize("chunk", [4096, 65536, 1 << 20])
def test_stream_aggregation_folds_chunk_partials(ctx, chunk):
    """qe_pipeline_stream with a GlobalAggregation pipeline: every chunk is a partition, partials folded in chunk
    order -- COUNT / MIN / MAX identical to the one-shot run, SUM exact on integer-valued terms; a filter pipeline
    rides along in the same call."""
    from queryengine_b200.engine import stream_pipelines
    rng = np.random.default_rng(chunk + 1)
    n = 300_007
    q = rng.integers(0, 50, n).astype(np.float64)
    x = rng.integers(-1000, 1000, n).astype(np.float64)
    xv = rng.random(n) > 0.3
    fields = [Field("q", D, False), Field("x", D, True)]
    ref_t = ColumnarTable.from_arrays(ctx, fields, [q, x], [None, xv])
    filt = fn(Function.CMP_LT, col(0), 24.0)
    projs = [fn(Function.MUL, col(0), col(1)), col(1), col(1), col(1), col(0)]
    fns = [A.SUM, A.COUNT, A.MIN, A.MAX, A.AVG]
    whole = Pipeline(ref_t, filt, projs, fns).run()
    t = ColumnarTable(ctx, fields, n)
    xbits = np.packbits(xv.astype(np.uint8), bitorder="little")
    sel = np.zeros(n, np.uint32)
    pipes = [Pipeline(t, filt, projs, fns), Pipeline(t, filt, [col(0)])]
    res = stream_pipelines(t, pipes, [q.ctypes.data, x.ctypes.data], [None, xbits.ctypes.data], chunk_rows=chunk,
                           host_selection=[None, sel.ctypes.data], host_out_values=[None, [None]], host_out_validity=[None, [None]])
    for k in range(len(fns)):
        assert res[0].agg(k) == whole.agg(k), (fns[k], res[0].agg(k), whole.agg(k))
    keep = q < 24.0
    assert res[0].agg(1) == float(np.count_nonzero(keep & xv))
    assert res[0].agg(0) == float(np.sum((q * x)[keep & xv]))
    assert res[1].num_rows == int(keep.sum()) and np.array_equal(sel[:res[1].num_rows], np.nonzero(keep)[0])
    for r in res + [whole]:
        r.close()
    t.close()
    ref_t.close()
