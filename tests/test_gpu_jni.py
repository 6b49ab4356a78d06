"""The JNI natives of bindings/jni/qegpu_jni.c EXECUTED on the GPU through a mock JNIEnv (tests/jni_mock.c): context ->
programs -> table -> upload from direct buffers -> pipelineRun / groupByRun -> result columns, compared with the
oracle; and the error path (status != 0 -> one pending IllegalStateException with qe_last_error()).  This is the
drop-in SURVEY.md 8f next-1 names, minus the JVM: the Kotlin adapter (bindings/kotlin/) calls exactly these natives.
A real JVM run stays "unavailable on this box" (no JDK in the image)."""
import numpy as np
import pytest

from jni_driver import JavaException, QeGpu
from oracle_binding import OraclePlan, OracleTable
from parity import f64_equal, random_table
from queryengine_b200.ast import AggregationFunction as A, DataType, Function, Mode, col, fn
from queryengine_b200.program import serialize

pytestmark = pytest.mark.gpu
D, B = DataType.DOUBLE, DataType.BOOLEAN


@pytest.fixture(scope="module")
def jni():
    return QeGpu()


def _program(jni, ctx, expr):
    words = jni.int_array(serialize(expr).view(np.int32))
    h = jni.call("programCreate", ctx, words)
    jni.free(words)
    return h


def _table(jni, ctx, spec):
    types = jni.int_array([int(f.type) for f in spec.fields])
    nullable = jni.boolean_array([1 if f.nullable else 0 for f in spec.fields])
    t = jni.call("tableCreate", ctx, types, nullable, spec.nrows)
    jni.free(types, nullable)
    keep = []
    for c, f in enumerate(spec.fields):
        data = np.packbits(spec.columns[c], bitorder="little") if f.type == B else np.ascontiguousarray(spec.columns[c], np.float64)
        vb = np.packbits(spec.valid[c], bitorder="little") if f.nullable else None
        bv, bn = jni.direct_buffer(data), (jni.direct_buffer(vb) if vb is not None else None)
        jni.call("tableUpload", t, c, bv, bn, 0, spec.nrows)
        keep += [data, vb]
        jni.free(bv, bn)
    return t, keep


def _column(jni, result, k, n):
    tv = jni.call("resultColumnType", result, k)
    typ = tv & 0xff
    vals = np.zeros(max(n, 1), np.float64) if typ == int(D) else np.zeros((n + 7) // 8 + 1, np.uint8)
    vb = np.zeros((n + 7) // 8 + 1, np.uint8)
    b1, b2 = jni.direct_buffer(vals), jni.direct_buffer(vb)
    jni.call("resultColumn", result, k, b1, b2)
    jni.free(b1, b2)
    valid = np.unpackbits(vb, bitorder="little")[:n].astype(bool)
    if typ == int(B):
        return np.unpackbits(vals, bitorder="little")[:n].astype(bool), valid
    return vals[:n], valid


def test_filter_projection_aggregation_and_groupby_through_the_jni_natives(jni):
    rng = np.random.default_rng(21)
    spec = random_table(rng, 200_000, "dDbd")
    ctx = jni.call("ctxCreate", 0)
    table, keep = _table(jni, ctx, spec)
    filt = fn(Function.AND, fn(Function.CMP_LT, col(0), 250.0), col(2, B))
    projs = [fn(Function.MUL, col(0), col(1)), col(2, B), col(3)]
    pf = _program(jni, ctx, filt)
    pp = [_program(jni, ctx, e) for e in projs]
    # Scan -> Filter -> Projection
    proj_arr, none = jni.long_array(pp), jni.int_array([])
    res = jni.call("pipelineRun", ctx, table, pf, proj_arr, none, 0)
    n = jni.call("resultNumRows", res)
    plan = OraclePlan(spec.oracle(), [0, 1, 2, 3], filter=filt, projections=projs, mode=Mode.BYTECODE_COMPILER)
    n_ref, ref_vals, ref_valid = plan.columns()
    assert n == n_ref and jni.call("resultNumColumns", res) == 3
    for k in range(3):
        v, ok = _column(jni, res, k, n)
        assert np.array_equal(ok, ref_valid[k])
        assert (f64_equal(v.astype(np.float64), ref_vals[k]) | ~ok).all()
    # the result handle is reused by the next run (previous != 0), as GpuPipelineOperator.open does
    res = jni.call("pipelineRun", ctx, table, pf, proj_arr, none, res)
    assert jni.call("resultNumRows", res) == n
    # Scan -> Filter -> Projection -> GlobalAggregation
    fns = jni.int_array([int(A.COUNT), int(A.MIN), int(A.MAX)])
    three = jni.long_array([pp[2]] * 3)
    agg = jni.call("pipelineRun", ctx, table, pf, three, fns, 0)
    ref_row, = OraclePlan(spec.oracle(), [0, 1, 2, 3], filter=filt, projections=[projs[2]] * 3, aggregates=[A.COUNT, A.MIN, A.MAX],
                          mode=Mode.BYTECODE_COMPILER).rows()
    flag = jni.boolean_array([0])
    for k in range(3):
        v = jni.call("resultAgg", agg, k, flag)
        assert not jni.L.mock_boolean_array_get(flag, 0)
        assert f64_equal(np.array([v]), np.array([float(ref_row[k])])).all()
    # group-by: boolean key, SUM / COUNT of a nullable column
    keys, inputs = jni.long_array([pp[1]]), jni.long_array([_program(jni, ctx, col(1)), _program(jni, ctx, col(1))])
    gfns = jni.int_array([int(A.MAX), int(A.COUNT)])
    grp = jni.call("groupByRun", ctx, table, 0, keys, inputs, gfns, 0, 0)
    ng = jni.call("resultNumRows", grp)
    want = OraclePlan(spec.oracle(), [0, 1, 2, 3], projections=[col(2, B), col(1), col(1)], aggregates=[A.MAX, A.COUNT], group_count=1).rows()
    assert ng == len(want) == 2
    kcol = _column(jni, grp, 0, ng)
    mx, cnt = _column(jni, grp, 1, ng), _column(jni, grp, 2, ng)
    for g in range(ng):
        assert bool(kcol[0][g]) == want[g][0] and cnt[0][g] == want[g][2]
        assert f64_equal(np.array([mx[0][g]]), np.array([float(want[g][1])])).all()
    # ---- the error path: a type error inside the library surfaces as ONE IllegalStateException with the message of
    # the reference's TypeCheck (M/evaluator/TypeCheck.kt:123-125), and the native has released its arrays
    bad = _program(jni, ctx, fn(Function.ADD, col(0), col(2, B)))
    one = jni.long_array([bad])
    with pytest.raises(JavaException) as ei:
        jni.call("pipelineRun", ctx, table, 0, one, none, 0)
    assert ei.value.cls == "java/lang/IllegalStateException"
    assert "Invalid operand types for [ADD] [DOUBLE, BOOLEAN]" in ei.value.message
    with pytest.raises(JavaException) as ei:
        jni.call("programCreate", ctx, jni.int_array([999]))
    assert "unknown opcode" in ei.value.message
    with pytest.raises(JavaException):
        jni.call("resultColumn", res, 17, None, None)
    assert jni.outstanding == 0                                       # every Get...ArrayElements was released
    for r in (res, agg, grp):
        jni.call("resultDestroy", r)
    for p in [pf, bad] + pp:
        jni.call("programDestroy", p)
    jni.call("tableDestroy", table)
    jni.call("ctxDestroy", ctx)
    jni.free(proj_arr, none, fns, three, flag, keys, inputs, gfns, one)
