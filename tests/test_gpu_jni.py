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


def test_selection_only_and_every_gpu_of_the_box_through_the_jni_natives(jni):
    """QE_RUN_NO_VALUES + resultSelection, and the qe_mgpu natives a JVM adapter drives several GPUs with (two contexts
    on one device when the box has one GPU): filter / projection concatenated in scan order, global row indices,
    aggregates equal to the single-context run bit for bit, group-by merged by key."""
    import torch
    rng = np.random.default_rng(22)
    n = 3 * (1 << 20) + 12_345                                        # partitions split at multiples of 2^20 rows
    spec = random_table(rng, n, "dDb")
    filt = fn(Function.OR, fn(Function.CMP_LT, col(0), -400.0), col(2, B))
    projs = [fn(Function.ADD, col(0), col(1)), col(2, B)]
    n_ref, ref_vals, ref_valid = OraclePlan(spec.with_row_index().oracle(), [0, 1, 2, 3], filter=filt, projections=projs + [col(3)],
                                            mode=Mode.BYTECODE_COMPILER).columns()
    ref_rows = ref_vals[2].astype(np.int64)
    # ---- one context: the selection vector alone
    ctx = jni.call("ctxCreate", 0)
    table, keep = _table(jni, ctx, spec)
    pf = _program(jni, ctx, filt)
    pp = [_program(jni, ctx, e) for e in projs]
    proj_arr, none = jni.long_array(pp), jni.int_array([])
    res = jni.call("pipelineRunFlags", ctx, table, pf, proj_arr, none, 8, 0)
    assert jni.call("resultNumRows", res) == n_ref and jni.call("resultNumColumns", res) == 0
    sel = np.zeros(n_ref, np.uint32)
    b = jni.direct_buffer(sel)
    jni.call("resultSelection", res, b)
    jni.free(b)
    assert np.array_equal(sel.astype(np.int64), ref_rows)
    fns = jni.int_array([int(A.SUM), int(A.COUNT)])
    two = jni.long_array([pp[0], pp[0]])
    agg1 = jni.call("pipelineRun", ctx, table, pf, two, fns, 0)
    flag = jni.boolean_array([0])
    want = [jni.call("resultAgg", agg1, k, flag) for k in range(2)]
    # ---- the group: the same programs run on every context
    devices = jni.int_array([0, 1] if torch.cuda.device_count() >= 2 else [0, 0])
    g = jni.call("mgpuCreate", devices)
    types = jni.int_array([int(f.type) for f in spec.fields])
    nullable = jni.boolean_array([1 if f.nullable else 0 for f in spec.fields])
    mt = jni.call("mtableCreate", g, types, nullable, n)
    for c, f in enumerate(spec.fields):
        data = np.packbits(spec.columns[c], bitorder="little") if f.type == B else np.ascontiguousarray(spec.columns[c], np.float64)
        vb = np.packbits(spec.valid[c], bitorder="little") if f.nullable else None
        bv, bn = jni.direct_buffer(data), (jni.direct_buffer(vb) if vb is not None else None)
        jni.call("mtableUpload", mt, c, bv, bn)
        jni.free(bv, bn)
    gctx = jni.call("mgpuCtx", g, 0)
    gf = _program(jni, gctx, filt)
    gp = [_program(jni, gctx, e) for e in projs]
    garr = jni.long_array(gp)
    mres = jni.call("mgpuPipelineRun", g, mt, gf, garr, none, 0)
    assert jni.call("mresultNumRows", mres) == n_ref and jni.call("mresultNumColumns", mres) == 2
    vals, vb = np.zeros(n_ref, np.float64), np.zeros((n_ref + 7) // 8 + 1, np.uint8)
    b1, b2 = jni.direct_buffer(vals), jni.direct_buffer(vb)
    jni.call("mresultColumn", mres, 0, b1, b2)
    jni.free(b1, b2)
    ok = np.unpackbits(vb, bitorder="little")[:n_ref].astype(bool)
    assert np.array_equal(ok, ref_valid[0]) and (f64_equal(vals, ref_vals[0]) | ~ok).all()
    rows = np.zeros(n_ref, np.uint64)
    b = jni.direct_buffer(rows)
    jni.call("mresultSelection", mres, b)
    jni.free(b)
    assert np.array_equal(rows.astype(np.int64), ref_rows)
    gtwo = jni.long_array([gp[0], gp[0]])
    magg = jni.call("mgpuPipelineRun", g, mt, gf, gtwo, fns, 0)
    got = [jni.call("mresultAgg", magg, k, flag) for k in range(2)]
    assert np.array(got).tobytes() == np.array(want).tobytes()        # SUM has the same bits for any number of partitions
    keys, inputs = jni.long_array([gp[1]]), jni.long_array([_program(jni, gctx, col(1)), _program(jni, gctx, col(1))])
    gfns = jni.int_array([int(A.MAX), int(A.COUNT)])
    mgrp = jni.call("mgpuGroupByRun", g, mt, 0, keys, inputs, gfns, 0)
    wantg = OraclePlan(spec.oracle(), [0, 1, 2], projections=[col(2, B), col(1), col(1)], aggregates=[A.MAX, A.COUNT], group_count=1).rows()
    assert jni.call("mresultNumRows", mgrp) == len(wantg)
    cnt = np.zeros(len(wantg), np.float64)
    vbits = np.zeros(2, np.uint8)
    b1, b2 = jni.direct_buffer(cnt), jni.direct_buffer(vbits)
    jni.call("mresultColumn", mgrp, 2, b1, b2)
    jni.free(b1, b2)
    assert [int(c) for c in cnt] == [w[2] for w in wantg]
    assert jni.outstanding == 0
    for r in (mres, magg, mgrp):
        jni.call("mresultDestroy", r)
    jni.call("mtableDestroy", mt)
    jni.call("mgpuDestroy", g)
    for r in (res, agg1):
        jni.call("resultDestroy", r)
    jni.call("tableDestroy", table)
    jni.call("ctxDestroy", ctx)
    jni.free(proj_arr, none, fns, two, flag, devices, types, nullable, garr, gtwo, keys, inputs, gfns)
