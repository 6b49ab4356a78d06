"""GPU parity tests of the round-2 runtime, all through the C ABI:

* literals as kernel parameters (one kernel per literal class; results against the oracle for every literal)
* QE_RUN_NO_VALUES (the selection vector as the whole output of a filter over a host-resident table)
* bounded-memory streaming: a table larger than its device window, chunk slots and output buffer sets reused round
  robin, compacted bit columns spliced across chunks, chunk-relative selection vectors with a 64-bit chunk base
* the fixed-shape aggregation: SUM / AVG bit-identical for any grid, batch split, chunk size and partition count
* several contexts driven by one process (qe_mgpu_*): on a one-GPU box two contexts share the device, so the
  multi-context path runs under the driver's single-GPU test pass too
* the ahead-of-time table and the no-compiler fallback

The checker is the oracle (tests/oracle_binding.py), never the thing under test.
"""
import ctypes as C
import math
import os
import subprocess
import sys

import numpy as np
import pytest

from oracle_binding import OraclePlan, OracleTable
from parity import TableSpec, compare_pipeline, f64_equal, random_table
from queryengine_b200 import capi, sharding
from queryengine_b200.ast import AggregationFunction as A, DataType, Function, Mode, col, fn, lit
from queryengine_b200.engine import ColumnarTable, Context, Field, GroupByPipeline, Pipeline, jit_stats, stream_pipelines
from queryengine_b200.multigpu import MultiGpu, ShardedGroupBy, ShardedPipeline, ShardedTable, partition_rows

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
D, B, I, T, S = DataType.DOUBLE, DataType.BOOLEAN, DataType.INT64, DataType.DATE32, DataType.STRING
SUPER = capi.QE_AGG_SUPER_ROWS


@pytest.fixture(scope="module")
def ctx():
    c = Context(0)
    yield c
    c.close()


class Pinned:
    def __init__(self, nbytes, dtype):
        self.ptr = C.c_void_p()
        capi.check(capi.lib().qe_host_alloc(max(nbytes, 64), C.byref(self.ptr)))
        self.nbytes = max(nbytes, 64)
        self.array = np.frombuffer((C.c_char * self.nbytes).from_address(self.ptr.value), dtype=dtype)

    def free(self):
        if self.ptr:
            self.array = None
            capi.lib().qe_host_free(self.ptr)
            self.ptr = C.c_void_p()


# ------------------------------------------------------------------ literals as parameters

def test_one_kernel_serves_every_literal_of_a_class(ctx):
    rng = np.random.default_rng(1)
    spec = random_table(rng, 50_000, "idDt")
    before = jit_stats()
    # integer column against integral literals, a double column against plain literals, MOD by a literal, date compare
    for k in (-3.0, 0.0, 17.0, 2.0 ** 40, -(2.0 ** 39)):
        compare_pipeline(ctx, spec, fn(Function.CMP_LT, col(0, I), k), [col(0, I)])
    for k in (0.5, -123.25, 999.0):
        compare_pipeline(ctx, spec, fn(Function.CMP_GE, col(1), k), [fn(Function.MOD, col(1), k + 7.5), col(3, T)])
    for k in (8500.0, 10999.0):
        compare_pipeline(ctx, spec, fn(Function.CMP_LE, col(3, T), k), [col(2)])
    after = jit_stats()
    # three query shapes (+ literal classes: 0.0 against an integer column is still integral) -> a handful of cubins
    assert (after["nvrtc"] + after["disk"] + after["aot"]) - (before["nvrtc"] + before["disk"] + before["aot"]) <= 4
    # classes that DO change the code: non-integral literal against an integer column, NaN / zero against a double
    for k in (0.5, float("nan"), -0.0, 1e300, -1e300):
        compare_pipeline(ctx, spec, fn(Function.CMP_EQ, col(0, I), k), [col(0, I)])
        compare_pipeline(ctx, spec, fn(Function.CMP_GT, col(1), k), [col(1)])
        compare_pipeline(ctx, spec, fn(Function.CMP_NE, k, col(0, I)), [col(0, I)])
    # more literals than parameter slots
    e = col(1)
    for i in range(30):
        e = fn(Function.ADD, e, float(i) * 0.125)
    compare_pipeline(ctx, spec, fn(Function.CMP_LT, e, 100.0), [e])


def test_int64_columns_up_to_the_2p53_contract(ctx):
    """INT64 columns stand for the reference's DOUBLE values: exact below 2^53 (include/qegpu.h).  At the edge of the
    contract the generated code (integer-domain compare) and the interpreter kernels (compare after widening) still
    agree with the oracle; a generator range beyond it is refused."""
    top = (1 << 53) - 1
    vals = np.array([top, -top, top - 1, -(top - 1), 0, 1, -1, 1 << 52, -(1 << 52), (1 << 52) + 1], dtype=np.int64)
    vals = np.tile(vals, 1000)
    spec = TableSpec([Field("x", I, True)], [vals], [np.tile(np.array([1, 1, 1, 0, 1, 1, 1, 1, 1, 1], bool), 1000)])
    for k in (float(top), float(top - 1), -float(top), float(1 << 52), 0.5, float(1 << 52) + 1.0):
        for op in (Function.CMP_LT, Function.CMP_LE, Function.CMP_EQ, Function.CMP_GE, Function.CMP_NE):
            for interp in (False, True):
                compare_pipeline(ctx, spec, fn(op, col(0, I), k), [col(0, I), fn(Function.ADD, col(0, I), 1.0)], interpreter=interp)
    t = ColumnarTable(ctx, [Field("x", I, False)], 1024)
    with pytest.raises(capi.QeError) as ei:
        t.generate(0, capi.QE_GEN_UNIFORM_INT, 1, 0.0, float(1 << 53))
    assert "2^53" in ei.value.message
    t.close()


def test_loaded_kernels_are_bounded_by_an_lru(ctx):
    c2 = Context(0)
    c2.set_option("jit_cache_entries", 2)
    rng = np.random.default_rng(2)
    spec = random_table(rng, 20_000, "dD")
    shapes = [fn(Function.CMP_LT, col(0), 1.0), fn(Function.CMP_GT, fn(Function.ADD, col(0), col(1)), 1.0),
              fn(Function.CMP_LE, fn(Function.MUL, col(0), col(1)), 1.0)]
    for _ in range(3):                                   # every kernel is evicted and reloaded a few times
        for f in shapes:
            compare_pipeline(c2, spec, f, [col(1)])
    c2.close()


# ------------------------------------------------------------------ selection-only output

def test_no_values_is_the_same_selection_vector(ctx):
    rng = np.random.default_rng(3)
    spec = random_table(rng, 300_000, "dDb")
    t = spec.gpu(ctx)
    f = fn(Function.OR, fn(Function.CMP_LT, col(0), 0.0), col(2, B))
    full = Pipeline(t, f, [col(0), col(1)]).run()
    only = Pipeline(t, f, [col(0), col(1)], values=False).run()
    assert only.num_rows == full.num_rows and only.num_columns == 0 and full.num_columns == 2
    assert np.array_equal(only.selection(), full.selection())
    full.close(); only.close(); t.close()


# ------------------------------------------------------------------ bounded streaming

def _oracle_filter(spec, filt, projs):
    full = spec.with_row_index()
    row = col(len(full.fields) - 1)
    plan = OraclePlan(full.oracle(), list(range(len(full.fields))), filter=filt, projections=list(projs) + [row], mode=Mode.BYTECODE_COMPILER)
    return plan.columns()


@pytest.mark.parametrize("chunk_relative", [False, True])
def test_bounded_stream_equals_the_oracle(ctx, chunk_relative):
    """Table of 11.4 chunks streamed through a window of 3 chunk slots; outputs: selection vector, a nullable f64
    column, a BOOLEAN column (compacted bit columns: bytes shared by two chunks) and an int64 column."""
    rng = np.random.default_rng(4)
    chunk = 1 << 14
    n = 11 * chunk + 5000
    spec = random_table(rng, n, "dDbI")
    filt = fn(Function.OR, fn(Function.CMP_LT, col(0), -500.0), fn(Function.AND, col(2, B), fn(Function.CMP_GT, col(1), 900.0)))
    projs = [fn(Function.ADD, col(0), col(1)), col(2, B), col(3, I)]
    n_ref, ref_vals, ref_valid = _oracle_filter(spec, filt, projs)

    window = ColumnarTable(ctx, spec.fields, 3 * chunk)
    cap = n_ref + 64
    h_cols, h_valid = [], []
    for c, f in enumerate(spec.fields):
        if f.type == B:
            buf = Pinned((n + 7) // 8, np.uint8); buf.array[:(n + 7) // 8] = np.packbits(spec.columns[c], bitorder="little")
        else:
            buf = Pinned(n * 8, np.float64 if f.type == D else np.int64); buf.array[:n] = spec.columns[c]
        h_cols.append(buf)
        if f.nullable:
            vb = Pinned((n + 7) // 8, np.uint8); vb.array[:(n + 7) // 8] = np.packbits(spec.valid[c], bitorder="little")
        else:
            vb = None
        h_valid.append(vb)
    h_sel = Pinned(cap * 4, np.uint32)
    h_out = [Pinned(cap * 8, np.float64), Pinned(cap // 8 + 8, np.uint8), Pinned(cap * 8, np.int64)]
    h_on = [Pinned(cap // 8 + 8, np.uint8) for _ in range(3)]
    for b in h_out + h_on:
        b.array[:] = 0xAB if b.array.dtype == np.uint8 else 7          # garbage: every byte must be written or OR-ed correctly
    for b in (h_out[1], *h_on):
        b.array[:] = 0                                                  # bit columns are OR-ed into at chunk seams
    nchunks = -(-n // chunk)
    kept = np.zeros(nchunks, np.int64)
    pipe = Pipeline(window, filt, projs, out_capacity_rows=cap)
    res = None
    for rep in range(2):                                               # the second call reuses every buffer set
        res = stream_pipelines(window, [pipe], [b.ptr.value for b in h_cols], [b.ptr.value if b else None for b in h_valid],
                               chunk_rows=chunk, results=[res[0]] if res else None,
                               host_selection=[h_sel.ptr.value], host_out_values=[[b.ptr.value for b in h_out]],
                               host_out_validity=[[b.ptr.value for b in h_on]], total_rows=n, chunk_relative=chunk_relative,
                               chunk_kept=[kept])
        r = res[0]
        assert r.num_rows == n_ref
        sel = h_sel.array[:n_ref].astype(np.int64)
        if chunk_relative:
            assert kept[-1] == n_ref and (np.diff(kept) >= 0).all()
            chunk_of = np.searchsorted(kept, np.arange(n_ref), side="right")
            assert (sel < chunk).all()
            sel = sel + chunk_of * chunk
        assert np.array_equal(sel.astype(np.float64), ref_vals[-1]), "selection vector differs"
        v0 = np.unpackbits(h_on[0].array, bitorder="little")[:n_ref].astype(bool)
        assert np.array_equal(v0, ref_valid[0])
        assert (f64_equal(h_out[0].array[:n_ref], ref_vals[0]) | ~v0).all()
        b1 = np.unpackbits(h_out[1].array, bitorder="little")[:n_ref].astype(bool)
        v1 = np.unpackbits(h_on[1].array, bitorder="little")[:n_ref].astype(bool)
        assert np.array_equal(v1, ref_valid[1]) and np.array_equal(b1 & v1, (ref_vals[1] != 0) & v1)
        v2 = np.unpackbits(h_on[2].array, bitorder="little")[:n_ref].astype(bool)
        assert np.array_equal(v2, ref_valid[2])
        assert np.array_equal(h_out[2].array[:n_ref][v2].astype(np.float64), ref_vals[2][v2])
        with pytest.raises(capi.QeError):
            r.column(0)                                                # the rows are in the host buffers only
    # capacity of the host buffers is enforced
    small = Pipeline(window, filt, projs, out_capacity_rows=max(1, n_ref // 2))
    with pytest.raises(capi.QeError) as ei:
        stream_pipelines(window, [small], [b.ptr.value for b in h_cols], [b.ptr.value if b else None for b in h_valid],
                         chunk_rows=chunk, host_selection=[h_sel.ptr.value], total_rows=n)
    assert ei.value.status == capi.QE_ERR_CAPACITY
    res[0].close()
    window.close()
    for b in h_cols + [x for x in h_valid if x] + [h_sel] + h_out + h_on:
        b.free()


@pytest.mark.parametrize("frac", [0.004, 0.05, 0.3])
def test_bounded_stream_of_a_narrow_nullable_column_takes_the_sparse_path(ctx, frac):
    """One nullable f64 column streamed through a window of three chunk slots, `x < k` returning x: the chained chunks run
    on the row lists + sweeper warps (8-byte rows), the column's validity is a run of ones per tile whose ends meet inside
    bytes shared by two chunks, and the second stream picks its shape from what the first one kept (the whole stream, not
    the chunk).  Against numpy, twice."""
    rng = np.random.default_rng(int(frac * 1000) + 5)
    chunk = 1 << 16
    n = 9 * chunk + 12_345
    x = rng.uniform(0.0, 1.0, n)
    xv = rng.random(n) < 0.9
    keep = xv & (x < frac)
    want = np.nonzero(keep)[0]
    window = ColumnarTable(ctx, [Field("x", D, True)], 3 * chunk)
    h_x = Pinned(n * 8, np.float64); h_x.array[:n] = x
    h_v = Pinned((n + 7) // 8, np.uint8); h_v.array[:(n + 7) // 8] = np.packbits(xv, bitorder="little")
    cap = want.size + 64
    h_sel = Pinned(cap * 4, np.uint32)
    h_out = Pinned(cap * 8, np.float64)
    h_on = Pinned(cap // 8 + 8, np.uint8)
    pipe = Pipeline(window, fn(Function.CMP_LT, col(0), float(frac)), [col(0)], out_capacity_rows=cap)
    res = None
    for rep in range(3):
        h_out.array[:] = -1.0
        h_on.array[:] = 0
        res = stream_pipelines(window, [pipe], [h_x.ptr.value], [h_v.ptr.value], chunk_rows=chunk, results=[res[0]] if res else None,
                               host_selection=[h_sel.ptr.value], host_out_values=[[h_out.ptr.value]],
                               host_out_validity=[[h_on.ptr.value]], total_rows=n)
        assert res[0].num_rows == want.size, rep
        assert np.array_equal(h_sel.array[:want.size].astype(np.int64), want), rep
        assert np.array_equal(h_out.array[:want.size], x[want]), rep
        bits = np.unpackbits(h_on.array, bitorder="little")
        assert bits[:want.size].all() and not bits[want.size:].any(), rep          # ones as far as the output goes, nothing beyond
    res[0].close()
    window.close()
    for b in (h_x, h_v, h_sel, h_out, h_on):
        b.free()


def test_bounded_stream_dense_projection_and_aggregation(ctx):
    """Dense projection (values + validity per chunk) and a global aggregation through the bounded window; the
    aggregation over streamed chunks of 2^20 rows has the bits of one run over the resident table."""
    rng = np.random.default_rng(5)
    chunk = SUPER
    n = 5 * chunk + 12_345
    a = rng.uniform(-1000, 1000, n)
    b = rng.uniform(-10, 10, n)
    vb = rng.random(n) > 0.2
    fields = [Field("a", D, False), Field("b", D, True)]
    spec = TableSpec(fields, [a, b], [None, vb])
    expr = fn(Function.MUL, col(0), fn(Function.SUB, 1.0, col(1)))
    filt = fn(Function.CMP_GT, col(0), -900.0)
    resident = spec.gpu(ctx)
    want = Pipeline(resident, filt, [expr, expr, col(1)], [A.SUM, A.AVG, A.MIN]).run()
    want_vals = [want.agg(k) for k in range(3)]
    dense = Pipeline(resident, None, [expr]).run()
    dv, dn = dense.column(0)

    window = ColumnarTable(ctx, fields, 2 * chunk)
    ha, hb = Pinned(n * 8, np.float64), Pinned(n * 8, np.float64)
    hvb = Pinned((n + 7) // 8, np.uint8)
    ha.array[:n] = a; hb.array[:n] = b; hvb.array[:(n + 7) // 8] = np.packbits(vb, bitorder="little")
    ho, hon = Pinned(n * 8, np.float64), Pinned((n + 7) // 8 + 8, np.uint8)
    p_agg = Pipeline(window, filt, [expr, expr, col(1)], [A.SUM, A.AVG, A.MIN])
    p_dense = Pipeline(window, None, [expr])
    res = stream_pipelines(window, [p_agg, p_dense], [ha.ptr.value, hb.ptr.value], [None, hvb.ptr.value], chunk_rows=chunk,
                           host_out_values=[None, [ho.ptr.value]], host_out_validity=[None, [hon.ptr.value]], total_rows=n)
    got_vals = [res[0].agg(k) for k in range(3)]
    assert got_vals == want_vals, (got_vals, want_vals)              # bit for bit, SUM and AVG included
    assert res[1].num_rows == n
    gn = np.unpackbits(hon.array, bitorder="little")[:n].astype(bool)
    assert np.array_equal(gn, dn) and (f64_equal(ho.array[:n], dv) | ~dn).all()
    # resident streaming in chunks gives the same bits as well
    res2 = stream_pipelines(resident, [Pipeline(resident, filt, [expr, expr, col(1)], [A.SUM, A.AVG, A.MIN])], [None, None], chunk_rows=2 * chunk)
    assert [res2[0].agg(k) for k in range(3)] == want_vals
    for r in (want, dense, res[0], res[1], res2[0]):
        r.close()
    window.close(); resident.close()
    for x in (ha, hb, hvb, ho, hon):
        x.free()


# ------------------------------------------------------------------ fixed-shape aggregation

def test_sum_bits_do_not_depend_on_grid_batches_or_partitions(ctx):
    rng = np.random.default_rng(6)
    n = 7 * SUPER + 333_333
    a = rng.uniform(-1e6, 1e6, n)                                     # mixed signs: the order of additions matters
    b = rng.uniform(0, 1, n)
    fields = [Field("a", D, False), Field("b", D, False)]
    t = ColumnarTable.from_arrays(ctx, fields, [a, b])
    filt = fn(Function.CMP_LT, col(1), 0.7)
    fns = [A.SUM, A.AVG, A.MIN, A.MAX, A.COUNT]
    pipe = Pipeline(t, filt, [col(0)] * 5, fns)
    whole = pipe.run()
    want = [whole.agg(k) for k in range(5)]
    lists = np.stack([whole.agg_partials(k) for k in range(5)], axis=1)
    assert lists.shape[0] == -(-n // SUPER)
    assert sharding.combine_lists(fns, lists) == want
    # against the oracle: COUNT / MIN / MAX exact, SUM within the stated tolerance
    ot = OracleTable([D, D], n); ot.set_column(0, a); ot.set_column(1, b)
    ref, = OraclePlan(ot, [0, 1], filter=filt, projections=[col(0)] * 5, aggregates=fns, mode=Mode.BYTECODE_COMPILER).rows()
    assert want[4] == ref[4] and want[2] == ref[2] and want[3] == ref[3]
    terms = a[b < 0.7]
    exact = math.fsum(terms.tolist())
    assert abs(want[0] - exact) <= max(abs(ref[0] - exact), 4 * math.ulp(exact))
    # another grid
    c2 = Context(0)
    c2.set_option("grid_per_sm", 1)
    t2 = ColumnarTable.from_arrays(c2, fields, [a, b])
    r2 = Pipeline(t2, filt, [col(0)] * 5, fns).run()
    assert [r2.agg(k) for k in range(5)] == want
    r2.close(); t2.close(); c2.close()
    # row-range batches at super-tile boundaries, lists concatenated: 2, 3 and 8 "partitions"
    for parts in (2, 3, 8):
        pieces = []
        for p in range(parts):
            lo, hi = partition_rows(n, parts, p)
            if hi > lo:
                r = pipe.run(None, lo, hi)
                pieces.append(np.stack([r.agg_partials(k) for k in range(5)], axis=1))
                r.close()
        assert sharding.combine_lists(fns, np.concatenate(pieces, axis=0)) == want, parts
    # the interpreter kernels fold in the same shape (another tile size: the sub-tile partial differs in its last bits
    # at most, COUNT / MIN / MAX not at all)
    ri = Pipeline(t, filt, [col(0)] * 5, fns, interpreter=True).run()
    gi = [ri.agg(k) for k in range(5)]
    assert gi[2:] == want[2:] and abs(gi[0] - exact) <= max(abs(ref[0] - exact), 4 * math.ulp(exact))
    ri.close(); whole.close(); t.close()


# ------------------------------------------------------------------ several contexts, one process

def _devices():
    import torch
    return [0, 1] if torch.cuda.device_count() >= 2 else [0, 0]        # one GPU: two contexts share it


def test_mgpu_filter_projection_concatenates_in_row_order(ctx):
    devs = _devices()
    mg = MultiGpu(devs)
    rng = np.random.default_rng(7)
    n = 3 * SUPER + 77_777                                             # partitions of 2 Mi and 1 Mi + tail rows
    spec = random_table(rng, n, "dDbi")
    st = ShardedTable(mg, spec.fields, n)
    assert st.ranges == [(0, 2 * SUPER), (2 * SUPER, n)]
    for c in range(4):
        st.upload(c, spec.columns[c], spec.valid[c])
    filt = fn(Function.AND, fn(Function.CMP_LT, col(0), -800.0), fn(Function.CMP_GE, col(3, I), 0.0))
    projs = [fn(Function.SUB, col(0), col(1)), col(2, B), col(3, I)]
    n_ref, ref_vals, ref_valid = _oracle_filter(spec, filt, projs)
    sp = ShardedPipeline(st, filt, projs)
    res = sp.run()
    assert res.num_rows == n_ref and res.part_rows(0) + res.part_rows(1) == n_ref
    assert np.array_equal(res.selection().astype(np.float64), ref_vals[-1])      # global, 64-bit row indices
    v, ok = res.column(0, D)
    assert np.array_equal(ok, ref_valid[0]) and (f64_equal(v, ref_vals[0]) | ~ok).all()
    bv, bok = res.column(1, B)
    assert np.array_equal(bok, ref_valid[1]) and np.array_equal(bv & bok, (ref_vals[1] != 0) & bok)
    iv, iok = res.column(2, I)
    assert iok.all() and np.array_equal(iv.astype(np.float64), ref_vals[2])
    # the single-context result is the same thing
    t1 = spec.gpu(ctx)
    r1 = Pipeline(t1, filt, projs).run()
    assert np.array_equal(r1.selection().astype(np.uint64), res.selection())
    r1.close(); t1.close()
    # errors of one partition surface on the calling thread with their message
    with pytest.raises(capi.QeError) as ei:
        ShardedPipeline(st, fn(Function.ADD, col(0), 1.0), [col(0)]).run()
    assert "BOOLEAN" in ei.value.message and "partition 0" in ei.value.message
    res.close(); st.close(); mg.close()


def test_mgpu_aggregates_and_groupby_equal_one_context(ctx):
    devs = _devices()
    mg = MultiGpu(devs)
    rng = np.random.default_rng(8)
    n = 2 * SUPER + 500_000
    key = rng.integers(0, 5, n).astype(np.float64)
    key[rng.random(n) < 0.01] = np.nan                                # NaN is one key (Double.equals)
    flag = np.array(["R", "A", "N"], dtype=object)[rng.integers(0, 3, n)]
    x = np.round(rng.uniform(-100, 100, n), 2)
    vx = rng.random(n) > 0.1
    fields = [Field("k", D, False), Field("f", S, True), Field("x", D, True)]
    vf = rng.random(n) > 0.05
    st = ShardedTable(mg, fields, n)
    st.upload(0, key)
    st.upload(1, [s if ok else None for s, ok in zip(flag, vf)])
    st.upload(2, x, vx)
    t1 = ColumnarTable(ctx, fields, n)
    t1.upload(0, key)
    t1.upload(1, [s if ok else None for s, ok in zip(flag, vf)])
    t1.upload(2, x, vx)
    # global aggregation: the same bits as one context, for SUM and AVG too
    filt = fn(Function.CMP_GT, col(2), -50.0)
    fns = [A.SUM, A.AVG, A.MIN, A.MAX, A.COUNT]
    one = Pipeline(t1, filt, [col(2)] * 5, fns).run()
    many = ShardedPipeline(st, filt, [col(2)] * 5, fns).run()
    assert [many.agg(k) for k in range(5)] == [one.agg(k) for k in range(5)]
    assert many.num_rows == 1
    # group-by: merged by key on the host, groups in order of first appearance
    g1 = GroupByPipeline(t1, None, [col(0), col(1, S)], [col(2), col(2), col(2), col(1, S)], [A.SUM, A.MIN, A.MAX, A.COUNT]).run()
    want_rows = []
    cols = [g1.column(k) for k in range(6)]
    for i in range(g1.num_rows):
        row = []
        for k, (v, ok) in enumerate(cols):
            row.append(None if not ok[i] else (ctx.dictionary.decode(v[i]) if k == 1 else float(v[i])))
        want_rows.append(row)
    sg = ShardedGroupBy(st, None, [col(0), col(1, S)], [col(2), col(2), col(2), col(1, S)], [A.SUM, A.MIN, A.MAX, A.COUNT])
    got_rows = sg.rows(sg.run(), [D, S])
    assert len(got_rows) == len(want_rows) == 24                      # (5 values + NaN) x (3 flags + null)
    for g, w in zip(got_rows, want_rows):
        for k, (a_, b_) in enumerate(zip(g, w)):
            if isinstance(a_, float) and isinstance(b_, float):
                if k == 2:                                            # SUM: two partial sums folded, not one sequence
                    assert abs(a_ - b_) <= 1e-6 * max(1.0, abs(b_)), (g, w)
                else:
                    assert (math.isnan(a_) and math.isnan(b_)) or a_ == b_, (g, w)
            else:
                assert a_ == b_, (g, w)
    one.close(); many.close(); g1.close(); t1.close(); st.close(); mg.close()


def test_mgpu_streams_every_partition_at_once(ctx):
    mg = MultiGpu(_devices())
    n = 2 * SUPER + 100_000
    rng = np.random.default_rng(9)
    a = rng.integers(0, 1 << 40, n, dtype=np.int64)
    st = ShardedTable(mg, [Field("col", I, False)], n)
    h_in = Pinned(n * 8, np.int64)
    h_in.array[:n] = a
    k = float(1 << 39)
    want = np.nonzero(a < k)[0]
    caps = [int(((e - b) * 0.55)) + 4096 for b, e in st.ranges]
    offs = [0, caps[0]]
    h_sel, h_val = Pinned(sum(caps) * 4, np.uint32), Pinned(sum(caps) * 8, np.int64)
    sp = ShardedPipeline(st, fn(Function.CMP_LT, col(0, I), k), [col(0, I)], out_capacity_rows=max(caps))
    res = sp.stream([h_in.ptr.value], [8], 1 << 18, h_sel.ptr.value, [h_val.ptr.value], offs)
    rows = [r.num_rows for r in res]
    assert sum(rows) == want.size
    sel = np.concatenate([h_sel.array[offs[p]:offs[p] + rows[p]].astype(np.int64) + st.ranges[p][0] for p in range(2)])
    val = np.concatenate([h_val.array[offs[p]:offs[p] + rows[p]] for p in range(2)])
    assert np.array_equal(sel, want) and np.array_equal(val, a[want])
    for r in res:
        r.close()
    st.close(); mg.close()
    h_in.free(); h_sel.free(); h_val.free()


# ------------------------------------------------------------------ ahead-of-time table, no compiler

def test_baseline_pipelines_run_without_a_compiler():
    """QE_DISABLE_NVRTC=1, no disk cache: configs[1] runs generated code out of the ahead-of-time table; a pipeline
    that is not in it falls back to the interpreter kernels (still on the GPU), with the same results."""
    code = r"""
import sys
sys.path.insert(0, %r); sys.path.insert(0, %r)
import numpy as np
from queryengine_b200 import capi, workloads as W
from queryengine_b200.ast import DataType, Function, col, fn
from queryengine_b200.engine import ColumnarTable, Context, Field, Pipeline, jit_stats
ctx = Context(0)
I = DataType.INT64
n = 1 << 21
t = ColumnarTable(ctx, [Field("col", I, False)], n)
t.generate(0, capi.QE_GEN_UNIFORM_U40, 42)
import bench
vals = bench.synthetic_column(n)
r = Pipeline(t, W.single_column_filter(0.37), [col(0, I, "col")]).run()
assert r.used_jit and np.array_equal(r.selection(), np.nonzero(vals < float(int(0.37 * 2 ** 40)))[0])
s = jit_stats()
assert s["aot"] >= 1 and s["nvrtc"] == 0 and s["disk"] == 0, s
odd = Pipeline(t, fn(Function.CMP_GT, fn(Function.MOD, col(0, I), 1000.0), 500.0), [col(0, I, "col")]).run()
assert not odd.used_jit and np.array_equal(odd.selection(), np.nonzero(np.fmod(vals.astype(np.float64), 1000.0) > 500.0)[0])
print("OK", s)
""" % (ROOT, os.path.join(ROOT, "tests"))
    env = dict(os.environ, QE_DISABLE_NVRTC="1", QE_CACHE_DIR="off")
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True)
    assert out.returncode == 0 and "OK" in out.stdout, out.stdout + out.stderr
