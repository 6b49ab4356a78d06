#!/usr/bin/env python
"""A reduced pass over every kernel family at small sizes, for compute-sanitizer (tools/sanitize.sh):

  compute-sanitizer --tool memcheck|racecheck|synccheck python tools/sanitizer_subset.py

Ring filter shapes (consumer warps, stages, defer, bit columns, capacity overflow), append and both streaming modes,
dense projection (grid-stride and ring), global aggregation, every group-by variant, the interpreter kernels.  Results
are checked against numpy, so a kernel the sanitizer perturbs still has to be right.  SURVEY.md section 5 asked for
this on the shared-memory protocol (five mbarrier arrays, relaxed gpu-scope descriptors, CAS spin loops)."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from queryengine_b200 import capi  # noqa: E402
from queryengine_b200.ast import AggregationFunction as A, DataType, Function, col, fn  # noqa: E402
from queryengine_b200.engine import ColumnarTable, Context, Field, GroupByPipeline, Pipeline, stream_pipelines  # noqa: E402

D, B, I = DataType.DOUBLE, DataType.BOOLEAN, DataType.INT64
rng = np.random.default_rng(0)
done = []


def ring_filters():
    for opts in ({}, {"ncw": 16}, {"ncw": 20, "stages": 4, "defer": 1}, {"stage_kb": 16, "nlb": 2}, {"staging": 0}, {"staging": 1},
                 {"list_cap": -1}, {"list_cap": 32, "nsw": -1}, {"list_cap": 64, "nsw": 3}, {"list_cap": 32, "nsw": 1, "stage_kb": 40}):
        c = Context(0)
        for k, v in opts.items():
            c.set_option(k, v)
        for n in (0, 1, 4097, 70_001):
            v = rng.integers(0, 1000, n).astype(np.int64)
            nb = rng.random(n) < 0.8
            x = rng.uniform(-1, 1, n)
            t = ColumnarTable.from_arrays(c, [Field("col", I, False), Field("x", D, True)], [v, x], [None, nb])
            for k in (10.0, 500.0, 990.0):
                r = Pipeline(t, fn(Function.CMP_LT, col(0, I), k), [col(0, I), col(1), fn(Function.CMP_GT, col(1), 0.0)]).run()
                want = np.nonzero(v < k)[0]
                assert r.num_rows == want.size and np.array_equal(r.selection(), want.astype(np.uint32))
                vals, ok = r.column(1)
                assert np.array_equal(ok, nb[want]) and np.array_equal(vals[ok], x[want][ok])
                bv, bok = r.column(2)
                assert np.array_equal(bok, nb[want]) and np.array_equal(bv[bok], (x[want] > 0.0)[bok])
                r.close()
            # capacity overflow: counted, not written
            if n > 4000:
                r = Pipeline(t, fn(Function.CMP_LT, col(0, I), 990.0), [col(0, I)], out_capacity_rows=100).run()
                try:
                    r.num_rows
                    raise AssertionError("overflow not reported")
                except capi.QeError as e:
                    assert e.status == capi.QE_ERR_CAPACITY
                r.close()
            t.close()
        c.close()
    done.append("ring / staged / register filter shapes")


def append_and_streams():
    c = Context(0)
    n, chunk = 9 * 4096 + 777, 4096
    v = rng.integers(0, 1000, n).astype(np.int64)
    want = np.nonzero(v < 400.0)[0]
    t = ColumnarTable.from_arrays(c, [Field("col", I, False)], [v])
    p = Pipeline(t, fn(Function.CMP_LT, col(0, I), 400.0), [col(0, I)], out_capacity_rows=n)
    r = p.run(None, 0, 4096)
    for lo in range(4096, n, 8192):
        r = p.run_append(r, lo, min(n, lo + 8192))
    assert np.array_equal(r.selection(), want.astype(np.uint32))
    r.close()
    hp = [C.c_void_p() for _ in range(3)]
    for h, nb in zip(hp, (n * 8, n * 4, n * 8)):
        capi.check(capi.lib().qe_host_alloc(nb, C.byref(h)))
    C.memmove(hp[0], v.ctypes.data, n * 8)
    for bounded in (False, True):
        tt = ColumnarTable(c, [Field("col", I, False)], 3 * chunk) if bounded else t
        pp = Pipeline(tt, fn(Function.CMP_LT, col(0, I), 400.0), [col(0, I)], out_capacity_rows=n)
        pa = Pipeline(tt, fn(Function.CMP_LT, col(0, I), 400.0), [col(0, I)], [A.SUM])
        res = stream_pipelines(tt, [pp, pa], [hp[0].value], chunk_rows=chunk, host_selection=[hp[1].value, None],
                               host_out_values=[[hp[2].value], None], total_rows=n if bounded else 0)
        assert res[0].num_rows == want.size
        sel = np.frombuffer((C.c_char * (want.size * 4)).from_address(hp[1].value), np.uint32)
        out = np.frombuffer((C.c_char * (want.size * 8)).from_address(hp[2].value), np.int64)
        assert np.array_equal(sel, want.astype(np.uint32)) and np.array_equal(out, v[want])
        assert res[1].agg(0) == float(v[want].sum())
        for x in res:
            x.close()
        if bounded:
            tt.close()
    for h in hp:
        capi.lib().qe_host_free(h)
    t.close()
    c.close()
    done.append("append, resident and bounded streaming")


def projection_and_aggregation():
    for opts in ({}, {"staging": 2}):
        c = Context(0)
        for k, v in opts.items():
            c.set_option(k, v)
        n = 50_003
        a, b = rng.uniform(-10, 10, n), rng.uniform(1, 2, n)
        vb = rng.random(n) < 0.9
        t = ColumnarTable.from_arrays(c, [Field("a", D, False), Field("b", D, True)], [a, b], [None, vb])
        r = Pipeline(t, None, [fn(Function.DIV, col(0), col(1)), fn(Function.CMP_LT, col(0), 0.0)]).run()
        vals, ok = r.column(0)
        assert np.array_equal(ok, vb) and np.array_equal(vals[ok], (a / b)[ok])
        assert np.array_equal(r.column(1)[0], a < 0.0)
        r.close()
        for interp in (False, True):
            r = Pipeline(t, fn(Function.CMP_GT, col(0), 0.0), [col(1), col(1), col(1)], [A.COUNT, A.MIN, A.MAX], interpreter=interp).run()
            m = (a > 0.0) & vb
            assert r.agg(0) == float(m.sum()) and r.agg(1) == b[m].min() and r.agg(2) == b[m].max()
            r.close()
        t.close()
        c.close()
    done.append("dense projection (grid-stride, ring), aggregation (generated, interpreter)")


def group_by():
    c = Context(0)
    n = 60_000
    seen = set()
    for ngroups, flags in ((4, {}), (4, {"registers": 0}), (40, {}), (200, {}), (1000, {}), (5000, {}), (6, {"atomic": True})):
        c.set_option("groupby_registers", flags.get("registers", 1))
        k = rng.integers(0, ngroups, n).astype(np.float64)
        cols = [k] + [np.round(rng.uniform(-5, 5, n), 1) for _ in range(4)]
        t = ColumnarTable.from_arrays(c, [Field("k", D, False)] + [Field(x, D, False) for x in "xyzw"], cols)
        wide = ngroups <= 8
        inputs = [col(1), col(2), col(3), col(4)] if wide else [col(1), col(1)]
        fns = [A.SUM, A.SUM, A.SUM, A.COUNT] if wide else [A.SUM, A.COUNT]
        r = GroupByPipeline(t, None, [col(0)], inputs, fns, atomic=flags.get("atomic", False)).run()
        seen.add(capi.lib().qe_result_group_variant(r.handle))
        keys = r.column(0)[0]
        counts = r.column(len(inputs))[0]
        order = {}
        for i, kk in enumerate(k):
            order.setdefault(kk, i)
        assert list(keys) == sorted(order, key=order.get)                # order of first appearance
        assert np.array_equal(counts, np.array([np.count_nonzero(k == kk) for kk in keys], dtype=np.float64))
        sums = r.column(1)[0]
        assert np.allclose(sums, [cols[1][k == kk].sum() for kk in keys], atol=1e-6)
        r.close()
        t.close()
    c.close()
    done.append(f"group-by variants {sorted(seen)} (3 registers, 0 private, 4 lane copies, 5 eight copies, 2 CTA-shared, 1 atomics)")


if __name__ == "__main__":
    which = sys.argv[1:] or ["ring", "streams", "project", "groupby"]
    if "ring" in which:
        ring_filters()
    if "streams" in which:
        append_and_streams()
    if "project" in which:
        projection_and_aggregation()
    if "groupby" in which:
        group_by()
    print("sanitizer subset ok:", "; ".join(done))
