"""Several GPUs driven by ONE process, as a JVM adapter would ("Several GPUs in one JVM process"):
one context per device, contiguous row ranges, worker threads running concurrently, results combined on the host
in partition order.  Needs two GPUs; skipped on a single-GPU box (run with `gpurun --gpus 2`).
"""
import os
import sys
import threading

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from queryengine_b200 import capi  # noqa: E402
from queryengine_b200.ast import AggregationFunction as A, DataType, Function, col, fn  # noqa: E402
from queryengine_b200.engine import ColumnarTable, Context, Field, Pipeline, combine_aggregate  # noqa: E402

pytestmark = pytest.mark.gpu
I = DataType.INT64


def _device_count():
    import torch
    return torch.cuda.device_count()


@pytest.fixture(scope="module")
def contexts():
    n = _device_count()
    if n < 2:
        pytest.skip("needs two GPUs")
    cs = [Context(d) for d in range(min(n, 4))]
    yield cs
    for c in cs:
        c.close()


def _run_threads(fns):
    out, errs = [None] * len(fns), []

    def work(i):
        try:
            out[i] = fns[i]()
        except Exception as e:  # noqa: BLE001
            errs.append(e)
    ts = [threading.Thread(target=work, args=(i,)) for i in range(len(fns))]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errs, errs
    return out


def test_row_range_partitions_on_several_devices_equal_one_device(contexts):
    g = len(contexts)
    n = 8_000_000                                             # a multiple of 128 per partition
    per = n // g // 128 * 128
    bounds = [(r * per, n if r == g - 1 else (r + 1) * per) for r in range(g)]
    k = float(int(0.3 * 2 ** 40))

    # one device, the whole table
    whole = ColumnarTable(contexts[0], [Field("col", I, False)], n)
    whole.generate(0, capi.QE_GEN_UNIFORM_U40, 42)
    res = Pipeline(whole, fn(Function.CMP_LT, col(0, I), k), [col(0, I)]).run()
    sel_whole, vals_whole = res.selection().astype(np.int64), res.column(0)[0].copy()
    res.close()
    whole.close()

    def part(r):
        lo, hi = bounds[r]
        t = ColumnarTable(contexts[r], [Field("col", I, False)], hi - lo)
        t.generate(0, capi.QE_GEN_UNIFORM_U40, 42, row_offset=lo)       # values are a function of the global row index
        out = []
        for _ in range(3):                                             # repeated, concurrently with the other devices
            rr = Pipeline(t, fn(Function.CMP_LT, col(0, I), k), [col(0, I)]).run()
            out = [rr.selection().astype(np.int64) + lo, rr.column(0)[0].copy()]
            rr.close()
        t.close()
        return out
    parts = _run_threads([lambda r=r: part(r) for r in range(g)])
    assert np.array_equal(np.concatenate([p[0] for p in parts]), sel_whole)      # rank-order concatenation = scan order
    assert np.array_equal(np.concatenate([p[1] for p in parts]), vals_whole)


def test_partial_aggregates_of_several_devices_combine_in_partition_order(contexts):
    g = len(contexts)
    n = 6_000_128
    per = n // g // 128 * 128
    bounds = [(r * per, n if r == g - 1 else (r + 1) * per) for r in range(g)]
    t0, pipe0 = bench.q6_pipeline(contexts[0], n)
    r0 = pipe0.run()
    want_sum, want_count = r0.agg(0), r0.agg_partial(0).count
    r0.close()
    t0.close()

    def part(r):
        lo, hi = bounds[r]
        t, pipe = bench.q6_pipeline(contexts[r], hi - lo, row_offset=lo)
        rr = pipe.run()
        p = rr.agg_partial(0)
        rr.close()
        t.close()
        return p
    partials = _run_threads([lambda r=r: part(r) for r in range(g)])
    assert sum(p.count for p in partials) == want_count                         # COUNT is exact
    got = combine_aggregate(A.SUM, partials)
    assert abs(got - want_sum) <= 16 * np.spacing(want_sum)                      # positive terms, different grouping of the adds
    again = combine_aggregate(A.SUM, _run_threads([lambda r=r: part(r) for r in range(g)]))
    assert again == got                                                          # reproducible: fixed order on every level
