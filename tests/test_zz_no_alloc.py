"""qe_agg_column_noalloc (include/qegpu.h, "no device allocation") against the existing entry point.

The function has no counterpart in the reference (it was added for callers that
cannot call cudaMalloc); its checker is therefore the library's own global aggregation through qe_pipeline_run /
qe_result_agg on the same device column, with the tolerance of tests/parity.py: COUNT / MIN / MAX bit-exact,
|SUM - exact| <= max(|sequential - exact|, 4 ulp(exact)), and bit-exact SUMs where every partial sum is exact.

Collected last on purpose; passes alone in a fresh process and after the other tests.
"""
import ctypes as C
import math

import numpy as np
import pytest

from queryengine_b200 import capi
from queryengine_b200.ast import AggregationFunction as A, DataType, col
from queryengine_b200.engine import ColumnarTable, Context, Field, Pipeline

pytestmark = pytest.mark.gpu
D = DataType.DOUBLE


@pytest.fixture(scope="module")
def ctx():
    c = Context(0)
    yield c
    c.close()


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    torch.cuda.set_device(0)
    return torch


def _existing(table):
    """[SUM, COUNT, MIN, MAX] of column 0 through the existing entry point."""
    r = Pipeline(table, None, [col(0), col(0), col(0), col(0)], aggregates=[A.SUM, A.COUNT, A.MIN, A.MAX]).run()
    got = [r.agg(k) for k in range(4)]
    r.close()
    return got


def _noalloc(torch, table, nrows, nullable, stream=None):
    out = torch.full((4,), 123.0, dtype=torch.float64, device="cuda:0")
    torch.cuda.synchronize()
    rc = capi.lib().qe_agg_column_noalloc(0, table.device_ptr(0, 0), table.device_ptr(0, 1) if nullable else None, nrows,
                                          out.data_ptr(), stream)
    capi.check(rc)
    torch.cuda.synchronize()
    return [float(x) for x in out.cpu().tolist()]


def _check(new, old, values, valid):
    kept = values[valid]
    if kept.size == 0:
        assert new == [0.0, 0.0, math.inf, -math.inf]
        return
    assert new[1] == float(kept.size) and old[1] == float(kept.size)
    assert np.float64(new[2]).tobytes() == np.float64(old[2]).tobytes() == np.float64(kept.min()).tobytes()
    assert np.float64(new[3]).tobytes() == np.float64(old[3]).tobytes() == np.float64(kept.max()).tobytes()
    exact = math.fsum(kept.tolist())
    sequential = float(np.cumsum(kept)[-1])
    tol = max(abs(sequential - exact), 4 * math.ulp(exact))
    assert abs(new[0] - exact) <= tol
    assert abs(old[0] - exact) <= tol
    assert abs(new[0] - old[0]) <= 2 * tol


def test_noalloc_equals_the_pipeline_aggregation_two_calls_in_a_row(ctx, torch_cuda):
    torch = torch_cuda
    rng = np.random.default_rng(20)
    cases = []
    # first the larger column, then a smaller one: the second call finds the first call's partials in the scratch array
    for n, null_rate in ((3 * capi.QE_AGG_SUPER_ROWS + 12_345, 0.1), (70_001, 0.5)):
        values = rng.uniform(-1000.0, 1000.0, n)
        valid = rng.random(n) >= null_rate
        t = ColumnarTable(ctx, [Field("x", D, True)], n)
        t.upload(0, values, valid)
        ctx.synchronize()
        cases.append((t, n, values, valid))
    news = [_noalloc(torch, t, n, True) for t, n, _, _ in cases]           # the two calls, back to back
    news_again = [_noalloc(torch, t, n, True) for t, n, _, _ in cases]
    assert news == news_again                                                # same bits whatever the array held before
    for (t, n, values, valid), new in zip(cases, news):
        _check(new, _existing(t), values, valid)
        t.close()


def test_noalloc_is_bit_identical_where_sums_are_exact(ctx, torch_cuda):
    torch = torch_cuda
    rng = np.random.default_rng(21)
    n = capi.QE_AGG_SUPER_ROWS + 16_384 + 7
    values = rng.integers(-(1 << 20), 1 << 20, n).astype(np.float64)           # every partial sum is an integer < 2^53
    t = ColumnarTable(ctx, [Field("x", D, False)], n)
    t.upload(0, values)
    ctx.synchronize()
    side = torch.cuda.Stream(device=0)                                         # a caller's own stream
    new = _noalloc(torch, t, n, False, stream=C.c_void_p(side.cuda_stream))
    old = _existing(t)
    assert new == [float(values.sum()), float(n), float(values.min()), float(values.max())]
    assert new == old
    # no rows
    assert _noalloc(torch, t, 0, False) == [0.0, 0.0, math.inf, -math.inf]
    # the steady state allocates nothing on the device (the kernels' code is loaded by now)
    free0, _ = torch.cuda.mem_get_info(0)
    for _ in range(3):
        assert _noalloc_into(torch, t, n) is None
    free1, _ = torch.cuda.mem_get_info(0)
    assert free1 >= free0
    t.close()


def _noalloc_into(torch, table, nrows, _out=[]):
    if not _out:
        _out.append(torch.zeros(4, dtype=torch.float64, device="cuda:0"))
    capi.check(capi.lib().qe_agg_column_noalloc(0, table.device_ptr(0, 0), None, nrows, _out[0].data_ptr(), None))
    torch.cuda.synchronize()
    return None


def test_noalloc_refuses_more_rows_than_the_scratch_array_holds(ctx, torch_cuda):
    torch = torch_cuda
    out = torch.zeros(4, dtype=torch.float64, device="cuda:0")
    t = ColumnarTable(ctx, [Field("x", D, False)], 1024)
    lib = capi.lib()
    rc = lib.qe_agg_column_noalloc(0, t.device_ptr(0, 0), None, capi.QE_NOALLOC_MAX_ROWS + 1, out.data_ptr(), None)
    assert rc == capi.QE_ERR_INVALID_ARGUMENT
    assert "QE_NOALLOC_MAX_ROWS" in lib.qe_last_error().decode()
    assert lib.qe_agg_column_noalloc(0, t.device_ptr(0, 0) + 8, None, 16, out.data_ptr(), None) == capi.QE_ERR_INVALID_ARGUMENT
    assert lib.qe_agg_column_noalloc(0, t.device_ptr(0, 0), None, 16, None, None) == capi.QE_ERR_INVALID_ARGUMENT
    assert lib.qe_agg_column_noalloc(0, t.device_ptr(0, 0), None, -1, out.data_ptr(), None) == capi.QE_ERR_INVALID_ARGUMENT
    t.close()
