"""BASELINE.json's configs at their FULL sizes (10^9 / 6·10^8 rows), checked on the device.

The oracle needs seconds per 10^7 rows, so at these sizes the kernels are checked
against an independent restatement of the SAME query in plain PyTorch ops on the
result buffers where they lie in HBM (`qe_result_device_ptr`; nothing is copied to
the host), plus the size-independent properties of the domain (sorted unique
selection vector, count + complement count = N, checksum of checksums).  torch is
the checker here, never the thing measured.  Per-row outputs are compared bit for
bit: torch's elementwise f64 kernels are one IEEE operation each (no contraction
across ops), which is also what the JVM's DADD/DSUB/DMUL/DDIV/DREM do
(M/evaluator/BytecodeCompiler.kt:328-332).

Needs a B200-sized HBM (the largest case holds ~70 GB); skipped below 100 GB.
"""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402  (the pipelines under test are the ones bench.py times)
from queryengine_b200 import capi  # noqa: E402
from queryengine_b200.ast import DataType, Function, col, fn  # noqa: E402
from queryengine_b200.engine import ColumnarTable, Context, Field, Pipeline  # noqa: E402

pytestmark = pytest.mark.gpu
I = DataType.INT64
CHUNK = 1 << 27                      # rows per torch pass: bounds the checker's scratch memory


class _Dev:
    """A device buffer of the library seen through __cuda_array_interface__ (no copy)."""

    def __init__(self, ptr, n, typestr):
        assert ptr, "null device pointer"
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": typestr, "data": (int(ptr), False), "version": 2}


def dev(torch, ptr, n, typestr):
    if n == 0:
        return torch.empty(0, device="cuda")
    return torch.as_tensor(_Dev(ptr, n, typestr), device="cuda")


@pytest.fixture(scope="module")
def torch_mod():
    import torch
    if torch.cuda.get_device_properties(0).total_memory < 100e9:
        pytest.skip("full-size configs need a B200-sized HBM")
    return torch


@pytest.fixture(scope="module")
def ctx():
    c = Context(0)
    yield c
    c.close()


def _bits(torch, ptr, row0, n):
    """rows [row0, row0+n) of an LSB-first bitmap as a bool tensor (row0 % 8 == 0)."""
    nb = (n + 7) // 8
    by = dev(torch, ptr + row0 // 8, nb, "|u1")
    sh = torch.arange(8, device="cuda", dtype=torch.uint8)
    return ((by[:, None] >> sh[None, :]) & 1).bool().reshape(-1)[:n]


@pytest.mark.parametrize("held_as", ["int64", "f64"])
def test_config_b_single_column_filter_1e9(ctx, torch_mod, held_as):
    """configs[1]: one column, 10^9 rows, `col < k` at 1 / 50 / 99 %: selection vector and compacted column.  The column is
    held as int64 (BASELINE.json) and as f64, the reference's own numeric type: same values."""
    torch = torch_mod
    n = 1_000_000_000
    ctype = I if held_as == "int64" else DataType.DOUBLE
    t = ColumnarTable(ctx, [Field("col", ctype, False)], n)
    t.generate(0, capi.QE_GEN_UNIFORM_U40, 42)
    ctx.synchronize()
    column = dev(torch, t.device_ptr(0, 0), n, "<i8" if held_as == "int64" else "<f8")
    bits = column.view(torch.int64)
    total = int(bits.sum())                                                # int64 wrap-around sum of the value bits
    for s in (0.01, 0.5, 0.99):
        k = int(s * 2 ** 40)
        res = Pipeline(t, fn(Function.CMP_LT, col(0, ctype), float(k)), [col(0, ctype)]).run()
        m = res.num_rows
        assert res.used_jit and abs(m / n - s) < 1e-4
        ctx.synchronize()
        sel = dev(torch, res.device_ptr(0, 2), m, "<i4")                   # row ids < 2^31
        vals = dev(torch, res.device_ptr(0, 0), m, "<i8")                  # value bits
        # the independent restatement, chunk by chunk: nonzero(col < k) and col[those rows]
        done = 0
        for r0 in range(0, n, CHUNK):
            c = column[r0:r0 + CHUNK]
            idx = torch.nonzero(c < k).reshape(-1)
            got = sel[done:done + idx.numel()]
            assert got.numel() == idx.numel() and bool((got.long() == idx + r0).all())
            assert bool((vals[done:done + idx.numel()] == bits[r0:r0 + CHUNK][idx]).all())
            done += idx.numel()
            del idx, got
        assert done == m
        # size-independent properties
        assert bool((sel[1:] > sel[:-1]).all())                            # ascending, unique
        assert bool((vals.view(column.dtype) < k).all())
        res2 = Pipeline(t, fn(Function.CMP_GE, col(0, ctype), float(k)), [col(0, ctype)]).run()
        m2 = res2.num_rows
        ctx.synchronize()
        assert m + m2 == n                                                 # complement keeps exactly the rest
        vals2 = dev(torch, res2.device_ptr(0, 0), m2, "<i8")
        assert (int(vals.sum()) + int(vals2.sum()) - total) % 2 ** 64 == 0   # checksum of checksums
        torch.cuda.synchronize()
        res.close()
        res2.close()
    t.close()


def test_config_c_q6_sum_600m(ctx, torch_mod):
    """configs[2]: Q6-shaped filter + SUM(extendedprice * discount) over 600 037 902 rows."""
    torch = torch_mod
    n = 600_037_902
    t, pipe = bench.q6_pipeline(ctx, n)
    res = pipe.run()
    part = res.agg_partial(0)
    ctx.synchronize()
    qty, ep, disc, sd = (dev(torch, t.device_ptr(c, 0), n, "<f8") for c in range(4))
    count, sums, abs_sum = 0, [], 0.0
    for r0 in range(0, n, CHUNK):
        sl = slice(r0, r0 + CHUNK)
        keep = (sd[sl] >= 8766.0) & (sd[sl] < 9131.0) & (disc[sl] >= 0.05) & (disc[sl] <= 0.07) & (qty[sl] < 24.0)
        prod = (ep[sl] * disc[sl])[keep]
        count += int(keep.sum())
        sums.append(float(prod.sum()))
        abs_sum += float(prod.abs().sum())
    torch.cuda.synchronize()
    assert part.count == count and 0.015 < count / n < 0.022              # COUNT is exact
    expect = float(np.sum(np.array(sums)))
    # any summation order of `count` terms is within (count-1)·2^-53·Σ|x| of the exact sum; two such sums
    # (the kernel's and torch's) differ by at most twice that.  Observed: a few ulp.
    bound = 2.0 * (count - 1) * 2.0 ** -53 * abs_sum
    assert abs(res.agg(0) - expect) <= bound
    assert abs(res.agg(0) - expect) <= 64 * np.spacing(expect)             # what is actually seen (terms are all positive)
    res.close()
    t.close()


def _tree_torch(torch, a, b, c, d):
    """bench.tree_expr() one IEEE operation at a time; CMP_GT is Double.compare (NaN > everything, Interpreter.kt:104-107)."""
    inner = ((a + b) * (c - d)) / ((a * 0.5) + (torch.fmod(c, 7.0) - (-b)))
    cond = (inner > 0.0) | torch.isnan(inner)
    then = (inner * (a - d)) + (c / (b + 1.5))
    other = (-inner) - (d * 0.25)
    return torch.where(cond, then, other)


def test_config_d_nullable_tree_1e9(ctx, torch_mod):
    """configs[3]: depth-8 nullable f64 tree with validity bitmaps over 10^9 rows: values and validity bit for bit."""
    torch = torch_mod
    n = 1_000_000_000
    D = DataType.DOUBLE
    t = ColumnarTable(ctx, [Field(x, D, True) for x in "abcd"], n)
    for c, rate in enumerate((0.0, 0.01, 0.10, 0.50)):
        t.generate(c, capi.QE_GEN_UNIFORM_REAL, bench.SEED + 10 + c, -1000.0, 1000.0, null_rate=rate)
    res = Pipeline(t, None, [bench.tree_expr()]).run()
    assert res.num_rows == n and res.used_jit
    ctx.synchronize()
    cols = [dev(torch, t.device_ptr(c, 0), n, "<f8") for c in range(4)]
    out = dev(torch, res.device_ptr(0, 0), n, "<f8")
    nulls = 0
    for r0 in range(0, n, CHUNK):
        m = min(CHUNK, n - r0)
        valid = _bits(torch, t.device_ptr(0, 1), r0, m)
        for c in range(1, 4):
            valid &= _bits(torch, t.device_ptr(c, 1), r0, m)
        got_valid = _bits(torch, res.device_ptr(0, 1), r0, m)
        assert bool((got_valid == valid).all())                            # null iff any leaf is null (every leaf is reached)
        exp = _tree_torch(torch, *(x[r0:r0 + m] for x in cols))
        got = out[r0:r0 + m]
        same = (got.view(torch.int64) == exp.view(torch.int64)) | (torch.isnan(got) & torch.isnan(exp))
        assert bool((same | ~valid).all())
        nulls += int((~valid).sum())
        del exp, same, valid, got_valid
    torch.cuda.synchronize()
    assert 0.54 < nulls / n < 0.57                                         # 1 - 0.99 * 0.9 * 0.5
    res.close()
    t.close()


def test_config_e_partition_filter_and_projections_1e9(ctx, torch_mod):
    """configs[4], one row-range partition of 10^9 rows: five-column AND/OR filter, two projections, selection vector."""
    torch = torch_mod
    n = 1_000_000_000
    t, pipe = bench.sf_filter_pipeline(ctx, n)
    res = pipe.run()
    m = res.num_rows
    ctx.synchronize()
    qty, ep, disc, tax, sd = (dev(torch, t.device_ptr(c, 0), n, "<f8") for c in range(5))
    sel = dev(torch, res.device_ptr(0, 2), m, "<i4")
    net = dev(torch, res.device_ptr(0, 0), m, "<f8")
    gross = dev(torch, res.device_ptr(1, 0), m, "<f8")
    done = 0
    for r0 in range(0, n, CHUNK):
        sl = slice(r0, r0 + CHUNK)
        keep = ((sd[sl] >= 8766.0) & (sd[sl] < 9131.0) & (qty[sl] < 24.0)) | ((disc[sl] >= 0.05) & (tax[sl] <= 0.02))
        idx = torch.nonzero(keep).reshape(-1)
        k = idx.numel()
        assert done + k <= m and bool((sel[done:done + k].long() == idx + r0).all())
        e_net = ep[sl][idx] * (1.0 - disc[sl][idx])
        e_gross = e_net * (1.0 + tax[sl][idx])
        assert bool((net[done:done + k].view(torch.int64) == e_net.view(torch.int64)).all())
        assert bool((gross[done:done + k].view(torch.int64) == e_gross.view(torch.int64)).all())
        done += k
        del idx, keep, e_net, e_gross
    torch.cuda.synchronize()
    assert done == m and 0.1 < m / n < 0.3
    res.close()
    t.close()
