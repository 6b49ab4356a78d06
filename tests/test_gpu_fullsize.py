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

On top of that, three random 32 Mi-row windows of every full-size result are checked against the ORACLE itself
(oracle/qe_oracle.c, the restatement pinned to the reference's vectors): the window's input rows are copied to the
host, the oracle runs the same plan over them on all host threads, and its rows must equal the slice of the
full-size GPU result that belongs to the window, bit for bit.

Needs a B200-sized HBM (the largest case holds ~70 GB); skipped below 100 GB.
"""
import ctypes as C
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


WINDOW = 32 << 20


def _windows(n, seed):
    """three random window starts (multiples of 2^20 rows)"""
    rng = np.random.default_rng(seed)
    return [int(x) << 20 for x in rng.integers(0, (n - WINDOW) >> 20, 3)]


def oracle_window(dtypes, host_cols, host_valid, filt, projs, aggregates=None):
    """The oracle's plan over the rows of one window, split over the host threads (ctypes releases the GIL).  Returns
    (row index inside the window, values[k], valid[k]) of the rows it emits, in scan order -- or, with aggregates, the
    list of per-slice (count, terms) needed to check a SUM."""
    from concurrent.futures import ThreadPoolExecutor
    from oracle_binding import OraclePlan, OracleTable
    from queryengine_b200.ast import Mode
    n = len(host_cols[0])
    threads = max(1, min(32, os.cpu_count() or 1))
    step = -(-n // threads) // 8 * 8 + 8
    D = DataType.DOUBLE

    def one(lo):
        hi = min(n, lo + step)
        t = OracleTable(list(dtypes) + [D], hi - lo)
        for c in range(len(dtypes)):
            v = None if host_valid[c] is None else np.packbits(host_valid[c][lo:hi].astype(np.uint8), bitorder="little")
            t.set_column(c, host_cols[c][lo:hi], v)
        t.set_column(len(dtypes), np.arange(lo, hi, dtype=np.float64))
        row = col(len(dtypes))
        plan = OraclePlan(t, list(range(len(dtypes) + 1)), filter=filt, projections=list(projs) + [row], mode=Mode.BYTECODE_COMPILER)
        return plan.columns()

    with ThreadPoolExecutor(threads) as ex:
        parts = list(ex.map(one, range(0, n, step)))
    k = len(projs)
    rows = np.concatenate([p[1][k] for p in parts]).astype(np.int64)
    vals = [np.concatenate([p[1][j] for p in parts]) for j in range(k)]
    valid = [np.concatenate([p[2][j] for p in parts]) for j in range(k)]
    return rows, vals, valid


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
        # the oracle itself over three random windows of the full-size result
        for a in _windows(n, 100 + int(s * 100)):
            host = column[a:a + WINDOW].cpu().numpy()
            rows_o, vals_o, valid_o = oracle_window([ctype], [host], [None], fn(Function.CMP_LT, col(0, ctype), float(k)), [col(0, ctype)])
            lo = int(torch.searchsorted(sel, torch.tensor([a], device="cuda", dtype=sel.dtype)).item())
            hi = int(torch.searchsorted(sel, torch.tensor([a + WINDOW], device="cuda", dtype=sel.dtype)).item())
            assert hi - lo == rows_o.size
            assert np.array_equal(sel[lo:hi].cpu().numpy().astype(np.int64), rows_o + a)
            got = vals[lo:hi].cpu().numpy()
            got = got.astype(np.float64) if held_as == "int64" else got.view(np.float64)
            assert np.array_equal(got, vals_o[0]) and valid_o[0].all()
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
    # the oracle itself: the same aggregation over three random row-range batches (windows) of the table
    import math
    from queryengine_b200 import workloads as W
    for a in _windows(n, 7):
        host = [x[a:a + WINDOW].cpu().numpy() for x in (qty, ep, disc, sd)]
        rows_o, vals_o, valid_o = oracle_window([DataType.DOUBLE] * 4, host, [None] * 4, W.q6_filter(), [W.q6_projection()])
        rw = pipe.run(None, a, a + WINDOW)
        pw = rw.agg_partial(0)
        assert pw.count == rows_o.size                                     # COUNT exact
        seq = 0.0
        for x in vals_o[0]:
            seq += float(x)                                                # the reference's own order (Accumulators.kt:44)
        exact = math.fsum(vals_o[0].tolist())
        assert abs(rw.agg(0) - exact) <= max(abs(seq - exact), 4 * math.ulp(exact))
        rw.close()
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
    # the oracle itself over three random windows (8 Mi rows each: the tree costs the oracle ~1 us per row and thread)
    wn = 8 << 20
    for a in _windows(n, 11):
        host = [x[a:a + wn].cpu().numpy() for x in cols]
        hv = [_bits(torch, t.device_ptr(c, 1), a, wn).cpu().numpy() for c in range(4)]
        rows_o, vals_o, valid_o = oracle_window([D] * 4, host, hv, None, [bench.tree_expr()])
        assert rows_o.size == wn
        gv = _bits(torch, res.device_ptr(0, 1), a, wn).cpu().numpy()
        assert np.array_equal(gv, valid_o[0])
        g = out[a:a + wn].cpu().numpy()
        same = (g.view(np.uint64) == vals_o[0].view(np.uint64)) | (np.isnan(g) & np.isnan(vals_o[0]))
        assert (same | ~gv).all()
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
    # the oracle itself over three random windows
    from queryengine_b200 import workloads as W
    for a in _windows(n, 13):
        host = [x[a:a + WINDOW].cpu().numpy() for x in (qty, ep, disc, tax, sd)]
        rows_o, vals_o, valid_o = oracle_window([DataType.DOUBLE] * 5, host, [None] * 5, W.lineitem_filter(), W.lineitem_projections())
        lo = int(torch.searchsorted(sel, torch.tensor([a], device="cuda", dtype=sel.dtype)).item())
        hi = int(torch.searchsorted(sel, torch.tensor([a + WINDOW], device="cuda", dtype=sel.dtype)).item())
        assert hi - lo == rows_o.size and np.array_equal(sel[lo:hi].cpu().numpy().astype(np.int64), rows_o + a)
        assert np.array_equal(net[lo:hi].cpu().numpy().view(np.uint64), vals_o[0].view(np.uint64))
        assert np.array_equal(gross[lo:hi].cpu().numpy().view(np.uint64), vals_o[1].view(np.uint64))
    res.close()
    t.close()


def test_config_e_full_sf1000_through_a_bounded_window(ctx, torch_mod):
    """configs[4] at its FULL size on one GPU: 5 999 989 709 rows (more than 2^32, more than the HBM holds) stream
    through a three-chunk device window (qe_pipeline_stream, bounded mode, columns generated chunk by chunk).  The same
    filter over the first chunks, run resident, must keep the same rows; counts add up chunk by chunk."""
    torch = torch_mod
    from queryengine_b200 import workloads as W
    from queryengine_b200.engine import stream_pipelines
    total, chunk = 5_999_989_709, 1 << 27
    fields = [Field(nm, tp, nl) for nm, tp, nl in W.lineitem_schema()]
    window = ColumnarTable(ctx, fields, 3 * chunk)
    gen = [(capi.QE_GEN_UNIFORM_INT, bench.SEED + 1, 1, 50), (capi.QE_GEN_CENTS, bench.SEED + 2, 90000, 10499999),
           (capi.QE_GEN_CENTS, bench.SEED + 3, 0, 10), (capi.QE_GEN_CENTS, bench.SEED + 5, 0, 8),
           (capi.QE_GEN_UNIFORM_INT, bench.SEED + 4, 8036, 10561)]
    nchunks = -(-total // chunk)
    kept = np.zeros(nchunks, np.int64)
    # the selection vector is downloaded (chunk-relative, ~1.3e9 x 4 B): pinned host buffer
    cap = int(total * 0.25)
    hp = C.c_void_p()
    capi.check(capi.lib().qe_host_alloc(cap * 4, C.byref(hp)))
    pipe = Pipeline(window, W.lineitem_filter(), W.lineitem_projections(), out_capacity_rows=cap)
    res = stream_pipelines(window, [pipe], [None] * 5, chunk_rows=chunk, total_rows=total, chunk_relative=True, gen=gen,
                           chunk_kept=[kept], host_selection=[hp.value])
    m = res[0].num_rows
    assert kept[-1] == m and (np.diff(kept) > 0).all() and 0.1 < m / total < 0.3
    sel = np.frombuffer((C.c_char * (m * 4)).from_address(hp.value), dtype=np.uint32)
    # chunk by chunk: ascending inside the chunk, below the chunk's row count
    bounds = np.concatenate([[0], kept])
    for c in (0, 1, nchunks // 2, nchunks - 2, nchunks - 1):
        sc = sel[bounds[c]:bounds[c + 1]].astype(np.int64)
        rows_c = min(chunk, total - c * chunk)
        assert (np.diff(sc) > 0).all() and sc[-1] < rows_c
    # against a resident run over the same global rows (first two chunks and the last, short one)
    for c in (0, 1, nchunks - 1):
        rows_c = min(chunk, total - c * chunk)
        t, p1 = bench.sf_filter_pipeline(ctx, rows_c, row_offset=c * chunk)
        r1 = p1.run()
        assert r1.num_rows == bounds[c + 1] - bounds[c]
        assert np.array_equal(r1.selection(), sel[bounds[c]:bounds[c + 1]])
        r1.close()
        t.close()
    res[0].close()
    window.close()
    capi.lib().qe_host_free(hp)


@pytest.mark.parametrize("ngroups,variant", [(1024, (2, 5)), (65536, (7,))])
def test_groupby_600m_against_torch(ctx, torch_mod, ngroups, variant):
    """SURVEY.md 8f next-2 at BASELINE size: SELECT k, SUM(x), COUNT(x), MIN(x), MAX(x) GROUP BY k over 6e8 rows, 1 024
    groups (shared-memory copies) and 65 536 groups (two passes), against an independent restatement in
    torch ops over the table's columns where they lie in HBM.  x holds integers, so every SUM is exact in any order and
    is compared bit for bit; groups must come out in order of first appearance (GroupByAggregationOperator.kt:22)."""
    torch = torch_mod
    from queryengine_b200.ast import AggregationFunction as A
    from queryengine_b200.engine import GroupByPipeline
    D = DataType.DOUBLE
    n = 600_037_902
    t = ColumnarTable(ctx, [Field("k", D, False), Field("x", D, False)], n)
    t.generate(0, capi.QE_GEN_UNIFORM_INT, 71, 0, ngroups - 1)
    t.generate(1, capi.QE_GEN_UNIFORM_INT, 72, -5000, 5000)
    ctx.synchronize()
    pipe = GroupByPipeline(t, None, [col(0)], [col(1)] * 4, [A.SUM, A.COUNT, A.MIN, A.MAX], max_groups=2 * ngroups)
    res = pipe.run()
    res = pipe.run(res)                                            # the run that knows the group count
    assert capi.lib().qe_result_group_variant(res.handle) in variant and res.num_rows == ngroups
    keys = res.column(0)[0]
    got = [res.column(1 + j)[0] for j in range(4)]
    first = np.zeros(ngroups, np.int64)
    capi.check(capi.lib().qe_result_group_first_rows(res.handle, first.ctypes.data))
    # torch: per-group sums / counts / extrema / first rows, chunk by chunk
    sums = torch.zeros(ngroups, dtype=torch.int64, device="cuda")
    cnts = torch.zeros(ngroups, dtype=torch.int64, device="cuda")
    mins = torch.full((ngroups,), 1 << 40, dtype=torch.int64, device="cuda")
    maxs = torch.full((ngroups,), -(1 << 40), dtype=torch.int64, device="cuda")
    firsts = torch.full((ngroups,), 1 << 40, dtype=torch.int64, device="cuda")
    for r0 in range(0, n, CHUNK):
        m = min(CHUNK, n - r0)
        k = dev(torch, t.device_ptr(0, 0) + 8 * r0, m, "<f8").long()
        x = dev(torch, t.device_ptr(1, 0) + 8 * r0, m, "<f8").long()
        sums.scatter_add_(0, k, x)
        cnts += torch.bincount(k, minlength=ngroups)
        mins.scatter_reduce_(0, k, x, "amin")
        maxs.scatter_reduce_(0, k, x, "amax")
        firsts.scatter_reduce_(0, k, torch.arange(r0, r0 + m, device="cuda"), "amin")
        del k, x
    order = torch.argsort(firsts).cpu().numpy()
    assert np.array_equal(keys.astype(np.int64), order)
    assert np.array_equal(first, firsts.cpu().numpy()[order])
    for g, want in zip(got, (sums, cnts, mins, maxs)):
        assert np.array_equal(g, want.cpu().numpy()[order].astype(np.float64))
    res.close()
    t.close()
