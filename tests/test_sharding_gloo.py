"""Multi-partition host logic (queryengine_b200/sharding.py) on CPU: world_size 2 over gloo.

The per-partition pipeline results are produced by the oracle here (no GPU in this suite); what is under
test is the partitioning, the all-gather of partial aggregates, the rank-order combine through the C ABI
(qe_agg_combine needs no device) and the selection-vector rebasing.
"""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from queryengine_b200 import capi, sharding  # noqa: E402
from queryengine_b200.ast import AggregationFunction as A, DataType, Function, col, fn  # noqa: E402

D = DataType.DOUBLE
N = 100_003
FNS = [A.SUM, A.COUNT, A.MIN, A.MAX, A.AVG]


def _table(n=N):
    rng = np.random.default_rng(7)
    a = np.round(rng.uniform(-1000, 1000, n), 2)
    b = rng.integers(0, 11, n).astype(np.float64) / 100.0
    vb = rng.random(n) > 0.1                       # 10 % nulls in b
    return a, b, vb


def _filter():
    return fn(Function.AND, fn(Function.CMP_GE, col(0), -250.0), fn(Function.CMP_LT, col(1), 0.07))


def _proj():
    return fn(Function.MUL, col(0), fn(Function.SUB, 1.0, col(1)))


def _oracle_rows(a, b, vb, aggregates=None):
    from oracle_binding import OraclePlan, OracleTable
    t = OracleTable([D, D], len(a))
    t.set_column(0, a)
    t.set_column(1, b, np.packbits(np.asarray(vb, dtype=np.uint8), bitorder="little"))
    nproj = len(aggregates) if aggregates else 1          # accumulator k consumes projection k
    return t, OraclePlan(t, [0, 1], _filter(), [_proj()] * nproj, aggregates)


def _partial_of_shard(a, b, vb):
    """What qe_result_agg_partial returns for one partition, computed with the oracle's per-row outputs."""
    t, plan = _oracle_rows(a, b, vb)
    n, vals, valid = plan.columns()
    v = vals[0][valid[0]]
    s = 0.0
    for x in v:                                    # sequential, like SumAccumulator (Accumulators.kt:44)
        s += float(x)
    if v.size == 0:
        return capi.AggPartial(0.0, 0, float("inf"), float("-inf"))
    return capi.AggPartial(s, int(v.size), float(v.min()), float(v.max()))


def test_partition_covers_rows_once_and_is_aligned():
    for n in (0, 1, 127, 128, 129, 1000, N, 6_000_000_000):
        for world in (1, 2, 3, 4, 8):
            parts = sharding.partitions(n, world)
            assert parts[0][0] == 0 and parts[-1][1] == n
            for (b0, e0), (b1, e1) in zip(parts, parts[1:]):
                assert e0 == b1 and b0 <= e0
            assert all(b % sharding.ALIGN == 0 or b == n for b, _ in parts)
            sizes = [e - b for b, e in parts]
            assert max(sizes) <= -(-n // world) + sharding.ALIGN


def test_single_process_combine_matches_global_oracle():
    a, b, vb = _table()
    parts = sharding.partitions(N, 4)
    by_rank = [[_partial_of_shard(a[lo:hi], b[lo:hi], vb[lo:hi])] * len(FNS) for lo, hi in parts]
    got = sharding.combine(FNS, by_rank)
    _, plan = _oracle_rows(a, b, vb, FNS)
    want = plan.rows()[0]
    assert got[1] == want[1] and got[2] == want[2] and got[3] == want[3]
    assert abs(got[0] - want[0]) <= 1e-9 * abs(want[0]) and abs(got[4] - want[4]) <= 1e-9 * abs(want[4])
    # empty partitions are skipped, an all-empty input is NULL (COUNT is 0)
    empty = capi.AggPartial(0.0, 0, float("inf"), float("-inf"))
    assert sharding.combine([A.SUM, A.COUNT, A.MIN], [[empty] * 3, [empty] * 3]) == [None, 0.0, None]


def test_super_tile_lists_fold_to_the_same_bits_for_any_partition_count():
    """SUM over lists of super-tile partials (qe_result_agg_partials) concatenated in partition order: the fold is the
    same sequence of additions for 1, 2, 3 or 8 partitions, so the bits agree -- unlike a fold of per-partition totals."""
    rng = np.random.default_rng(3)
    n = 4096                                                   # super-tiles of the whole table
    lists = np.zeros((n, 2, 4))
    lists[:, 0, 0] = rng.uniform(-1e9, 1e9, n)
    lists[:, 0, 1] = np.full(n, 1000, np.int64).view(np.float64)
    lists[:, 1, :] = lists[:, 0, :]
    lists[:, 0, 2] = lists[:, 1, 2] = -5.0
    lists[:, 0, 3] = lists[:, 1, 3] = 7.0
    want = sharding.combine_lists([A.SUM, A.AVG], lists)
    seq = 0.0
    for x in lists[:, 0, 0]:
        seq += float(x)
    assert want[0] == seq and want[1] == seq / (1000 * n)
    for world in (2, 3, 8):
        per = -(-n // world)
        chunks = [lists[r * per:(r + 1) * per] for r in range(world)]
        assert sharding.combine_lists([A.SUM, A.AVG], np.concatenate(chunks, axis=0)) == want
    totals = [capi.AggPartial(float(np.sum(lists[r * 512:(r + 1) * 512, 0, 0])), 512000, -5.0, 7.0) for r in range(8)]
    assert sharding.combine([A.SUM], [[t] for t in totals])[0] != want[0]       # what round 1 did: depends on the split


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    try:
        import torch.distributed as dist
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        a, b, vb = _table()
        lo, hi = sharding.partition(N, world, rank)
        mine = [_partial_of_shard(a[lo:hi], b[lo:hi], vb[lo:hi])] * len(FNS)
        by_rank = sharding.all_gather_partials(mine)
        got = sharding.combine(FNS, by_rank)
        # super-tile partial lists of unequal length: gathered in rank order = row order
        whole = np.random.default_rng(5).uniform(-1, 1, (8, len(FNS), 4))
        cut = 5
        lists = sharding.all_gather_partial_lists(whole[:cut] if rank == 0 else whole[cut:])
        assert lists.shape == whole.shape and np.array_equal(lists, whole)
        # selection vectors: partition-relative + base, concatenated in rank order
        keep = (a[lo:hi] >= -250.0) & (b[lo:hi] < 0.07) & vb[lo:hi]
        local = np.nonzero(keep)[0].astype(np.uint32)
        import torch
        g = sharding.global_selection(local, lo)
        sizes = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(sizes, torch.tensor([g.size], dtype=torch.int64))
        q.put((rank, got, [int(s.item()) for s in sizes], int(g[0]) if g.size else -1, int(g[-1]) if g.size else -1,
               [(p.count, p.sum) for p in (r[0] for r in by_rank)]))
        dist.destroy_process_group()
    except Exception as e:            # surfaced by the parent
        q.put((rank, repr(e)))


@pytest.mark.timeout(300)
def test_world_size_2_gloo_all_gather_and_combine():
    import torch.multiprocessing as mp
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(60)
    res.sort(key=lambda r: r[0])
    for r in res:
        assert len(r) > 2, r
    a, b, vb = _table()
    _, plan = _oracle_rows(a, b, vb, FNS)
    want = plan.rows()[0]
    for rank, got, sizes, first, last, seen in res:
        assert got == res[0][1]                                   # every rank derives the same result, bit for bit
        assert got[1] == want[1] and got[2] == want[2] and got[3] == want[3]
        assert abs(got[0] - want[0]) <= 1e-9 * abs(want[0])
        assert sum(sizes) == int(np.count_nonzero((a >= -250.0) & (b < 0.07) & vb))
        assert seen == res[0][5]                                  # gathered in rank order on every rank
    keep = (a >= -250.0) & (b < 0.07) & vb
    sel = np.nonzero(keep)[0]
    lo1, _ = sharding.partition(N, 2, 1)
    assert res[0][3] == sel[0] and res[1][4] == sel[-1]
    assert res[1][3] == sel[sel >= lo1][0]
    assert sum(res[0][2]) == sel.size


# ---------------------------------------------------------------- group-by across partitions

def _group_rows():
    rng = np.random.default_rng(11)
    n = 4000
    flag = [None if rng.random() < 0.1 else ["A", "N", "R"][int(rng.integers(3))] for _ in range(n)]
    k = [[0.0, -0.0, float("nan"), 2.0][int(rng.integers(4))] for _ in range(n)]
    x = [None if rng.random() < 0.2 else float(int(rng.integers(-100, 100))) for _ in range(n)]
    return [[flag[i], k[i], x[i]] for i in range(n)]


def _shard_groups(rows, lo, hi, dictionary):
    """What sharding.group_arrays returns for rows [lo, hi): fixed-width arrays, computed here with plain Python
    (no GPU in this suite).  `dictionary`: the process's StringDictionary (codes are private to a process)."""
    S = DataType.STRING
    groups = {}
    for i in range(lo, hi):
        flag, k, x = rows[i]
        kid = (None if flag is None else dictionary.encode(flag), sharding.canonical_key_bits(k, D))
        g = groups.setdefault(kid, [i, [0.0, 0, float("inf"), float("-inf")]])
        if x is not None:
            a = g[1]
            a[0] += x; a[1] += 1; a[2] = min(a[2], x); a[3] = max(a[3], x)
    n = len(groups)
    keys = np.zeros((n, 2), np.uint64)
    nulls = np.zeros((n, 2), np.uint8)
    first = np.zeros(n, np.int64)
    parts = np.zeros((n, 4, 4), np.float64)
    for g, ((code, kbits), (i, a)) in enumerate(groups.items()):
        keys[g] = (0 if code is None else code, kbits)
        nulls[g, 0] = code is None
        first[g] = i
        parts[g, :] = (a[0], np.array([a[1]], np.int64).view(np.float64)[0], a[2], a[3])
    return keys, nulls, first, parts


GFNS = [A.SUM, A.COUNT, A.MIN, A.MAX]


def _oracle_groups(rows):
    from oracle_binding import OraclePlan, OracleTable
    S = DataType.STRING
    t = OracleTable.from_rows([S, D, D], rows)
    from queryengine_b200.ast import col as c
    return OraclePlan(t, [0, 1, 2], projections=[c(0, S), c(1), c(2), c(2), c(2), c(2)], aggregates=GFNS, group_count=2).rows()


def _rows_equal(got, want):
    import math
    assert len(got) == len(want), (len(got), len(want))
    for g, w in zip(got, want):
        for a, b in zip(g, w):
            if isinstance(a, float) and isinstance(b, float):
                assert (math.isnan(a) and math.isnan(b)) or (a == b and math.copysign(1, a) == math.copysign(1, b)), (g, w)
            else:
                assert a == b, (g, w)


def test_group_partials_merge_by_key_in_first_appearance_order():
    from queryengine_b200.engine import StringDictionary
    rows = _group_rows()
    parts = sharding.partitions(len(rows), 3)
    dic = StringDictionary()
    by_rank = [_shard_groups(rows, lo, hi, dic) for lo, hi in parts]
    merged = sharding.merge_group_arrays(GFNS, 2, by_rank)                       # qe_groupby_merge, fixed-width arrays
    got = sharding.merged_rows(GFNS, [DataType.STRING, D], merged, dic.decode)
    _rows_equal(got, _oracle_groups(rows))                                       # integer-valued x: sums are exact
    assert list(merged[2]) == sorted(merged[2])                                  # ordered by first appearance


def _group_worker(rank, world, port, q):
    try:
        import torch.distributed as dist
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        from queryengine_b200.engine import StringDictionary
        rows = _group_rows()
        lo, hi = sharding.partition(len(rows), world, rank)
        dic = StringDictionary()
        if rank == 1:
            dic.encode("zz")                                     # the ranks' codes differ: the merge must re-code them
        keys, nulls, first, parts = _shard_groups(rows, lo, hi, dic)

        class FakeResult:                                        # stands in for a qe_groupby_run result of this partition
            pass
        orig = sharding.group_arrays
        sharding.group_arrays = lambda result, nkeys, nagg, row_offset=0: (keys, nulls, first, parts, [DataType.STRING, D])
        try:
            got = sharding.sharded_groupby(FakeResult(), 2, GFNS, dic, 0)
        finally:
            sharding.group_arrays = orig
        q.put((rank, got))
        dist.destroy_process_group()
    except Exception as e:
        q.put((rank, repr(e)))


@pytest.mark.timeout(300)
def test_world_size_2_gloo_groupby_merge():
    import torch.multiprocessing as mp
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_group_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=240) for _ in range(world)], key=lambda r: r[0])
    for p in procs:
        p.join(60)
    want = _oracle_groups(_group_rows())
    for rank, got in res:
        assert isinstance(got, list), got
        _rows_equal(got, want)
