"""Group-by aggregation on the GPU against the CPU oracle's restatement of
GroupByAggregationOperator (M/operator/GroupByAggregationOperator.kt:21-75).  Needs a B200.

Checked: the groups and their ORDER (first appearance, LinkedHashMap), key equality rules (null is a key,
NaN == NaN, -0.0 != 0.0, strings by dictionary code), null skipping, COUNT / MIN / MAX bit-exact,
SUM / AVG within the tolerance of SURVEY.md 8d (widened to the rigorous re-ordering bound (n-1) u sum|x| for
the mixed-sign test data), for the private-accumulator and the atomic variant.
"""
import math

import numpy as np
import pytest

from oracle_binding import OraclePlan, OracleTable
from parity import ulp
from queryengine_b200 import capi, sharding
from queryengine_b200.ast import AggregationFunction, DataType, Function, StringLiteralExpression, agg, col, fn
from queryengine_b200.engine import ColumnarTable, Context, Field, GpuGroupByOperator, GroupByPipeline

pytestmark = pytest.mark.gpu
S, D, B = DataType.STRING, DataType.DOUBLE, DataType.BOOLEAN
A = AggregationFunction


@pytest.fixture(scope="module")
def ctx():
    c = Context(0)
    yield c
    c.close()


def _gpu_table(ctx, fields, rows):
    cols, valid = [], []
    for c, f in enumerate(fields):
        vals = [r[c] for r in rows]
        valid.append(np.array([v is not None for v in vals], dtype=bool))
        if f.type == B:
            cols.append(np.array([bool(v) for v in vals], dtype=bool))
        elif f.type == S:
            cols.append(vals)
        else:
            cols.append(np.array([0.0 if v is None else v for v in vals], dtype=np.float64))
    if not rows:
        return ColumnarTable(ctx, fields, 0)
    return ColumnarTable.from_arrays(ctx, fields, cols, valid)


@pytest.mark.parametrize("atomic", [False, "private", True])
def test_querytest_group_by_multiple_columns_golden(ctx, atomic):
    # T/evaluator/QueryTest.kt:14-32, as written there: SELECT bar, SUM(num), foo FROM table
    fields = [Field("foo", S, True), Field("bar", S, True), Field("num", D, True)]
    rows = [["a", "A", 1.0], ["a", "B", 2.0], ["a", "B", 3.0], ["b", "B", 4.0], ["b", "B", None], ["c", None, None]]
    t = _gpu_table(ctx, fields, rows)
    ctx.set_option("groupby_registers", 0 if atomic == "private" else 1)
    try:
        op = GpuGroupByOperator(t, None, [col(1, S), agg(A.SUM, col(2)), col(0, S)], atomic=atomic is True)
        assert op.map() == [["A", 1.0, "a"], ["B", 5.0, "a"], ["B", 4.0, "b"], [None, None, "c"]]
        assert op.map() == [["A", 1.0, "a"], ["B", 5.0, "a"], ["B", 4.0, "b"], [None, None, "c"]]      # re-open
    finally:
        ctx.set_option("groupby_registers", 1)
    t.close()


def _same(a, b):
    if a is None or b is None:
        return a is None and b is None
    if isinstance(a, float) and isinstance(b, float):
        return (math.isnan(a) and math.isnan(b)) or (a == b and math.copysign(1.0, a) == math.copysign(1.0, b))
    return a == b


def _random_rows(rng, n, nstr, cardinality):
    vocab = ["g%d" % i for i in range(cardinality)]
    specials = [0.0, -0.0, float("nan"), 1.0, float("inf")]
    rows = []
    for _ in range(n):
        s = None if rng.random() < 0.1 else vocab[int(rng.integers(cardinality))]
        b = None if rng.random() < 0.1 else bool(rng.random() < 0.5)
        k = specials[int(rng.integers(len(specials)))] if rng.random() < 0.7 else None
        x = float(np.round(rng.uniform(-100, 100), 2))
        y = None if rng.random() < 0.3 else float(rng.uniform(-1e6, 1e6))
        rows.append([s, b, k, x, y])
    return rows


FIELDS = [Field("s", S, True), Field("b", B, True), Field("k", D, True), Field("x", D, False), Field("y", D, True)]


def _check_against_oracle(ctx, rows, keys, inputs, fns, filter_, atomic, expect_variant=None, max_groups=0, fields=None):
    if atomic == "private":                   # start with the shared-memory private variant instead of the register one
        ctx.set_option("groupby_registers", 0)
        try:
            return _check_against_oracle(ctx, rows, keys, inputs, fns, filter_, False, expect_variant, max_groups, fields)
        finally:
            ctx.set_option("groupby_registers", 1)
    nkeys = len(keys)
    FIELDS = fields if fields is not None else globals()["FIELDS"]
    t = _gpu_table(ctx, FIELDS, rows)
    shared = atomic == "shared"
    twopass = atomic == "twopass"
    atomic = atomic is True
    pipe = GroupByPipeline(t, filter_, keys, inputs, fns, atomic=atomic, shared=shared, twopass=twopass, max_groups=max_groups)
    res = pipe.run()
    if expect_variant is not None:
        allowed = expect_variant if isinstance(expect_variant, tuple) else (expect_variant,)
        assert capi.lib().qe_result_group_variant(res.handle) in allowed
    elif atomic or shared or twopass:
        assert capi.lib().qe_result_group_variant(res.handle) in ((1,) if atomic else (7,) if twopass else (2, 4, 5, 6))
    op = GpuGroupByOperator(t, filter_, list(keys) + [agg(f, e) for f, e in zip(fns, inputs)], atomic=atomic, shared=shared, twopass=twopass, max_groups=max_groups)
    got = op.map()
    if not rows:
        assert got == []
        t.close()
        return got
    otable = OracleTable.from_rows([f.type for f in FIELDS], rows)
    scan = list(range(len(FIELDS)))
    want = OraclePlan(otable, scan, filter=filter_, projections=list(keys) + list(inputs), aggregates=list(fns), group_count=nkeys).rows()
    # the terms every accumulator saw, per group, for the SUM / AVG tolerance
    per_row = OraclePlan(otable, scan, filter=filter_, projections=list(keys) + list(inputs)).rows()
    terms = {}
    for r in per_row:
        key = tuple("nan" if isinstance(v, float) and math.isnan(v) else (("-0" if math.copysign(1.0, v) < 0 else "+0") if v == 0.0 and isinstance(v, float) else v)
                    for v in r[:nkeys])
        terms.setdefault(key, [[] for _ in fns])
        for j in range(len(fns)):
            if r[nkeys + j] is not None:
                terms[key][j].append(r[nkeys + j])
    assert len(got) == len(want), (len(got), len(want))
    for g, w in zip(got, want):
        for k in range(nkeys):
            assert _same(g[k], w[k]), ("key / order", g, w)
        key = tuple("nan" if isinstance(v, float) and math.isnan(v) else (("-0" if math.copysign(1.0, v) < 0 else "+0") if v == 0.0 and isinstance(v, float) else v)
                    for v in w[:nkeys])
        for j, f in enumerate(fns):
            gv, wv = g[nkeys + j], w[nkeys + j]
            if wv is None or gv is None:
                assert gv is None and wv is None, (f, g, w)
            elif f in (A.COUNT, A.MIN, A.MAX):
                assert _same(float(gv), float(wv)), (f, g, w)
            else:
                tj = terms[key][j]
                if not all(math.isfinite(v) for v in tj):
                    assert not math.isfinite(gv) or _same(gv, wv)
                    continue
                div = len(tj) if f == A.AVG else 1
                exact = math.fsum(tj) / div
                # SURVEY.md 8d tolerance, widened for terms of mixed sign (these cancel: |sum| << sum|x|) to the
                # rigorous bound of ANY summation order, (n-1) * 2^-53 * sum|x|: the reference adds sequentially,
                # no parallel order can match that bit for bit
                tol = max(abs(wv - exact), 4 * ulp(exact), len(tj) * 2.0 ** -53 * math.fsum(abs(v) for v in tj) / div)
                assert abs(gv - exact) <= tol, (f, gv, wv, exact)
    res.close()
    t.close()
    return got


@pytest.mark.parametrize("atomic", [False, "private", "shared", "twopass", True])
@pytest.mark.parametrize("nrows", [0, 1, 300, 20000])
def test_groupby_matches_oracle_keys_order_and_aggregates(ctx, nrows, atomic):
    rng = np.random.default_rng(7 + nrows)
    rows = _random_rows(rng, nrows, 1, 3)
    s, b, k, x, y = (col(i, f.type) for i, f in enumerate(FIELDS))
    # keys: a nullable string, a nullable boolean, a double with NaN / +-0.0 / null
    _check_against_oracle(ctx, rows, [s, b], [x, y, y, x, y, s], [A.SUM, A.SUM, A.COUNT, A.MIN, A.AVG, A.COUNT], None, atomic)
    _check_against_oracle(ctx, rows, [k], [x, y, y], [A.MAX, A.MIN, A.AVG], fn(Function.CMP_GT, x, -50.0), atomic)
    # key expressions, not just columns: IF over strings and a comparison
    _check_against_oracle(ctx, rows, [fn(Function.IF, fn(Function.CMP_LT, x, 0.0), s, StringLiteralExpression("pos")),
                                      fn(Function.CMP_GE, y, 0.0)],
                          [fn(Function.MUL, x, y)], [A.SUM], fn(Function.CMP_NE, s, StringLiteralExpression("g1")), atomic)
    # no keys at all: one group (when any row arrives)
    # (one group in one partition outgrows the partition's space: the two-pass variant hands over to the atomics by design)
    _check_against_oracle(ctx, rows, [], [x, y], [A.SUM, A.COUNT], None, atomic if atomic != "twopass" else True) if nrows else None


@pytest.mark.parametrize("nrows", [1, 5000, 40000])
def test_register_variant_with_nulls_and_non_finite_inputs(ctx, nrows):
    # wide rows and few groups: the register variant (3).  SUM words are updated with a DFMA by a 0.0 / 1.0 mask, which
    # is only valid for finite inputs -- rows carrying Inf / NaN (null or not) must take the select path and must not
    # leak into the other groups
    fields = [Field("k", D, True), Field("x", D, True), Field("y", D, True), Field("z", D, False), Field("w", D, False),
              Field("s", S, False)]
    rng = np.random.default_rng(100 + nrows)
    specials = [float("inf"), float("-inf"), float("nan")]
    rows = []
    for i in range(nrows):
        k = None if rng.random() < 0.1 else float(rng.integers(0, 3))
        x = specials[int(rng.integers(3))] if (k == 2.0 and rng.random() < 0.01) else (None if rng.random() < 0.2 else float(np.round(rng.uniform(-100, 100), 2)))
        y = None if rng.random() < 0.3 else float(rng.uniform(-1e6, 1e6))
        z = float("inf") if (k == 1.0 and i % 977 == 5) else float(rng.integers(-1000, 1000))
        rows.append([k, x, y, z, float(rng.integers(0, 100)), "f%d" % int(rng.integers(2))])
    k, x, y, z, w, s = (col(i, f.type) for i, f in enumerate(fields))
    _check_against_oracle(ctx, rows, [k], [x, x, x, y, y, z, w, s], [A.SUM, A.AVG, A.MIN, A.MAX, A.COUNT, A.SUM, A.SUM, A.COUNT], None, False,
                          expect_variant=3, fields=fields)
    _check_against_oracle(ctx, rows, [k], [x, z, w, fn(Function.MUL, z, w)], [A.SUM, A.SUM, A.AVG, A.SUM], fn(Function.CMP_GE, w, 10.0), False,
                          expect_variant=3, fields=fields)


def test_eight_copies_variant_for_hundreds_of_groups(ctx):
    # more groups than the lane copies hold, few enough for eight copies per CTA (variant 5)
    rng = np.random.default_rng(808)
    n = 60000
    rows = [[None, bool(rng.random() < 0.5) if rng.random() > 0.1 else None, float(int(rng.integers(120))),
             float(np.round(rng.uniform(-10, 10), 1)), None if rng.random() < 0.2 else float(rng.uniform(-1e3, 1e3))] for _ in range(n)]
    b, k, x, y = col(1, B), col(2), col(3), col(4)
    # 360 groups of three words and two keys: more than the 32 lane copies hold (~250), few enough for eight copies
    got = _check_against_oracle(ctx, rows, [b, k], [x, y, x], [A.SUM, A.COUNT, A.AVG], None, False, expect_variant=5)
    assert len(got) == len({(r[1], r[2]) for r in rows})


def test_many_groups_fall_back_to_atomics_and_capacity_is_checked(ctx):
    rng = np.random.default_rng(3)
    n = 30000
    rows = [[None, None, float(int(rng.integers(5000))), float(np.round(rng.uniform(-10, 10), 1)), None] for _ in range(n)]
    k, x = col(2), col(3)
    # 5000 groups: no private variant holds them; a CTA of this small table meets only a few hundred, so the
    # CTA-shared variant takes over (a CTA that met more than its shared memory holds would fall back to atomics)
    got = _check_against_oracle(ctx, rows, [k], [x, x], [A.SUM, A.COUNT], None, False, expect_variant=(2, 4, 5, 6))
    assert len(got) == len({r[2] for r in rows})
    # ~1000 groups: more than the private copies hold, one shared-memory copy per CTA takes them
    rows1k = [[r[0], r[1], float(int(r[2]) % 1000), r[3], r[4]] for r in rows]
    got = _check_against_oracle(ctx, rows1k, [k], [x, x, x], [A.SUM, A.COUNT, A.MIN], None, False, expect_variant=(2, 4, 5, 6))
    assert len(got) == 1000
    t = _gpu_table(ctx, FIELDS, rows)
    with pytest.raises(capi.QeError) as e:
        GroupByPipeline(t, None, [k], [x], [A.SUM], max_groups=1000).run()
    assert "max_groups" in str(e.value)
    t.close()


def test_warp_copies_variant_for_up_to_a_thousand_groups(ctx):
    """One accumulator copy per warp, updated with plain loads and stores; lanes of a warp that meet the same group in
    one step take turns (match.any), so a group's rows are added in a fixed order: variant 6, reproducible.  Few distinct
    keys per step force the taking of turns (40 groups), many spread the accesses (700 groups, nulls, a filter)."""
    ctx.set_option("groupby_lanes", 0)                      # skip the lane-copies variants, which would hold these groups
    ctx.set_option("groupby_registers", 0)
    ctx.set_option("groupby_warps", 1)                      # off by default: it measured slower than the single shared copy
    try:
        for ngroups, seed in ((450, 1), (40, 2)):
            rng = np.random.default_rng(900 + seed)
            n = 120000
            rows = [[None, bool(rng.random() < 0.5) if rng.random() > 0.1 else None, float(int(rng.integers(ngroups))),
                     float(np.round(rng.uniform(-10, 10), 1)), None if rng.random() < 0.2 else float(rng.uniform(-1e3, 1e3))] for _ in range(n)]
            k, x, y = col(2), col(3), col(4)
            got = _check_against_oracle(ctx, rows, [k], [x, y, x, y, y], [A.SUM, A.COUNT, A.AVG, A.MIN, A.MAX], fn(Function.CMP_GT, x, -9.0), "shared",
                                        expect_variant=6)
            assert len(got) == ngroups
            again = _check_against_oracle(ctx, rows, [k], [x, y, x, y, y], [A.SUM, A.COUNT, A.AVG, A.MIN, A.MAX], fn(Function.CMP_GT, x, -9.0), "shared",
                                          expect_variant=6)
            assert all(_same(a, b) for ra, rb in zip(got, again) for a, b in zip(ra, rb))      # bit for bit, run to run
    finally:
        ctx.set_option("groupby_lanes", 1)
        ctx.set_option("groupby_registers", 1)
        ctx.set_option("groupby_warps", 0)


@pytest.mark.parametrize("ngroups", [30, 60])
def test_lane_copies_variant_for_tens_to_hundreds_of_groups(ctx, ngroups):
    # more groups than the thread-private copies hold, few enough for one copy per lane index (variant 4); then the same
    # query with that variant switched off (one copy per CTA, variant 2): same groups, same order
    rng = np.random.default_rng(50 + ngroups)
    n = 40000
    rows = [[("s%d" % int(rng.integers(3))) if rng.random() > 0.05 else None, None, float(int(rng.integers(ngroups // 3))),
             float(np.round(rng.uniform(-10, 10), 1)), None if rng.random() < 0.2 else float(rng.uniform(-1e3, 1e3))] for _ in range(n)]
    s, k, x, y = col(0, S), col(2), col(3), col(4)
    # three words per group (first row | rows, sum of x, count of y): up to 85 groups within the 64 KB of lane copies
    got = _check_against_oracle(ctx, rows, [s, k], [x, y, x], [A.SUM, A.COUNT, A.AVG], None, False, expect_variant=4)
    ctx.set_option("groupby_lanes", 0)
    try:
        got2 = _check_against_oracle(ctx, rows, [s, k], [x, y, x], [A.SUM, A.COUNT, A.AVG], None, False, expect_variant=2)
    finally:
        ctx.set_option("groupby_lanes", 1)
    assert [g[:2] for g in got] == [g[:2] for g in got2]


@pytest.mark.parametrize("registers", [1, 0])
def test_private_variant_is_reproducible_and_reports_partials(ctx, registers):
    # few groups: accumulators in registers (variant 3) by default, thread-private shared-memory copies (variant 0)
    # when that is switched off; both reduce in a fixed order
    ctx.set_option("groupby_registers", registers)
    try:
        _reproducible(ctx, 3 if registers else 0)
    finally:
        ctx.set_option("groupby_registers", 1)


def _reproducible(ctx, expect_variant):
    n = 2_000_000
    # 40-byte rows, five words per group: wide enough for the register variant's cost rule (plan_groupby)
    fields = [Field("g", D, False)] + [Field(c, D, False) for c in "xyzw"]
    t = ColumnarTable(ctx, fields, n)
    t.generate(0, capi.QE_GEN_UNIFORM_INT, 11, 0, 5)
    for c in range(1, 5):
        t.generate(c, capi.QE_GEN_UNIFORM_REAL, 11 + c, -1.0, 1.0)
    pipe = GroupByPipeline(t, None, [col(0)], [col(1), col(1), col(2), col(3), col(4)], [A.SUM, A.COUNT, A.SUM, A.SUM, A.SUM])
    r1 = pipe.run()
    assert capi.lib().qe_result_group_variant(r1.handle) == expect_variant
    a = [r1.column(k)[0].copy() for k in range(3)]
    r2 = pipe.run()
    b = [r2.column(k)[0].copy() for k in range(3)]
    assert r1.num_rows == 6
    for u, v in zip(a, b):
        assert u.tobytes() == v.tobytes()                     # same bits run to run
    assert a[2].sum() == n
    import ctypes as C
    first = np.zeros(6, np.int64)
    capi.check(capi.lib().qe_result_group_first_rows(r2.handle, first.ctypes.data))
    assert (np.diff(first) > 0).all() and first[0] == 0      # ordered by first appearance
    p = capi.AggPartial()
    capi.check(capi.lib().qe_result_group_partial(r2.handle, 0, 0, C.byref(p)))
    assert p.count == int(a[2][0]) and p.sum == a[1][0]
    r1.close(); r2.close(); t.close()


def test_row_range_partitions_merge_to_the_whole_table_result(ctx):
    # two partitions of one table grouped separately (what two GPUs would do), merged by key on the host
    from queryengine_b200 import sharding
    rng = np.random.default_rng(21)
    rows = _random_rows(rng, 6000, 1, 3)
    for r in rows:
        r[3] = float(int(r[3]))                              # integer-valued terms: every summation order is exact
    s, b, k, x, y = (col(i, f.type) for i, f in enumerate(FIELDS))
    fns = [A.SUM, A.COUNT, A.MIN, A.MAX]
    t = _gpu_table(ctx, FIELDS, rows)
    pipe = GroupByPipeline(t, fn(Function.CMP_GT, x, -90.0), [s, k], [x, y, x, x], fns)
    # fixed-width group records (key bits, null flags, global first row, partials) merged by qe_groupby_merge
    keys, nulls, first, partials, types = sharding.group_arrays(pipe.run(), 2, len(fns))
    whole = sharding.merged_rows(fns, types, sharding.merge_group_arrays(fns, 2, [(keys, nulls, first, partials)]), ctx.dictionary.decode)
    parts = []
    for lo, hi in ((0, 2432), (2432, 6000)):
        parts.append(sharding.group_arrays(pipe.run(row_begin=lo, row_end=hi), 2, len(fns), row_offset=0)[:4])
    merged = sharding.merged_rows(fns, types, sharding.merge_group_arrays(fns, 2, parts), ctx.dictionary.decode)
    assert len(whole) == len(merged) and len(whole) > 10
    for w, m in zip(whole, merged):
        assert all(_same(a, c) for a, c in zip(w, m)), (w, m)
    op = GpuGroupByOperator(t, fn(Function.CMP_GT, x, -90.0), [s, k, agg(A.SUM, x), agg(A.COUNT, y), agg(A.MIN, x), agg(A.MAX, x)])
    for w, m in zip(op.map(), merged):
        assert all(_same(a, c) for a, c in zip(w, m)), (w, m)
    t.close()


def test_int64_and_date32_key_columns(ctx):
    # INT64 / DATE32 have no reference counterpart: the oracle sees their exact f64 widening
    I, T = DataType.INT64, DataType.DATE32
    rng = np.random.default_rng(77)
    n = 10007
    ki = rng.integers(-3, 4, n).astype(np.int64) * (1 << 40)
    kd = rng.integers(9000, 9004, n).astype(np.int32)
    vd = rng.random(n) < 0.9
    x = np.round(rng.uniform(0, 10, n), 0)
    fields = [Field("ki", I, False), Field("kd", T, True), Field("x", D, False)]
    t = ColumnarTable.from_arrays(ctx, fields, [ki, kd, x], [None, vd, None])
    got = GpuGroupByOperator(t, None, [col(0, I), col(1, T), agg(A.SUM, col(2)), agg(A.COUNT, col(2))]).map()
    ot = OracleTable([D, D, D], n)
    ot.set_column(0, ki.astype(np.float64))
    ot.set_column(1, kd.astype(np.float64), np.packbits(vd.astype(np.uint8), bitorder="little"))
    ot.set_column(2, x)
    want = OraclePlan(ot, [0, 1, 2], projections=[col(0), col(1), col(2), col(2)], aggregates=[A.SUM, A.COUNT], group_count=2).rows()
    assert got == want and len(got) == 7 * 5
    t.close()


# five words per group (first row | rows, sum, count, min, max): 168 / 320 / 592 / 1024 / 1776 / 2048 groups per CTA with
# 32 / 16 / 8 / 4 / 2 / 1 copies.  First run: the ladder 32 -> 8 -> 1; second run: the most copies that hold the groups.
@pytest.mark.parametrize("ngroups,first_copies,again_copies", [(100, 32, 32), (250, 8, 16), (400, 8, 8), (800, 1, 4), (1200, 1, 2), (1800, 1, 1)])
def test_copies_variants_over_many_tiles_per_cta(ctx, ngroups, first_copies, again_copies):
    """The shared-memory variants with 32 / 16 / 8 / 4 / 2 copies per CTA on a table large enough that every CTA runs
    several epochs (32 tiles between two barriers): groups that first appear late are inserted in a later epoch, rows of
    older groups skip the first-row word.  Keys are not small integers (k * 0.37).  Checked against numpy: groups in order
    of first appearance, first rows, COUNT / MIN / MAX exact, SUM within the tolerance of tests/parity.py.  The second
    run reuses the result: it knows the group count and goes straight to the most copies that hold it."""
    rng = np.random.default_rng(ngroups)
    n = 148 * 2048 * 40 + 777
    k = rng.integers(0, ngroups // 2, n).astype(np.float64)
    late = n - n // 5
    k[late:] = rng.integers(0, ngroups, n - late).astype(np.float64)       # half of the groups first appear in the last fifth
    k[rng.integers(0, n, 5)] = np.nan
    k[rng.integers(0, n, 5)] = -0.0
    x = np.round(rng.uniform(-1000.0, 1000.0, n), 2)
    xv = rng.random(n) >= 0.1
    t = ColumnarTable.from_arrays(ctx, [Field("k", D, False), Field("x", D, True)], [k, x], [None, xv])
    key = fn(Function.MUL, col(0), 0.37)
    pipe = GroupByPipeline(t, None, [key], [col(1), col(1), col(1), col(1)], [A.SUM, A.COUNT, A.MIN, A.MAX])
    kk = k * 0.37
    bits = kk.view(np.uint64).copy()
    bits[np.isnan(kk)] = 0x7ff8000000000000
    uniq, first, inv = np.unique(bits, return_index=True, return_inverse=True)
    order = np.argsort(first, kind="stable")
    rank = np.empty(len(uniq), np.int64)
    rank[order] = np.arange(len(uniq))
    g_of_row = rank[inv]
    ng = len(uniq)
    cnt = np.bincount(g_of_row[xv], minlength=ng)
    res = None
    for copies in (first_copies, again_copies):
        res = pipe.run(res)
        assert capi.lib().qe_result_group_copies(res.handle) == copies
        assert res.num_rows == ng
        keys, nulls, first_row, parts, _ = sharding.group_arrays(res, 1, 4)
        assert np.array_equal(keys[:, 0], uniq[order]) and not nulls.any()
        assert np.array_equal(first_row, first[order])
        assert np.array_equal(parts[:, 1, 1].view(np.int64), cnt)
        xs = np.where(xv, x, np.inf)
        mn = np.full(ng, np.inf); np.minimum.at(mn, g_of_row, xs)
        mx = np.full(ng, -np.inf); np.maximum.at(mx, g_of_row, np.where(xv, x, -np.inf))
        assert np.array_equal(parts[:, 2, 2], mn) and np.array_equal(parts[:, 3, 3], mx)
        for g in rng.integers(0, ng, 40):
            terms = x[(g_of_row == g) & xv]
            exact = math.fsum(terms.tolist())
            seq = float(np.cumsum(terms)[-1]) if len(terms) else 0.0
            tol = max(abs(seq - exact), 4 * ulp(exact), len(terms) * 2.0 ** -53 * math.fsum(np.abs(terms).tolist()))
            assert abs(parts[g, 0, 0] - exact) <= tol
    res.close()
    t.close()


def _numpy_groups(keys_bits, x, xv):
    uniq, first, inv = np.unique(keys_bits, return_index=True, return_inverse=True)
    order = np.argsort(first, kind="stable")
    rank = np.empty(len(uniq), np.int64)
    rank[order] = np.arange(len(uniq))
    g = rank[inv]
    ng = len(uniq)
    cnt = np.bincount(g[xv], minlength=ng)
    mn = np.full(ng, np.inf); np.minimum.at(mn, g, np.where(xv, x, np.inf))
    mx = np.full(ng, -np.inf); np.maximum.at(mx, g, np.where(xv, x, -np.inf))
    return uniq[order], first[order], g, cnt, mn, mx


@pytest.mark.parametrize("ngroups", [20_000, 300_000])
def test_two_pass_variant_for_more_groups_than_a_cta_holds(ctx, ngroups):
    """More groups than any shared-memory variant holds: rows are partitioned by key hash (pass 1), one CTA per partition
    aggregates its records (pass 2): variant 7.  Keys are not small integers, inputs are nullable, NaN and -0.0 are keys.
    Against numpy: groups in order of first appearance, first rows, COUNT / MIN / MAX exact, SUM within the tolerance of
    tests/parity.py.  The second run knows the group count and uses more copies per partition."""
    rng = np.random.default_rng(ngroups)
    n = (1 << 22) + 1_234_567
    k = rng.integers(0, ngroups, n).astype(np.float64)
    k[rng.integers(0, n, 5)] = np.nan
    k[rng.integers(0, n, 5)] = -0.0
    x = np.round(rng.uniform(-1000.0, 1000.0, n), 2)
    xv = rng.random(n) >= 0.1
    t = ColumnarTable.from_arrays(ctx, [Field("k", D, False), Field("x", D, True)], [k, x], [None, xv])
    key = fn(Function.MUL, col(0), 0.37)
    pipe = GroupByPipeline(t, None, [key], [col(1), col(1), col(1), col(1)], [A.SUM, A.COUNT, A.MIN, A.MAX], max_groups=2 * ngroups)
    kk = k * 0.37
    bits = kk.view(np.uint64).copy()
    bits[np.isnan(kk)] = 0x7ff8000000000000
    uniq, first, g_of_row, cnt, mn, mx = _numpy_groups(bits, x, xv)
    ng = len(uniq)
    res = None
    copies = []
    for _ in range(2):
        res = pipe.run(res)
        assert capi.lib().qe_result_group_variant(res.handle) == 7
        copies.append(capi.lib().qe_result_group_copies(res.handle))
        assert res.num_rows == ng
        keys, nulls, first_row, parts, _ = sharding.group_arrays(res, 1, 4) if ng <= 30_000 else _group_arrays_fast(res)
        assert np.array_equal(keys[:, 0], uniq) and not nulls.any()
        assert np.array_equal(first_row, first)
        if parts is not None:
            assert np.array_equal(parts[:, 1, 1].view(np.int64), cnt)
            assert np.array_equal(parts[:, 2, 2], mn) and np.array_equal(parts[:, 3, 3], mx)
            for g in rng.integers(0, ng, 40):
                terms = x[(g_of_row == g) & xv]
                exact = math.fsum(terms.tolist())
                seq = float(np.cumsum(terms)[-1]) if len(terms) else 0.0
                tol = max(abs(seq - exact), 4 * ulp(exact), len(terms) * 2.0 ** -53 * math.fsum(np.abs(terms).tolist()))
                assert abs(parts[g, 0, 0] - exact) <= tol
        else:
            vals = [res.column(1 + j) for j in range(4)]          # SUM, COUNT, MIN, MAX as the reference's rows hold them
            assert np.array_equal(vals[1][0][cnt > 0].astype(np.int64), cnt[cnt > 0])
            ok = cnt > 0
            assert np.array_equal(vals[2][0][ok], mn[ok]) and np.array_equal(vals[3][0][ok], mx[ok])
            assert not vals[2][1][~ok].any()                      # MIN of no rows is null
            sums = np.bincount(g_of_row[xv], weights=x[xv], minlength=ng)
            assert np.allclose(vals[0][0][ok], sums[ok], rtol=1e-12, atol=1e-6)
    assert copies[1] >= copies[0]
    res.close()
    t.close()


def _group_arrays_fast(res):
    """keys / first rows of a large group-by result without one C call per group"""
    n = res.num_rows
    vals, valid = res.column(0)
    bits = np.ascontiguousarray(vals, np.float64).view(np.uint64).copy()
    bits[np.isnan(vals)] = 0x7ff8000000000000
    first = np.zeros(max(n, 1), np.int64)
    capi.check(capi.lib().qe_result_group_first_rows(res.handle, first.ctypes.data))
    return bits.reshape(-1, 1), ~np.asarray(valid, bool).reshape(-1, 1), first[:n], None, None


def test_two_pass_falls_back_to_atomics_when_one_key_has_most_rows(ctx):
    rng = np.random.default_rng(99)
    n = (1 << 22) + 4096
    k = rng.integers(0, 50_000, n).astype(np.float64)
    k[rng.random(n) < 0.6] = 7.0                                   # its partition outgrows the space a partition gets
    x = rng.integers(-100, 100, n).astype(np.float64)
    t = ColumnarTable.from_arrays(ctx, [Field("k", D, False), Field("x", D, False)], [k, x])
    pipe = GroupByPipeline(t, None, [col(0)], [col(1), col(1)], [A.SUM, A.COUNT], max_groups=131072)
    res = pipe.run()
    assert capi.lib().qe_result_group_variant(res.handle) == 1
    uniq, first, g_of_row, cnt, _, _ = _numpy_groups(k.view(np.uint64), x, np.ones(n, bool))
    assert res.num_rows == len(uniq)
    keys, _, first_row, _, _ = _group_arrays_fast(res)
    assert np.array_equal(keys[:, 0], uniq) and np.array_equal(first_row, first)
    sums = np.bincount(g_of_row, weights=x, minlength=len(uniq))
    assert np.array_equal(res.column(1)[0], sums) and np.array_equal(res.column(2)[0].astype(np.int64), cnt)
    res.close()
    t.close()
