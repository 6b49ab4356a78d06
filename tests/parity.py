"""Shared helpers of the parity tests: build the same table for the oracle and
for the GPU path, run the same plan through both and compare.

Comparison rules ("parity"):
  * row counts, selection vectors, validity, booleans, integers: exact
  * doubles: bit-exact, except that every NaN is one value (the JVM does not
    canonicalise NaN payloads either; Double.equals / Double.compare treat all
    NaNs alike, M/evaluator/Interpreter.kt:102-107)
  * SUM / AVG: |gpu - exact| <= max(|sequential - exact|, 4 ulp(exact))  
"""
from __future__ import annotations

import math
from typing import List, Optional, Sequence

import numpy as np

from oracle_binding import OraclePlan, OracleTable
from queryengine_b200.ast import AggregationFunction, ColumnExpression, DataType, Expression, Mode
from queryengine_b200.engine import ColumnarTable, Context, Field, Pipeline

SPECIALS = np.array([0.0, -0.0, 1.0, -1.0, np.inf, -np.inf, np.nan, 5e-324, -5e-324, 1.7976931348623157e308,
                     2.0 ** 53, -(2.0 ** 53), 0.1, 0.5, 7.0, -7.0])


class TableSpec:
    """Columns as numpy arrays + optional validity masks, with the reference-visible dtypes."""

    def __init__(self, fields: Sequence[Field], columns: Sequence[np.ndarray], valid: Sequence[Optional[np.ndarray]]):
        self.fields = list(fields)
        self.columns = list(columns)
        self.valid = list(valid)
        self.nrows = len(columns[0]) if columns else 0

    def with_row_index(self) -> "TableSpec":
        """Append a non-null DOUBLE column holding the row index.  The reference's
        FilterOperator returns the scanned row unchanged, so filtering this column
        yields the selection vector in the reference's own terms."""
        idx = np.arange(self.nrows, dtype=np.float64)
        return TableSpec(self.fields + [Field("__row", DataType.DOUBLE, False)], self.columns + [idx], self.valid + [None])

    def oracle(self) -> OracleTable:
        t = OracleTable([f.type for f in self.fields], self.nrows)
        for c, f in enumerate(self.fields):
            v = None if self.valid[c] is None else np.packbits(self.valid[c].astype(np.uint8), bitorder="little")
            t.set_column(c, self.columns[c], v)
        return t

    def gpu(self, ctx: Context) -> ColumnarTable:
        return ColumnarTable.from_arrays(ctx, self.fields, self.columns, self.valid)


def random_table(rng: np.random.Generator, nrows: int, spec: str, special_rate: float = 0.05) -> TableSpec:
    """spec: one char per column: d/D double (non-null / nullable), b/B boolean, i/I int64, t/T date32."""
    fields, cols, valid = [], [], []
    for k, ch in enumerate(spec):
        nullable = ch.isupper()
        c = ch.lower()
        if c == "d":
            v = rng.uniform(-1000.0, 1000.0, nrows)
            # a few small integers so equality / MOD hit interesting cases
            small = rng.random(nrows) < 0.2
            v[small] = rng.integers(-5, 6, int(small.sum())).astype(np.float64)
            sp = rng.random(nrows) < special_rate
            v[sp] = rng.choice(SPECIALS, int(sp.sum()))
            dt = DataType.DOUBLE
        elif c == "b":
            v = rng.random(nrows) < 0.5
            dt = DataType.BOOLEAN
        elif c == "i":
            v = rng.integers(-(1 << 40), 1 << 40, nrows, dtype=np.int64)
            small = rng.random(nrows) < 0.3
            v[small] = rng.integers(-5, 6, int(small.sum()))
            dt = DataType.INT64
        elif c == "t":
            v = rng.integers(8000, 11000, nrows, dtype=np.int64).astype(np.int32)
            dt = DataType.DATE32
        else:
            raise ValueError(ch)
        fields.append(Field(f"c{k}", dt, nullable))
        cols.append(v)
        valid.append((rng.random(nrows) < 0.8) if nullable else None)
    return TableSpec(fields, cols, valid)


def f64_equal(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    """bit-exact, all NaNs equal"""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return (a.view(np.uint64) == b.view(np.uint64)) | (np.isnan(a) & np.isnan(b))


def ulp(x: float) -> float:
    return math.ulp(x) if math.isfinite(x) else float("inf")


def check_sum(gpu: float, terms: np.ndarray, sequential: float) -> None:
    if not np.isfinite(terms).all() or not math.isfinite(sequential):
        # Inf / NaN terms or an overflowing sequential sum: the order of additions decides between
        # Inf and NaN, so only "not finite" is comparable
        assert not math.isfinite(gpu) or not np.isfinite(terms).all(), (gpu, sequential)
        if np.isnan(terms).any():
            assert math.isnan(gpu)
        return
    exact = math.fsum(terms.tolist())
    tol = max(abs(sequential - exact), 4 * ulp(exact))
    assert abs(gpu - exact) <= tol, (gpu, exact, sequential, tol)


def compare_pipeline(ctx: Context, spec: TableSpec, filter: Optional[Expression],
                     projections: Sequence[Expression], interpreter: bool = False,
                     gpu_table: Optional[ColumnarTable] = None, mode: Mode = Mode.GPU) -> int:
    """Scan -> [Filter] -> Projection through the oracle and the GPU; asserts equality.

    mode: the evaluator whose semantics both sides follow (Mode.GPU = BYTECODE_COMPILER's; Mode.CLOSURE_COMPILER: IEEE
    ordering comparisons).

    The selection vector is checked by projecting the row-index column alongside."""
    full = spec.with_row_index()
    row_col = len(full.fields) - 1
    row_expr = ColumnExpression("__row", row_col, DataType.DOUBLE)
    projs = list(projections) + [row_expr]

    scan = list(range(len(full.fields)))
    plan = OraclePlan(full.oracle(), scan, filter=filter, projections=projs, mode=mode)
    n_ref, ref_vals, ref_valid = plan.columns()

    own_table = gpu_table is None
    table = gpu_table or full.gpu(ctx)
    pipe = Pipeline(table, filter, projs, interpreter=interpreter, mode=mode)
    res = pipe.run()
    assert res.used_jit == (not interpreter), "generated code was expected to run (is NVRTC missing?)"
    n_gpu = res.num_rows
    assert n_gpu == n_ref, f"row count: gpu {n_gpu} oracle {n_ref}"
    for k in range(len(projs)):
        vals, valid = res.column(k)
        assert np.array_equal(valid, ref_valid[k]), f"validity of output {k} differs"
        dt, _ = res.column_type(k)
        got = vals.astype(np.float64)      # BOOLEAN -> 0/1, INT64/DATE32 exact
        ok = f64_equal(got, ref_vals[k]) | ~valid
        assert ok.all(), f"output {k}: first mismatch at row {int(np.argmin(ok))}: gpu {got[np.argmin(ok)]!r} " \
                         f"oracle {ref_vals[k][np.argmin(ok)]!r}"
    if filter is not None:
        sel = res.selection()
        assert np.array_equal(sel.astype(np.float64), ref_vals[-1]), "selection vector differs"
        assert (np.diff(sel.astype(np.int64)) > 0).all(), "selection vector not strictly ascending"
    res.close()
    if own_table:
        table.close()
    return n_gpu


def compare_aggregation(ctx: Context, spec: TableSpec, filter: Optional[Expression],
                        projections: Sequence[Expression], aggregates: Sequence[AggregationFunction],
                        interpreter: bool = False) -> List:
    A = AggregationFunction
    scan = list(range(len(spec.fields)))
    otable = spec.oracle()
    ref_row, = OraclePlan(otable, scan, filter=filter, projections=list(projections), aggregates=list(aggregates),
                          mode=Mode.BYTECODE_COMPILER).rows()
    # the terms each accumulator saw, for the SUM tolerance
    n_terms, term_vals, term_valid = OraclePlan(otable, scan, filter=filter, projections=list(projections),
                                                mode=Mode.BYTECODE_COMPILER).columns()
    table = spec.gpu(ctx)
    res = Pipeline(table, filter, list(projections), list(aggregates), interpreter=interpreter).run()
    assert res.num_rows == 1
    out = []
    for k, fn in enumerate(aggregates):
        g = res.agg(k)
        r = ref_row[k]
        out.append(g)
        if r is None:
            assert g is None, (fn, g)
            continue
        assert g is not None, (fn, r)
        terms = term_vals[k][term_valid[k]]
        if fn == A.COUNT:
            assert int(g) == int(r) == len(terms)
        elif fn in (A.MIN, A.MAX):
            assert f64_equal(np.array([g]), np.array([float(r)])).all(), (fn, g, r)
        elif fn == A.SUM:
            check_sum(g, terms, float(r))
        elif fn == A.AVG:
            if math.isfinite(float(r)) and len(terms):
                exact = math.fsum(terms.tolist()) / len(terms)
                tol = max(abs(float(r) - exact), 4 * ulp(exact))
                assert abs(g - exact) <= tol, (g, exact, r)
    res.close()
    table.close()
    return out
