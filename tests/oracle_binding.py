"""ctypes binding of the CPU oracle (oracle/libqe_oracle.so).  TEST INFRASTRUCTURE ONLY.

The product package (queryengine_b200/) never imports this module.  bench.py
imports it for the `cpu_baseline` / `--impl reference` legs, and
__graft_entry__.smoke() for its check.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import List, Optional, Sequence

import numpy as np

from queryengine_b200.ast import (
    AggregationFunction, BooleanLiteralExpression, ColumnExpression, DataType, Expression,
    FunctionExpression, Mode, NumericLiteralExpression, StringLiteralExpression,
)

_ORACLE_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle")
_LIB_PATH = os.path.join(_ORACLE_DIR, "libqe_oracle.so")


def build_oracle() -> str:
    subprocess.run(["make", "-s", "-C", _ORACLE_DIR], check=True)
    return _LIB_PATH


class _Value(C.Structure):
    class _U(C.Union):
        _fields_ = [("d", C.c_double), ("b", C.c_int), ("s", C.c_char_p), ("i", C.c_int32)]
    _fields_ = [("tag", C.c_int), ("u", _U)]


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH) or \
            os.path.getmtime(_LIB_PATH) < os.path.getmtime(os.path.join(_ORACLE_DIR, "qe_oracle.c")):
        build_oracle()
    L = C.CDLL(_LIB_PATH)
    vp = C.c_void_p
    L.orc_column.restype = vp; L.orc_column.argtypes = [C.c_int, C.c_int]
    L.orc_numeric.restype = vp; L.orc_numeric.argtypes = [C.c_double]
    L.orc_boolean.restype = vp; L.orc_boolean.argtypes = [C.c_int]
    L.orc_string.restype = vp; L.orc_string.argtypes = [C.c_char_p]
    L.orc_function.restype = vp; L.orc_function.argtypes = [C.c_int, C.c_int, C.POINTER(vp)]
    L.orc_expr_free.argtypes = [vp]
    L.orc_eval.restype = C.POINTER(_Value)
    L.orc_eval.argtypes = [vp, C.POINTER(C.POINTER(_Value)), C.c_int, C.POINTER(C.c_int)]
    L.orc_table_create.restype = vp; L.orc_table_create.argtypes = [C.c_int, C.POINTER(C.c_int), C.c_int64]
    L.orc_table_set_double.argtypes = [vp, C.c_int, vp, vp]
    L.orc_table_set_boolean.argtypes = [vp, C.c_int, vp, vp]
    L.orc_table_set_string.argtypes = [vp, C.c_int, C.POINTER(C.c_char_p)]
    L.orc_table_free.argtypes = [vp]
    L.orc_memory_source.restype = vp; L.orc_memory_source.argtypes = [vp, C.POINTER(C.c_int), C.c_int]
    L.orc_memory_source_range.restype = vp
    L.orc_memory_source_range.argtypes = [vp, C.POINTER(C.c_int), C.c_int, C.c_int64, C.c_int64]
    L.orc_filter.restype = vp; L.orc_filter.argtypes = [vp, vp, C.c_int]
    L.orc_projection.restype = vp; L.orc_projection.argtypes = [vp, C.POINTER(vp), C.c_int, C.c_int]
    L.orc_global_aggregation.restype = vp; L.orc_global_aggregation.argtypes = [vp, C.POINTER(C.c_int), C.c_int]
    L.orc_groupby_aggregation.restype = vp
    L.orc_groupby_aggregation.argtypes = [vp, C.c_int, C.POINTER(C.c_int), C.c_int]
    L.orc_open.argtypes = [vp]
    L.orc_next.restype = C.POINTER(C.POINTER(_Value)); L.orc_next.argtypes = [vp]
    L.orc_close.argtypes = [vp]
    L.orc_width.restype = C.c_int; L.orc_width.argtypes = [vp]
    L.orc_error.restype = C.c_int; L.orc_error.argtypes = [vp]
    L.orc_op_free.argtypes = [vp]
    L.orc_drain.restype = C.c_int64
    L.orc_drain.argtypes = [vp, C.c_int, C.POINTER(vp), C.POINTER(vp), C.c_int64]
    L.orc_drain_count.restype = C.c_int64; L.orc_drain_count.argtypes = [vp, C.POINTER(C.c_double)]
    L.orc_timed_pipeline.restype = C.c_double
    L.orc_timed_pipeline.argtypes = [vp, C.POINTER(C.c_int), C.c_int, vp, C.POINTER(vp), C.c_int,
                                     C.POINTER(C.c_int), C.c_int, C.c_int, C.c_int,
                                     C.POINTER(C.c_double), C.POINTER(C.c_int), C.POINTER(C.c_int64)]
    _lib = L
    return L


def _oracle_dtype(dt: DataType) -> int:
    # INT64 / DATE32 have no reference counterpart: the oracle sees their exact f64 widening
    return int(DataType.DOUBLE) if dt in (DataType.INT64, DataType.DATE32) else int(dt)


def to_oracle_expr(e: Expression):
    """Build the oracle's tree from the host AST.  Caller owns the result (OracleExpr)."""
    L = lib()
    if isinstance(e, ColumnExpression):
        return L.orc_column(e.index, _oracle_dtype(e.dataType))
    if isinstance(e, NumericLiteralExpression):
        return L.orc_numeric(e.value)
    if isinstance(e, BooleanLiteralExpression):
        return L.orc_boolean(1 if e.value else 0)
    if isinstance(e, StringLiteralExpression):
        return L.orc_string(e.value.encode())
    if isinstance(e, FunctionExpression):
        ops = (C.c_void_p * len(e.operands))(*[to_oracle_expr(o) for o in e.operands])
        return L.orc_function(int(e.function), len(e.operands), ops)
    raise TypeError(type(e))


class OracleExpr:
    def __init__(self, e: Expression):
        self.handle = to_oracle_expr(e)

    def __del__(self):
        if getattr(self, "handle", None):
            lib().orc_expr_free(self.handle)
            self.handle = None


def _unbox(p):
    if not p:
        return None
    v = p.contents
    if v.tag == 1:
        return v.u.d
    if v.tag == 2:
        return bool(v.u.b)
    if v.tag == 0:
        return v.u.s.decode()
    if v.tag == 3:
        return int(v.u.i)
    raise ValueError(v.tag)


def _box(x, keep: list):
    if x is None:
        return C.POINTER(_Value)()
    v = _Value()
    if isinstance(x, bool):
        v.tag = 2; v.u.b = 1 if x else 0
    elif isinstance(x, str):
        b = x.encode(); keep.append(b)
        v.tag = 0; v.u.s = b
    else:
        v.tag = 1; v.u.d = float(x)
    keep.append(v)
    return C.pointer(v)


class EvalError(Exception):
    """The reference would have thrown for this input."""


def oracle_mode(mode: Mode) -> int:
    # Mode.GPU is graded against the BYTECODE_COMPILER semantics
    return {Mode.INTERPRETER: 0, Mode.CLOSURE_COMPILER: 1, Mode.BYTECODE_COMPILER: 2, Mode.GPU: 2}[Mode(mode)]


def compile_expression(expr: Expression, mode: Mode = Mode.BYTECODE_COMPILER):
    """compileExpression(expr, mode) -> callable(row) (M/evaluator/Compiler.kt:20-26)."""
    oe = OracleExpr(expr)
    m = oracle_mode(mode)

    def call(row: Sequence):
        keep: list = []
        arr = (C.POINTER(_Value) * max(1, len(row)))(*[_box(x, keep) for x in row])
        err = C.c_int(0)
        r = lib().orc_eval(oe.handle, arr, m, C.byref(err))
        if err.value:
            raise EvalError()
        return _unbox(r)
    call._keepalive = oe
    return call


class OracleTable:
    """MemoryTable: columns given as numpy arrays (+ optional Arrow-style validity bitmaps)."""

    def __init__(self, dtypes: Sequence[DataType], nrows: int):
        self.dtypes = [DataType(d) for d in dtypes]
        self.nrows = int(nrows)
        arr = (C.c_int * len(dtypes))(*[_oracle_dtype(d) for d in self.dtypes])
        self.handle = lib().orc_table_create(len(dtypes), arr, self.nrows)

    @staticmethod
    def _validity_ptr(validity: Optional[np.ndarray], n: int):
        if validity is None:
            return None, None
        v = np.ascontiguousarray(validity, dtype=np.uint8)
        assert v.size * 8 >= n
        return v, v.ctypes.data

    def set_column(self, col: int, values, validity: Optional[np.ndarray] = None):
        dt = self.dtypes[col]
        keep, vptr = self._validity_ptr(validity, self.nrows)
        if dt in (DataType.DOUBLE, DataType.INT64, DataType.DATE32):
            a = np.ascontiguousarray(np.asarray(values).astype(np.float64))
            assert a.size == self.nrows
            rc = lib().orc_table_set_double(self.handle, col, a.ctypes.data, vptr)
        elif dt == DataType.BOOLEAN:
            a = np.ascontiguousarray(np.asarray(values).astype(np.uint8))
            assert a.size == self.nrows
            rc = lib().orc_table_set_boolean(self.handle, col, a.ctypes.data, vptr)
        else:
            strs = (C.c_char_p * max(1, self.nrows))(*[None if s is None else s.encode() for s in values])
            rc = lib().orc_table_set_string(self.handle, col, strs)
        assert rc == 0

    @classmethod
    def from_rows(cls, dtypes: Sequence[DataType], rows: Sequence[Sequence]):
        """MemoryTable(schema, listOf(listOf(...), ...)) with None for null."""
        t = cls(dtypes, len(rows))
        n = len(rows)
        for c, dt in enumerate(t.dtypes):
            colv = [r[c] for r in rows]
            valid = np.packbits(np.array([x is not None for x in colv], dtype=np.uint8), bitorder="little") \
                if n else np.zeros(0, np.uint8)
            if dt == DataType.STRING:
                t.set_column(c, colv)
            elif dt == DataType.BOOLEAN:
                t.set_column(c, np.array([bool(x) if x is not None else False for x in colv], dtype=np.uint8), valid)
            else:
                t.set_column(c, np.array([x if x is not None else 0.0 for x in colv], dtype=np.float64), valid)
        return t

    def __del__(self):
        if getattr(self, "handle", None):
            lib().orc_table_free(self.handle)
            self.handle = None


class OraclePlan:
    """source -> [filter] -> [projection] -> [global | group-by aggregation], as
    buildPhysicalPlan assembles it (M/evaluator/Planner.kt:30-63)."""

    def __init__(self, table: OracleTable, scan_columns: Sequence[int],
                 filter: Optional[Expression] = None,
                 projections: Optional[Sequence[Expression]] = None,
                 aggregates: Optional[Sequence[AggregationFunction]] = None,
                 group_count: int = 0,
                 mode: Mode = Mode.BYTECODE_COMPILER):
        L = lib()
        self.table = table
        self._keep: List[OracleExpr] = []
        m = oracle_mode(mode)
        idx = (C.c_int * max(1, len(scan_columns)))(*scan_columns)
        op = L.orc_memory_source(table.handle, idx, len(scan_columns))
        if filter is not None:
            fe = OracleExpr(filter); self._keep.append(fe)
            op = L.orc_filter(op, fe.handle, m)
        if projections:
            pes = [OracleExpr(p) for p in projections]; self._keep += pes
            arr = (C.c_void_p * len(pes))(*[p.handle for p in pes])
            op = L.orc_projection(op, arr, len(pes), m)
        if aggregates:
            fns = (C.c_int * len(aggregates))(*[int(a) for a in aggregates])
            if group_count > 0:
                op = L.orc_groupby_aggregation(op, group_count, fns, len(aggregates))
            else:
                op = L.orc_global_aggregation(op, fns, len(aggregates))
        self.op = op

    def rows(self) -> List[list]:
        """Operator.map { it.toList() } (M/operator/Operators.kt:25-32)."""
        L = lib()
        out = []
        w = L.orc_width(self.op)
        L.orc_open(self.op)
        while True:
            r = L.orc_next(self.op)
            if not r:
                break
            out.append([_unbox(r[c]) for c in range(w)])
        L.orc_close(self.op)
        if L.orc_error(self.op):
            raise EvalError()
        return out

    def columns(self, capacity: Optional[int] = None):
        """Drain into (values[c] float64 array, valid[c] bool array) per output column."""
        L = lib()
        w = L.orc_width(self.op)
        cap = int(capacity if capacity is not None else max(1, self.table.nrows))
        vals = [np.zeros(cap, np.float64) for _ in range(w)]
        valid = [np.zeros(cap, np.uint8) for _ in range(w)]
        vp = (C.c_void_p * max(1, w))(*[v.ctypes.data for v in vals])
        bp = (C.c_void_p * max(1, w))(*[v.ctypes.data for v in valid])
        n = L.orc_drain(self.op, w, vp, bp, cap)
        if n < 0:
            raise EvalError()
        return n, [v[:n] for v in vals], [v[:n].astype(bool) for v in valid]

    def __del__(self):
        if getattr(self, "op", None):
            lib().orc_op_free(self.op)
            self.op = None


def timed_pipeline(table: OracleTable, scan_columns: Sequence[int], filter: Optional[Expression],
                   projections: Sequence[Expression], aggregates: Sequence[AggregationFunction],
                   threads: int, mode: Mode = Mode.BYTECODE_COMPILER):
    """Run the reference-shaped pull chain on `threads` host threads; returns
    (seconds, rows_at_sink, [(value, is_null), ...])."""
    L = lib()
    keep = []
    fe = None
    if filter is not None:
        fe = OracleExpr(filter); keep.append(fe)
    pes = [OracleExpr(p) for p in projections]; keep += pes
    parr = (C.c_void_p * max(1, len(pes)))(*[p.handle for p in pes])
    idx = (C.c_int * max(1, len(scan_columns)))(*scan_columns)
    fns = (C.c_int * max(1, len(aggregates)))(*[int(a) for a in aggregates])
    av = (C.c_double * max(1, len(aggregates)))()
    an = (C.c_int * max(1, len(aggregates)))()
    rows = C.c_int64(0)
    secs = L.orc_timed_pipeline(table.handle, idx, len(scan_columns), fe.handle if fe else None,
                                parr, len(pes), fns, len(aggregates), oracle_mode(mode), threads,
                                av, an, C.byref(rows))
    return secs, rows.value, [(av[i], bool(an[i])) for i in range(len(aggregates))]
