"""Host-side mirror of the reference's operator interface for the GPU path.

The reference assembles ``Scan -> [Filter] -> [Projection] -> [GlobalAggregation]``
in ``buildPhysicalPlan`` (src/main/java/net/jhorstmann/queryengine/evaluator/Planner.kt:30-63)
out of ``Operator`` objects with ``open() / next() / close()``
(.../operator/Operators.kt:5-11).  ``GpuPipelineOperator`` is that whole subtree
as one operator: it runs the pipeline on the GPU through the C ABI
(include/qegpu.h) when opened and then hands rows out one at a time, exactly
like the operators it replaces; ``next_batch()`` exposes the columnar result
beside ``next()``.

Everything here is plumbing over libqegpu.so.  Nothing is evaluated on the
host; if the library or a GPU is missing the calls raise.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import List, Optional, Sequence

import numpy as np

from . import capi
from .ast import AggregationFunction, AggregationFunctionExpression, ColumnExpression, DataType, Expression, Mode
from .program import serialize

_NP_OF = {DataType.DOUBLE: np.float64, DataType.INT64: np.int64, DataType.DATE32: np.int32, DataType.STRING: np.int32}


class StringDictionary:
    """String <-> int32 code, one per Context.  STRING columns live in HBM as codes and string literals
    are serialized as codes of the SAME dictionary, so equal codes <=> equal strings -- which is all the
    reference lets a query do with strings (CMP_EQ / CMP_NE, IF branches, group keys, COUNT:
    .../evaluator/TypeCheck.kt:65-71,86-94,118-120).  Codes are handed out in first-seen order."""

    def __init__(self):
        self._code = {}
        self._strings: List[str] = []

    def __len__(self) -> int:
        return len(self._strings)

    def encode(self, s: str) -> int:
        c = self._code.get(s)
        if c is None:
            c = len(self._strings)
            self._code[s] = c
            self._strings.append(s)
        return c

    def encode_column(self, values: Sequence[Optional[str]]):
        """list of str / None -> (int32 codes, bool validity mask)"""
        codes = np.zeros(len(values), np.int32)
        valid = np.ones(len(values), bool)
        for i, s in enumerate(values):
            if s is None:
                valid[i] = False
            else:
                codes[i] = self.encode(s)
        return codes, valid

    def decode(self, code: int) -> str:
        return self._strings[int(code)]


def pack_bits(mask) -> np.ndarray:
    """bool array -> Arrow bitmap (LSB first)."""
    return np.packbits(np.asarray(mask, dtype=np.uint8), bitorder="little")


def unpack_bits(bitmap: np.ndarray, n: int) -> np.ndarray:
    return np.unpackbits(bitmap, bitorder="little")[:n].astype(bool)


class Context:
    """qe_ctx: one device, one stream, one memory pool."""

    def __init__(self, device: int = 0, _handle=None, dictionary: Optional[StringDictionary] = None):
        self.handle = C.c_void_p()
        self._owned = _handle is None
        if _handle is None:
            capi.check(capi.lib().qe_ctx_create(device, C.byref(self.handle)))
        else:
            self.handle = C.c_void_p(_handle)               # a context of a MultiGpu group (owned by the group)
        self.device = device
        self.dictionary = dictionary if dictionary is not None else StringDictionary()

    def set_option(self, key: str, value: int) -> None:
        capi.check(capi.lib().qe_ctx_set_option(self.handle, key.encode(), int(value)))

    def synchronize(self) -> None:
        capi.check(capi.lib().qe_ctx_synchronize(self.handle))

    @property
    def stream(self) -> int:
        return capi.lib().qe_ctx_stream(self.handle) or 0

    @property
    def launch_count(self) -> int:
        return capi.lib().qe_ctx_launch_count(self.handle)

    def trim(self) -> None:
        """Return the device memory the context's pool holds but does not use to the driver."""
        capi.check(capi.lib().qe_ctx_trim(self.handle))

    def close(self) -> None:
        if self.handle and self._owned:
            capi.lib().qe_ctx_destroy(self.handle)
        self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def jit_stats() -> dict:
    """Where compiled pipelines came from, process-wide (qe_jit_stats)."""
    st = (C.c_double * 6)()
    capi.check(capi.lib().qe_jit_stats(st))
    return {"memory": int(st[0]), "aot": int(st[1]), "disk": int(st[2]), "nvrtc": int(st[3]), "nvrtc_ms": float(st[4]),
            "aot_table_entries": int(st[5])}


class Program:
    """compileExpression(expr, mode) (.../evaluator/Compiler.kt:20-26): the flattened tree, held by the library.

    `mode` picks the SEMANTICS the reference's three evaluators disagree on, not an implementation -- everything runs on
    the GPU.  Mode.GPU, INTERPRETER and BYTECODE_COMPILER order doubles like Double.compare (Interpreter.kt:104-107,
    BytecodeCompiler.kt:300); Mode.CLOSURE_COMPILER compares them unboxed, IEEE (ClosureCompiler.kt:127-130):
    QE_CMP_IEEE on every ordering comparison of the tree."""

    def __init__(self, ctx: Context, expr: Expression, mode: Mode = Mode.GPU):
        self.ctx = ctx
        self.expr = expr
        self.mode = Mode(mode)
        self.words = serialize(expr, ctx.dictionary, ieee_order=self.mode == Mode.CLOSURE_COMPILER)
        self.handle = C.c_void_p()
        capi.check(capi.lib().qe_program_create(ctx.handle, self.words.ctypes.data, self.words.size, C.byref(self.handle)))

    def close(self) -> None:
        if self.handle and self.ctx.handle:      # a closed context has already released everything
            capi.lib().qe_program_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def compileExpression(ctx: Context, expr: Expression, mode: Mode = Mode.GPU) -> Program:
    """compileExpression(expr, mode), .../evaluator/Compiler.kt:20-26 (the context stands for the JVM)."""
    return Program(ctx, expr, mode)


@dataclass
class Field:
    """data class Field(name, type) (.../data/Schema.kt:7) plus nullability."""
    name: str
    type: DataType
    nullable: bool = False


class ColumnarTable:
    """Columnar counterpart of MemoryTable (.../data/MemoryTable.kt:7-19), resident in HBM."""

    def __init__(self, ctx: Context, fields: Sequence[Field], nrows: int, _handle=None):
        self.ctx = ctx
        self.fields = list(fields)
        self.nrows = int(nrows)
        self._owned = _handle is None
        if _handle is not None:                             # a partition of a ShardedTable (owned by it)
            self.handle = C.c_void_p(_handle)
            return
        descs = (capi.ColDesc * len(self.fields))(*[capi.ColDesc(int(f.type), 1 if f.nullable else 0) for f in self.fields])
        self.handle = C.c_void_p()
        capi.check(capi.lib().qe_table_create(ctx.handle, len(self.fields), descs, self.nrows, C.byref(self.handle)))

    @classmethod
    def from_arrays(cls, ctx: Context, fields: Sequence[Field], columns: Sequence, validity: Optional[Sequence] = None):
        n = len(columns[0]) if len(columns) else 0
        t = cls(ctx, fields, n)
        for c, f in enumerate(fields):
            v = None if validity is None else validity[c]
            t.upload(c, columns[c], v)
        return t

    def upload(self, col: int, values, valid_mask=None, row0: int = 0) -> None:
        """values: array of the column's type (bool array for BOOLEAN); valid_mask: bool array or None."""
        f = self.fields[col]
        if f.type == DataType.STRING and not isinstance(values, np.ndarray):
            values, smask = self.ctx.dictionary.encode_column(values)      # str / None -> codes
            if valid_mask is None and not smask.all():
                valid_mask = smask
            elif valid_mask is not None:
                valid_mask = np.asarray(valid_mask, dtype=bool) & smask
        if f.type == DataType.BOOLEAN:
            n = len(values)
            data = pack_bits(values)
        else:
            data = np.ascontiguousarray(values, dtype=_NP_OF[f.type])
            n = data.size
        vbits = None
        if valid_mask is not None:
            if f.nullable:
                vbits = pack_bits(valid_mask)
            elif not np.asarray(valid_mask).all():
                raise ValueError(f"column {f.name} is declared non-nullable but has nulls")
        capi.check(capi.lib().qe_table_upload(
            self.handle, col, data.ctypes.data if n else None,
            vbits.ctypes.data if vbits is not None and n else None, row0, n))
        # uploads are asynchronous on the context's stream and may read `data` (pinned) later
        self.ctx.synchronize()

    def upload_raw(self, col: int, values_ptr: int, validity_ptr: Optional[int], row0: int, n: int) -> None:
        """Asynchronous upload from caller-managed (ideally pinned) memory."""
        capi.check(capi.lib().qe_table_upload(self.handle, col, values_ptr, validity_ptr, row0, n))

    def generate(self, col: int, kind: int, seed: int, p0: float = 0.0, p1: float = 0.0,
                 null_rate: float = 0.0, row_offset: int = 0) -> None:
        capi.check(capi.lib().qe_table_generate(self.handle, col, kind, seed, p0, p1, null_rate, row_offset))

    def device_ptr(self, col: int, which: int = 0) -> int:
        return capi.lib().qe_table_device_ptr(self.handle, col, which) or 0

    def getScanOperator(self, projection: Sequence[str]) -> "GpuPipelineOperator":
        """Table.getScanOperator(projection) (.../data/Table.kt:8): a scan of the named columns."""
        idx = []
        for name in projection:
            hits = [i for i, f in enumerate(self.fields) if f.name == name]
            if not hits:
                raise ValueError(f"Unknown field {name}")          # IllegalArgumentException, MemoryTable.kt:12
            idx.append(hits[0])
        exprs = [ColumnExpression(self.fields[i].name, i, self.fields[i].type) for i in idx]
        return GpuPipelineOperator(self, None, exprs, None)

    def close(self) -> None:
        if self.handle and self.ctx.handle and self._owned:
            capi.lib().qe_table_destroy(self.handle)
        self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PipelineResult:
    """qe_result: device buffers of one pipeline run."""

    def __init__(self, ctx: Context, _handle=None):
        self.ctx = ctx
        self._owned = _handle is None
        self.handle = C.c_void_p(_handle) if _handle is not None else C.c_void_p()

    @property
    def num_rows(self) -> int:
        n = C.c_int64(0)
        capi.check(capi.lib().qe_result_num_rows(self.handle, C.byref(n)))
        return n.value

    @property
    def num_columns(self) -> int:
        n = C.c_int(0)
        capi.check(capi.lib().qe_result_num_columns(self.handle, C.byref(n)))
        return n.value

    def column_type(self, k: int):
        t, hv = C.c_int(0), C.c_int(0)
        capi.check(capi.lib().qe_result_column_type(self.handle, k, C.byref(t), C.byref(hv)))
        return DataType(t.value), bool(hv.value)

    def column(self, k: int):
        """(values, valid): numpy arrays of the column's type and a bool mask."""
        n = self.num_rows
        t, _ = self.column_type(k)
        vbits = np.zeros((n + 7) // 8 + 1, np.uint8)
        if t == DataType.BOOLEAN:
            data = np.zeros((n + 7) // 8 + 1, np.uint8)
        else:
            data = np.zeros(max(n, 1), _NP_OF[t])
        capi.check(capi.lib().qe_result_column(self.handle, k, data.ctypes.data, vbits.ctypes.data))
        valid = unpack_bits(vbits, n)
        if t == DataType.BOOLEAN:
            return unpack_bits(data, n), valid
        return data[:n], valid

    def selection(self) -> np.ndarray:
        n = self.num_rows
        out = np.zeros(max(n, 1), np.uint32)
        capi.check(capi.lib().qe_result_selection(self.handle, out.ctypes.data))
        return out[:n]

    def agg(self, k: int):
        v, isnull = C.c_double(0.0), C.c_int(0)
        capi.check(capi.lib().qe_result_agg(self.handle, k, C.byref(v), C.byref(isnull)))
        return None if isnull.value else v.value

    def agg_partial(self, k: int) -> capi.AggPartial:
        p = capi.AggPartial()
        capi.check(capi.lib().qe_result_agg_partial(self.handle, k, C.byref(p)))
        return p

    def agg_partials(self, k: int) -> np.ndarray:
        """The partials of aggregate k per super-tile of QE_AGG_SUPER_ROWS rows, in row order: float64[n, 4] =
        sum, count (int64 bit pattern), min, max.  Concatenated over partitions in order and folded by
        qe_agg_combine they give the same bits for any number of partitions (qe_result_agg_partials)."""
        n = C.c_int64(0)
        capi.check(capi.lib().qe_result_agg_partials(self.handle, k, None, 0, C.byref(n)))
        out = np.zeros((max(n.value, 1), 4), np.float64)
        if n.value:
            capi.check(capi.lib().qe_result_agg_partials(self.handle, k, out.ctypes.data_as(C.POINTER(capi.AggPartial)), n.value, C.byref(n)))
        return out[:n.value]

    @property
    def kernel_ms(self) -> float:
        ms = C.c_float(0.0)
        capi.check(capi.lib().qe_result_kernel_ms(self.handle, C.byref(ms)))
        return ms.value

    def kernel_stats(self):
        """(total device ms of the main kernel over all completed runs into this result, number of runs)"""
        ms, runs = C.c_double(0.0), C.c_int64(0)
        capi.check(capi.lib().qe_result_kernel_stats(self.handle, C.byref(ms), C.byref(runs)))
        return ms.value, runs.value

    @property
    def used_jit(self) -> bool:
        return bool(capi.lib().qe_result_used_jit(self.handle))

    def device_ptr(self, k: int, which: int) -> int:
        return capi.lib().qe_result_device_ptr(self.handle, k, which) or 0

    def close(self) -> None:
        if self.handle and self.ctx.handle and self._owned:
            capi.lib().qe_result_destroy(self.handle)
        self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def combine_aggregate(fn: AggregationFunction, partials: Sequence[capi.AggPartial]):
    """Fixed-order combination of per-partition partial aggregates (rank order)."""
    arr = (capi.AggPartial * max(1, len(partials)))(*partials)
    v, isnull = C.c_double(0.0), C.c_int(0)
    capi.check(capi.lib().qe_agg_combine(int(fn), arr, len(partials), C.byref(v), C.byref(isnull)))
    return None if isnull.value else v.value


class Pipeline:
    """A bound Scan -> [Filter] -> [Projection] -> [GlobalAggregation] subtree, runnable repeatedly."""

    def __init__(self, table: ColumnarTable, filter: Optional[Expression] = None,
                 projections: Optional[Sequence[Expression]] = None,
                 aggregates: Optional[Sequence[AggregationFunction]] = None,
                 interpreter: bool = False, selection: bool = True, out_capacity_rows: int = 0, values: bool = True,
                 mode: Mode = Mode.GPU):
        """values=False (QE_RUN_NO_VALUES): a filter's selection vector is its whole output -- for callers that hold
        the table themselves, as FilterOperator hands back the row it was given (.../operator/FilterOperator.kt:20-22)."""
        self.table = table
        self.ctx = table.ctx
        self.filter = Program(self.ctx, filter, mode) if filter is not None else None
        self.projs = [Program(self.ctx, e, mode) for e in (projections or [])]
        self.aggs = [int(a) for a in (aggregates or [])]
        self.flags = ((capi.QE_RUN_INTERPRETER if interpreter else 0) | (0 if selection else capi.QE_RUN_NO_SELECTION) |
                      (0 if values else capi.QE_RUN_NO_VALUES))
        self.out_capacity_rows = out_capacity_rows
        self._proj_arr = (C.c_void_p * max(1, len(self.projs)))(*[p.handle for p in self.projs])
        self._agg_arr = (C.c_int32 * max(1, len(self.aggs)))(*self.aggs)

    def desc(self, row_begin: int = 0, row_end: int = -1, append: bool = False) -> "capi.PipelineDesc":
        d = capi.PipelineDesc()
        d.filter = self.filter.handle if self.filter else None
        d.proj = self._proj_arr
        d.nproj = len(self.projs)
        d.agg_fns = self._agg_arr
        d.nagg = len(self.aggs)
        d.row_begin = row_begin
        d.row_end = row_end
        d.out_capacity_rows = self.out_capacity_rows
        d.flags = self.flags | (capi.QE_RUN_APPEND if append else 0)
        return d

    def run_append(self, result: PipelineResult, row_begin: int, row_end: int) -> PipelineResult:
        """Run one more row-range batch and append its rows to `result` (QE_RUN_APPEND)."""
        d = self.desc(row_begin, row_end, append=True)
        capi.check(capi.lib().qe_pipeline_run_ex(self.ctx.handle, self.table.handle, C.byref(d), C.byref(result.handle)))
        return result

    def run(self, result: Optional[PipelineResult] = None, row_begin: int = 0, row_end: int = -1) -> PipelineResult:
        """Enqueue the pipeline on the context's stream; accessors of the result wait for it."""
        r = result or PipelineResult(self.ctx)
        d = capi.PipelineDesc()
        d.filter = self.filter.handle if self.filter else None
        d.proj = self._proj_arr
        d.nproj = len(self.projs)
        d.agg_fns = self._agg_arr
        d.nagg = len(self.aggs)
        d.row_begin = row_begin
        d.row_end = row_end
        d.out_capacity_rows = self.out_capacity_rows
        d.flags = self.flags
        capi.check(capi.lib().qe_pipeline_run_ex(self.ctx.handle, self.table.handle, C.byref(d), C.byref(r.handle)))
        return r


def stream_pipelines(table: ColumnarTable, pipelines: Sequence[Pipeline], host_values: Sequence[Optional[int]],
                     host_validity: Optional[Sequence[Optional[int]]] = None, chunk_rows: int = 0,
                     results: Optional[List[Optional[PipelineResult]]] = None,
                     host_selection: Optional[Sequence[Optional[int]]] = None,
                     host_out_values: Optional[Sequence[Sequence[Optional[int]]]] = None,
                     host_out_validity: Optional[Sequence[Sequence[Optional[int]]]] = None,
                     total_rows: int = 0, chunk_relative: bool = False, gen: Optional[Sequence[Optional[tuple]]] = None,
                     gen_row_offset: int = 0, chunk_kept: Optional[List[np.ndarray]] = None, _desc_only: bool = False):
    """qe_pipeline_stream: host columns in (addresses of caller-owned, ideally pinned buffers), host results
    out, chunk by chunk with uploads, kernels and downloads overlapped.  Returns the (complete) results.

    total_rows > 0: bounded mode -- `table` is a window of >= 2 chunks, the host table has total_rows rows (which
    may exceed the device memory and 2^32).  chunk_relative: selection vectors count from their chunk's first row;
    chunk_kept[q] (int64 array of one entry per chunk) receives the rows kept up to and including each chunk.
    gen[col] = (kind, seed, p0, p1, null_rate) generates that column on the device instead of uploading it."""
    ctx = table.ctx
    np_ = len(pipelines)
    res = [(results[q] if results and results[q] is not None else PipelineResult(ctx)) for q in range(np_)]
    ncols = len(table.fields)

    def ptr_array(vals, n):
        arr = (C.c_void_p * max(1, n))()
        for i in range(n):
            arr[i] = vals[i] if vals is not None and i < len(vals) and vals[i] else None
        return arr

    sd = capi.StreamDesc()
    sd.chunk_rows = chunk_rows
    hv = ptr_array(host_values, ncols)
    sd.host_values = hv
    hn = ptr_array(host_validity, ncols) if host_validity is not None else None
    if hn is not None:
        sd.host_validity = hn
    sd.npipelines = np_
    descs = (capi.PipelineDesc * np_)(*[p.desc() for p in pipelines])
    sd.pipelines = descs
    handles = (C.c_void_p * np_)(*[r.handle.value for r in res])
    sd.results = handles
    hs = ptr_array(host_selection, np_) if host_selection is not None else None
    if hs is not None:
        sd.host_selection = hs
    keep = []

    def ptr_matrix(rows):
        outer = (C.POINTER(C.c_void_p) * np_)()
        for q in range(np_):
            n = len(pipelines[q].projs)
            inner = ptr_array(rows[q] if rows[q] is not None else None, n)
            keep.append(inner)
            outer[q] = C.cast(inner, C.POINTER(C.c_void_p))
        return outer

    if host_out_values is not None:
        ov = ptr_matrix(host_out_values)
        sd.host_out_values = ov
    if host_out_validity is not None:
        on = ptr_matrix(host_out_validity)
        sd.host_out_validity = on
    sd.total_rows = int(total_rows)
    sd.flags = capi.QE_STREAM_CHUNK_RELATIVE if chunk_relative else 0
    sd.gen_row_offset = int(gen_row_offset)
    if gen is not None:
        garr = (capi.StreamGen * ncols)()
        for c in range(ncols):
            g = gen[c] if c < len(gen) else None
            if g is None:
                garr[c].kind = -1
            else:
                garr[c].kind, garr[c].seed, garr[c].p0, garr[c].p1 = int(g[0]), int(g[1]), float(g[2]), float(g[3])
                garr[c].null_rate = float(g[4]) if len(g) > 4 else 0.0
        keep.append(garr)
        sd.gen = garr
    if chunk_kept is not None:
        ck = (C.POINTER(C.c_int64) * np_)()
        for q in range(np_):
            ck[q] = chunk_kept[q].ctypes.data_as(C.POINTER(C.c_int64)) if chunk_kept[q] is not None else None
        keep.append(ck)
        sd.chunk_kept = ck
    if _desc_only:                                            # MultiGpu.stream runs the descriptors of all partitions at once
        return sd, res, handles, (hv, hn, descs, hs, keep)
    try:
        capi.check(capi.lib().qe_pipeline_stream(ctx.handle, table.handle, C.byref(sd)))
    finally:
        for q in range(np_):
            res[q].handle = C.c_void_p(handles[q])
    return res


class GpuPipelineOperator:
    """Operator { open(); next(): Array<Any?>?; close() } for the whole GPU subtree.

    Stands where buildPhysicalPlan would put
    ``[GlobalAggregationOperator(] [ProjectionOperator(] [FilterOperator(] scan [)])])``.
    A filter without projections returns the scanned row unchanged, like
    FilterOperator.next (.../operator/FilterOperator.kt:14-25).
    """

    def __init__(self, table: ColumnarTable, filter: Optional[Expression],
                 projections: Optional[Sequence[Expression]],
                 aggregates: Optional[Sequence[AggregationFunction]], interpreter: bool = False,
                 scan_columns: Optional[Sequence[int]] = None, mode: Mode = Mode.GPU):
        if not projections:
            cols = scan_columns if scan_columns is not None else range(len(table.fields))
            projections = [ColumnExpression(table.fields[i].name, i, table.fields[i].type) for i in cols]
        self.pipeline = Pipeline(table, filter, projections, aggregates, interpreter=interpreter, mode=mode)
        self.aggregates = list(aggregates or [])
        self._result: Optional[PipelineResult] = None
        self._cols = None
        self._idx = 0
        self._n = 0

    def open(self) -> None:
        self._result = self.pipeline.run(self._result)
        self._n = self._result.num_rows
        self._idx = 0
        self._cols = None

    def next_batch(self):
        """Columnar view: list of (values, valid) per output column, or None when consumed."""
        if self._result is None:
            raise RuntimeError("Operator not opened")          # IllegalStateException in the reference
        if self._cols is None:
            if self.aggregates:
                vals = [self._result.agg(k) for k in range(len(self.aggregates))]
                self._cols = [(np.array([0.0 if v is None else v]), np.array([v is not None])) for v in vals]
                self._types = [DataType.DOUBLE] * len(vals)
            else:
                self._cols = [self._result.column(k) for k in range(self._result.num_columns)]
                self._types = [self._result.column_type(k)[0] for k in range(self._result.num_columns)]
        return self._cols

    def next(self):
        if self._result is None:
            raise RuntimeError("Operator not opened")
        if self._idx >= self._n:
            return None
        cols = self.next_batch()
        i = self._idx
        self._idx += 1
        row = []
        for k, (vals, valid) in enumerate(cols):
            if not valid[i]:
                row.append(None)
            elif vals.dtype == np.bool_:
                row.append(bool(vals[i]))
            elif self._types[k] == DataType.STRING:
                row.append(self.pipeline.ctx.dictionary.decode(vals[i]))
            elif self.aggregates and self.aggregates[k] == AggregationFunction.COUNT:
                row.append(int(vals[i]))                       # CountAccumulator.finish() is an Int
            else:
                row.append(float(vals[i]))
        return row

    def close(self) -> None:
        self._cols = None
        self._idx = self._n = 0

    def map(self, mapper=lambda r: r) -> List:
        """Operator.map (.../operator/Operators.kt:25-32): open, pull until null, close."""
        out = []
        self.open()
        try:
            while True:
                r = self.next()
                if r is None:
                    break
                out.append(mapper(r))
        finally:
            self.close()
        return out


class GroupByPipeline:
    """A bound Scan -> [Filter] -> Projection(keys + inputs) -> GroupByAggregation subtree (qe_groupby_run)."""

    def __init__(self, table: ColumnarTable, filter: Optional[Expression], keys: Sequence[Expression],
                 inputs: Sequence[Expression], aggregates: Sequence[AggregationFunction],
                 max_groups: int = 0, atomic: bool = False, shared: bool = False, twopass: bool = False,
                 mode: Mode = Mode.GPU):
        self.table = table
        self.ctx = table.ctx
        self.filter = Program(self.ctx, filter, mode) if filter is not None else None
        self.keys = [Program(self.ctx, e, mode) for e in keys]
        self.inputs = [Program(self.ctx, e, mode) for e in inputs]
        self.aggs = [int(a) for a in aggregates]
        if len(self.aggs) != len(self.inputs):
            raise ValueError("one aggregation function per input")
        self.max_groups = max_groups
        self.flags = (capi.QE_GROUPBY_ATOMIC if atomic else 0) | (capi.QE_GROUPBY_SHARED if shared else 0) | (capi.QE_GROUPBY_TWOPASS if twopass else 0)
        self._key_arr = (C.c_void_p * max(1, len(self.keys)))(*[p.handle for p in self.keys])
        self._in_arr = (C.c_void_p * max(1, len(self.inputs)))(*[p.handle for p in self.inputs])
        self._agg_arr = (C.c_int32 * max(1, len(self.aggs)))(*self.aggs)

    def run(self, result: Optional[PipelineResult] = None, row_begin: int = 0, row_end: int = -1) -> PipelineResult:
        r = result or PipelineResult(self.ctx)
        d = capi.GroupByDesc()
        d.filter = self.filter.handle if self.filter else None
        d.keys = self._key_arr
        d.nkeys = len(self.keys)
        d.inputs = self._in_arr
        d.agg_fns = self._agg_arr
        d.nagg = len(self.aggs)
        d.row_begin = row_begin
        d.row_end = row_end
        d.max_groups = self.max_groups
        d.flags = self.flags
        capi.check(capi.lib().qe_groupby_run(self.ctx.handle, self.table.handle, C.byref(d), C.byref(r.handle)))
        return r


def rewrite_aggregates(select: Sequence[Expression]):
    """rewriteAggregates for one projection (.../evaluator/RewriteAggregates.kt:18-49): every select
    expression without an aggregate is a group key (in select order), every aggregate's operand an
    accumulator input.  Returns (keys, inputs, functions, finish) where finish[i] is the column of the
    aggregation's output row (keys first, then aggregates) that select expression i reads.  Only bare
    aggregates are handled here; expressions OVER aggregates stay with the caller (they run on the
    handful of result rows)."""
    keys = [e for e in select if not isinstance(e, AggregationFunctionExpression)]
    inputs, fns, finish = [], [], []
    ki = 0
    for e in select:
        if isinstance(e, AggregationFunctionExpression):
            finish.append(len(keys) + len(inputs))
            inputs.append(e.operands[0])
            fns.append(e.function)
        else:
            finish.append(ki)
            ki += 1
    return keys, inputs, fns, finish


class GpuGroupByOperator:
    """Operator for ProjectionOperator(GroupByAggregationOperator(ProjectionOperator([FilterOperator(] scan [)]))):
    `select` mixes key expressions and AggregationFunctionExpressions like a SELECT list
    (.../operator/GroupByAggregationOperator.kt:21-75).  Rows come out in order of first appearance."""

    def __init__(self, table: ColumnarTable, filter: Optional[Expression], select: Sequence[Expression],
                 max_groups: int = 0, atomic: bool = False, shared: bool = False, twopass: bool = False,
                 mode: Mode = Mode.GPU):
        self.select = list(select)
        keys, inputs, fns, self.finish = rewrite_aggregates(self.select)
        self.fns = fns
        self.nkeys = len(keys)
        self.pipeline = GroupByPipeline(table, filter, keys, inputs, fns, max_groups=max_groups, atomic=atomic, shared=shared, twopass=twopass,
                                        mode=mode)
        self._result: Optional[PipelineResult] = None
        self._rows: Optional[List[list]] = None
        self._idx = 0

    def open(self) -> None:
        self._result = self.pipeline.run(self._result)
        res = self._result
        n = res.num_rows
        cols = [res.column(k) for k in range(res.num_columns)]
        types = [res.column_type(k)[0] for k in range(res.num_columns)]
        dictionary = self.pipeline.ctx.dictionary
        rows = []
        for i in range(n):
            full = []
            for k, (vals, valid) in enumerate(cols):
                if not valid[i]:
                    full.append(None)
                elif types[k] == DataType.BOOLEAN:
                    full.append(bool(vals[i]))
                elif types[k] == DataType.STRING:
                    full.append(dictionary.decode(vals[i]))
                elif k >= self.nkeys and self.fns[k - self.nkeys] == AggregationFunction.COUNT:
                    full.append(int(vals[i]))                  # CountAccumulator.finish() is an Int
                else:
                    full.append(float(vals[i]))
            rows.append([full[c] for c in self.finish])
        self._rows = rows
        self._idx = 0

    def next(self):
        if self._rows is None:
            raise RuntimeError("Operator not opened")
        if self._idx >= len(self._rows):
            return None
        self._idx += 1
        return self._rows[self._idx - 1]

    def close(self) -> None:
        self._rows = None
        self._idx = 0

    def map(self, mapper=lambda r: r) -> List:
        out = []
        self.open()
        try:
            while True:
                r = self.next()
                if r is None:
                    break
                out.append(mapper(r))
        finally:
            self.close()
        return out
