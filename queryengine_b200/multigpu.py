"""Several GPUs driven by ONE process, through the C ABI (qe_mgpu_* in include/qegpu.h).

The reference is one process that builds one plan (src/main/java/net/jhorstmann/queryengine/Main.kt:11-19,
.../evaluator/Planner.kt:30), so this -- not torch.distributed -- is what a JVM adapter calls: a `MultiGpu` owns one
context and one worker thread per device, a `ShardedTable` is split into contiguous row ranges (boundaries are multiples
of QE_AGG_SUPER_ROWS, so SUM / AVG do not depend on the number of devices), every call fans out to all devices and
joins.  Filter / projection outputs stay on their device; concatenating the partitions in order is the reference's scan
order (.../operator/FilterOperator.kt:17-22).  Partial aggregates are folded on the calling thread in partition order,
group-by partitions are merged by key (qe_groupby_merge).

Plumbing over libqegpu.so only; nothing is evaluated here.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import numpy as np

from . import capi
from .ast import AggregationFunction, DataType, Expression, Mode
from .engine import (_NP_OF, ColumnarTable, Context, Field, GroupByPipeline, Pipeline, PipelineResult, StringDictionary,
                     pack_bits, stream_pipelines, unpack_bits)


def partition_rows(nrows: int, nparts: int, part: int):
    """Rows [begin, end) of partition `part` (qe_partition_rows: ceil(nrows / nparts) rounded up to QE_AGG_SUPER_ROWS)."""
    b, e = C.c_int64(0), C.c_int64(0)
    capi.check(capi.lib().qe_partition_rows(nrows, nparts, part, C.byref(b), C.byref(e)))
    return b.value, e.value


class MultiGpu:
    """qe_mgpu: one context + one worker thread per device.  All contexts share ONE string dictionary."""

    def __init__(self, devices: Sequence[int]):
        self.devices = list(devices)
        arr = (C.c_int * len(self.devices))(*self.devices)
        self.handle = C.c_void_p()
        capi.check(capi.lib().qe_mgpu_create(arr, len(self.devices), C.byref(self.handle)))
        self.dictionary = StringDictionary()
        self.contexts = [Context(d, _handle=capi.lib().qe_mgpu_ctx(self.handle, i), dictionary=self.dictionary)
                         for i, d in enumerate(self.devices)]

    def __len__(self) -> int:
        return len(self.devices)

    def set_option(self, key: str, value: int) -> None:
        capi.check(capi.lib().qe_mgpu_set_option(self.handle, key.encode(), int(value)))

    def synchronize(self) -> None:
        for c in self.contexts:
            c.synchronize()

    @property
    def launch_count(self) -> int:
        return sum(c.launch_count for c in self.contexts)

    def close(self) -> None:
        if self.handle:
            for c in self.contexts:
                c.handle = C.c_void_p()
            capi.lib().qe_mgpu_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class ShardedTable:
    """qe_mtable: a table partitioned by contiguous row ranges over the devices of a MultiGpu."""

    def __init__(self, group: MultiGpu, fields: Sequence[Field], nrows: int):
        self.group = group
        self.fields = list(fields)
        self.nrows = int(nrows)
        descs = (capi.ColDesc * len(self.fields))(*[capi.ColDesc(int(f.type), 1 if f.nullable else 0) for f in self.fields])
        self.handle = C.c_void_p()
        capi.check(capi.lib().qe_mtable_create(group.handle, len(self.fields), descs, self.nrows, C.byref(self.handle)))
        self.ranges = [partition_rows(self.nrows, len(group), p) for p in range(len(group))]
        self.parts = [ColumnarTable(group.contexts[p], self.fields, e - b, _handle=capi.lib().qe_mtable_part(self.handle, p))
                      for p, (b, e) in enumerate(self.ranges)]

    def upload(self, col: int, values, valid_mask=None) -> None:
        f = self.fields[col]
        if f.type == DataType.STRING and not isinstance(values, np.ndarray):
            values, smask = self.group.dictionary.encode_column(values)
            if valid_mask is None and not smask.all():
                valid_mask = smask
            elif valid_mask is not None:
                valid_mask = np.asarray(valid_mask, dtype=bool) & smask
        data = pack_bits(values) if f.type == DataType.BOOLEAN else np.ascontiguousarray(values, dtype=_NP_OF[f.type])
        vbits = pack_bits(valid_mask) if (valid_mask is not None and f.nullable) else None
        capi.check(capi.lib().qe_mtable_upload(self.handle, col, data.ctypes.data if data.size else None,
                                               vbits.ctypes.data if vbits is not None else None))

    def generate(self, col: int, kind: int, seed: int, p0: float = 0.0, p1: float = 0.0, null_rate: float = 0.0) -> None:
        capi.check(capi.lib().qe_mtable_generate(self.handle, col, kind, seed, p0, p1, null_rate))

    def close(self) -> None:
        if self.handle and self.group.handle:
            for t in self.parts:
                t.handle = C.c_void_p()
            capi.lib().qe_mtable_destroy(self.handle)
        self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class ShardedResult:
    """qe_mresult: one result per partition and the combined view."""

    def __init__(self, group: MultiGpu):
        self.group = group
        self.handle = C.c_void_p()

    @property
    def num_rows(self) -> int:
        n = C.c_int64(0)
        capi.check(capi.lib().qe_mresult_num_rows(self.handle, C.byref(n)))
        return n.value

    @property
    def num_columns(self) -> int:
        n = C.c_int(0)
        capi.check(capi.lib().qe_mresult_num_columns(self.handle, C.byref(n)))
        return n.value

    def part(self, p: int) -> PipelineResult:
        return PipelineResult(self.group.contexts[p], _handle=capi.lib().qe_mresult_part(self.handle, p))

    def part_rows(self, p: int) -> int:
        n = C.c_int64(0)
        capi.check(capi.lib().qe_mresult_part_rows(self.handle, p, C.byref(n)))
        return n.value

    def column(self, k: int, dtype: DataType):
        """(values, valid) of output column k over all partitions in order."""
        n = self.num_rows
        vbits = np.zeros((n + 7) // 8 + 1, np.uint8)
        if dtype == DataType.BOOLEAN:
            data = np.zeros((n + 7) // 8 + 1, np.uint8)
        else:
            data = np.zeros(max(n, 1), _NP_OF[dtype])
        capi.check(capi.lib().qe_mresult_column(self.handle, k, data.ctypes.data, vbits.ctypes.data))
        valid = unpack_bits(vbits, n)
        if dtype == DataType.BOOLEAN:
            return unpack_bits(data, n), valid
        return data[:n], valid

    def selection(self) -> np.ndarray:
        n = self.num_rows
        out = np.zeros(max(n, 1), np.uint64)
        capi.check(capi.lib().qe_mresult_selection(self.handle, out.ctypes.data))
        return out[:n]

    def agg(self, k: int):
        v, isnull = C.c_double(0.0), C.c_int(0)
        capi.check(capi.lib().qe_mresult_agg(self.handle, k, C.byref(v), C.byref(isnull)))
        return None if isnull.value else v.value

    def close(self) -> None:
        if self.handle and self.group.handle:
            capi.lib().qe_mresult_destroy(self.handle)
        self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class ShardedPipeline:
    """Scan -> [Filter] -> [Projection] -> [GlobalAggregation] over a ShardedTable (qe_mgpu_pipeline_run)."""

    def __init__(self, table: ShardedTable, filter: Optional[Expression] = None, projections: Optional[Sequence[Expression]] = None,
                 aggregates: Optional[Sequence[AggregationFunction]] = None, out_capacity_rows: int = 0, values: bool = True,
                 interpreter: bool = False, mode: Mode = Mode.GPU):
        self.table = table
        self.group = table.group
        # programs are host objects: created with the first context, run on all of them
        self.pipe = Pipeline(table.parts[0], filter, projections, aggregates, interpreter=interpreter, out_capacity_rows=out_capacity_rows,
                             values=values, mode=mode)

    def run(self, result: Optional[ShardedResult] = None) -> ShardedResult:
        r = result or ShardedResult(self.group)
        d = self.pipe.desc()
        capi.check(capi.lib().qe_mgpu_pipeline_run(self.group.handle, self.table.handle, C.byref(d), C.byref(r.handle)))
        return r

    def stream(self, host_values: Sequence[Optional[int]], widths: Sequence[int], chunk_rows: int, host_selection: Optional[int],
               host_out_values: Optional[Sequence[int]], out_offsets: Sequence[int],
               results: Optional[List[Optional[PipelineResult]]] = None) -> List[PipelineResult]:
        """Every device streams its row range of the host columns at the same time (qe_mgpu_pipeline_stream, resident
        mode).  host_values[c]: address of the whole host column c (widths[c] bytes per row); partition p's selection
        vector and 8-byte output columns land at row out_offsets[p] of host_selection / host_out_values[k] (the caller
        reserves every partition its capacity).  Returns the per-partition results."""
        n = len(self.group)
        keep, res_all, handles_all = [], [], []
        descs = (capi.StreamDesc * n)()
        for p in range(n):
            b, e = self.table.ranges[p]
            hv = [None if a is None else a + b * w for a, w in zip(host_values, widths)]
            pipe = Pipeline.__new__(Pipeline)
            pipe.__dict__.update(self.pipe.__dict__)
            pipe.table = self.table.parts[p]
            pipe.ctx = self.group.contexts[p]
            sd, res, handles, k = stream_pipelines(
                self.table.parts[p], [pipe], hv, chunk_rows=chunk_rows,
                results=[results[p]] if results and results[p] is not None else None,
                host_selection=[host_selection + 4 * out_offsets[p]] if host_selection else None,
                host_out_values=[[a + 8 * out_offsets[p] for a in host_out_values]] if host_out_values else None,
                _desc_only=True)
            descs[p] = sd
            keep.append((sd, k))
            res_all.append(res[0])
            handles_all.append(handles)
        try:
            capi.check(capi.lib().qe_mgpu_pipeline_stream(self.group.handle, self.table.handle, descs))
        finally:
            for p in range(n):
                res_all[p].handle = C.c_void_p(handles_all[p][0])
        return res_all


class ShardedGroupBy:
    """Scan -> [Filter] -> Projection -> GroupByAggregation over a ShardedTable (qe_mgpu_groupby_run)."""

    def __init__(self, table: ShardedTable, filter: Optional[Expression], keys: Sequence[Expression], inputs: Sequence[Expression],
                 aggregates: Sequence[AggregationFunction], max_groups: int = 0, mode: Mode = Mode.GPU):
        self.table = table
        self.group = table.group
        self.pipe = GroupByPipeline(table.parts[0], filter, keys, inputs, aggregates, max_groups=max_groups, mode=mode)
        self.nkeys = len(keys)
        self.fns = [AggregationFunction(a) for a in aggregates]

    def run(self, result: Optional[ShardedResult] = None) -> ShardedResult:
        r = result or ShardedResult(self.group)
        p = self.pipe
        d = capi.GroupByDesc()
        d.filter = p.filter.handle if p.filter else None
        d.keys, d.nkeys = p._key_arr, len(p.keys)
        d.inputs, d.agg_fns, d.nagg = p._in_arr, p._agg_arr, len(p.aggs)
        d.row_begin, d.row_end, d.max_groups, d.flags = 0, -1, p.max_groups, p.flags
        capi.check(capi.lib().qe_mgpu_groupby_run(self.group.handle, self.table.handle, C.byref(d), C.byref(r.handle)))
        return r

    def rows(self, result: ShardedResult, key_types: Sequence[DataType]) -> List[list]:
        """The merged groups as reference-shaped rows [key..., finish()...]."""
        n = result.num_rows
        cols = [result.column(k, key_types[k]) for k in range(self.nkeys)]
        cols += [result.column(self.nkeys + j, DataType.DOUBLE) for j in range(len(self.fns))]
        out = []
        for i in range(n):
            row = []
            for k, (vals, valid) in enumerate(cols):
                if not valid[i]:
                    row.append(None)
                elif k < self.nkeys and key_types[k] == DataType.STRING:
                    row.append(self.group.dictionary.decode(vals[i]))
                elif k < self.nkeys and key_types[k] == DataType.BOOLEAN:
                    row.append(bool(vals[i]))
                elif k >= self.nkeys and self.fns[k - self.nkeys] == AggregationFunction.COUNT:
                    row.append(int(vals[i]))
                else:
                    row.append(float(vals[i]))
            out.append(row)
        return out
