"""Row-range sharding of the filter/scan path across the GPUs of one box.

Every row is evaluated on its own (``RowCallable.invoke(row)`` reads one row,
src/main/java/net/jhorstmann/queryengine/evaluator/Compiler.kt:9-11), so the
path shards by contiguous row range with no data-path exchange:

* rank r owns rows ``[row_begin(r), row_end(r))`` of the logical table; range
  starts are multiples of 128 (the alignment qe_pipeline_run_ex wants);
* filter / projection outputs stay on their GPU.  Concatenating them in rank
  order reproduces the reference's scan order (FilterOperator.next preserves
  order, .../operator/FilterOperator.kt:17-22); a partition's selection vector
  is partition-relative, ``global_selection`` adds the partition's first row;
* the only exchange is the K partial aggregates (sum, count, min, max) of a
  trailing GlobalAggregationOperator (.../operator/GlobalAggregationOperator.kt:10-36):
  all-gathered (NCCL over NVLink on GPUs, gloo in the CPU tests) and combined
  in RANK ORDER by qe_agg_combine, so the result is bit-reproducible whatever
  reduction algorithm the collective library would have picked.

One process per GPU, torch.distributed for the plumbing.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np

from . import capi
from .ast import AggregationFunction

ALIGN = 128                      # what qe_pipeline_run_ex needs of a batch start
ALIGN_SUM = capi.QE_AGG_SUPER_ROWS   # partition boundaries at multiples of this make SUM / AVG independent of `world`


def partition(nrows: int, world: int, rank: int, align: int = ALIGN) -> Tuple[int, int]:
    """Rows [begin, end) of rank `rank`: ceil(nrows / world) rounded up to a multiple of `align` per rank.
    align=ALIGN_SUM is qe_partition_rows: the super-tiles of the fixed-shape aggregation then never straddle two
    partitions and a SUM has the same bits for any `world`."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError("bad rank / world size")
    per = -(-nrows // world)
    per = -(-per // align) * align
    begin = min(nrows, rank * per)
    end = min(nrows, begin + per)
    return begin, end


def partitions(nrows: int, world: int, align: int = ALIGN) -> List[Tuple[int, int]]:
    return [partition(nrows, world, r, align) for r in range(world)]


def global_selection(local_selection: np.ndarray, row_begin: int) -> np.ndarray:
    """Partition-relative selection vector (uint32) -> global row ids (int64)."""
    return local_selection.astype(np.int64) + np.int64(row_begin)


def _partials_to_array(partials: Sequence[capi.AggPartial]) -> np.ndarray:
    """K partials -> float64[K, 4]; the int64 count travels as its bit pattern."""
    a = np.zeros((len(partials), 4), dtype=np.float64)
    for k, p in enumerate(partials):
        a[k, 0] = p.sum
        a[k, 1] = np.array([p.count], dtype=np.int64).view(np.float64)[0]
        a[k, 2] = p.min
        a[k, 3] = p.max
    return a


def _array_to_partials(a: np.ndarray) -> List[capi.AggPartial]:
    out = []
    for k in range(a.shape[0]):
        count = int(np.array([a[k, 1]], dtype=np.float64).view(np.int64)[0])
        out.append(capi.AggPartial(float(a[k, 0]), count, float(a[k, 2]), float(a[k, 3])))
    return out


def all_gather_partials(partials: Sequence[capi.AggPartial], device=None, group=None) -> List[List[capi.AggPartial]]:
    """All-gather this rank's K partial aggregates; returns [rank][k].  Without an initialised process
    group (single process) the result is just [partials]."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return [list(partials)]
    world = dist.get_world_size(group)
    mine = torch.from_numpy(_partials_to_array(partials))
    if device is not None:
        mine = mine.to(device)
    gathered = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(gathered, mine, group=group)
    return [_array_to_partials(g.cpu().numpy()) for g in gathered]


def combine(fns: Sequence[AggregationFunction], by_rank: Sequence[Sequence[capi.AggPartial]]) -> List[Optional[float]]:
    """finish() of every aggregate over all partitions, partitions folded in rank order (qe_agg_combine)."""
    import ctypes as C
    out: List[Optional[float]] = []
    for k, fn in enumerate(fns):
        arr = (capi.AggPartial * max(1, len(by_rank)))(*[r[k] for r in by_rank])
        v, isnull = C.c_double(0.0), C.c_int(0)
        capi.check(capi.lib().qe_agg_combine(int(fn), arr, len(by_rank), C.byref(v), C.byref(isnull)))
        out.append(None if isnull.value else v.value)
    return out


def all_gather_partial_lists(mine: np.ndarray, device=None, group=None) -> np.ndarray:
    """float64[n_local, K, 4] super-tile partials of this rank -> float64[n_total, K, 4], the ranks' lists in rank order
    (= row order).  Lists have different lengths: the lengths are gathered first, the lists padded to the longest."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return mine
    world = dist.get_world_size(group)
    n = torch.tensor([mine.shape[0]], dtype=torch.int64, device=device)
    ns = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(ns, n, group=group)
    ns = [int(x.item()) for x in ns]
    width = int(np.prod(mine.shape[1:]))
    pad = np.zeros((max(max(ns), 1), width), np.float64)
    pad[:mine.shape[0]] = mine.reshape(mine.shape[0], width)
    t = torch.from_numpy(pad)
    if device is not None:
        t = t.to(device)
    out = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(out, t, group=group)
    parts = [o.cpu().numpy()[:k] for o, k in zip(out, ns)]
    return np.concatenate(parts, axis=0).reshape((-1,) + tuple(mine.shape[1:]))


def combine_lists(fns: Sequence[AggregationFunction], lists: np.ndarray) -> List[Optional[float]]:
    """finish() of every aggregate over float64[n, K, 4] super-tile partials folded one by one in row order
    (qe_agg_combine): the same additions in the same order whatever the number of partitions."""
    import ctypes as C
    out: List[Optional[float]] = []
    lists = np.ascontiguousarray(lists, dtype=np.float64)
    for k, fn in enumerate(fns):
        one = np.ascontiguousarray(lists[:, k, :])
        v, isnull = C.c_double(0.0), C.c_int(0)
        capi.check(capi.lib().qe_agg_combine(int(fn), one.ctypes.data_as(C.POINTER(capi.AggPartial)), one.shape[0],
                                             C.byref(v), C.byref(isnull)))
        out.append(None if isnull.value else v.value)
    return out


def sharded_aggregate(result, fns: Sequence[AggregationFunction], device=None, group=None) -> List[Optional[float]]:
    """This rank's aggregation result (PipelineResult of its partition) -> the global aggregates on every rank.
    The ranks exchange their super-tile partials (a few dozen bytes per 2^20 rows), not one total each, so the fold is
    over the same list for any number of ranks (partitions aligned to ALIGN_SUM)."""
    lists = [result.agg_partials(k) for k in range(len(fns))]                  # K x float64[n, 4]
    n = lists[0].shape[0] if lists else 0
    mine = np.stack(lists, axis=1) if lists else np.zeros((0, 0, 4))
    assert mine.shape == (n, len(fns), 4)
    return combine_lists(fns, all_gather_partial_lists(mine, device=device, group=group))


# ---------------------------------------------------------------- group-by across partitions
# GroupByAggregationOperator over a row-range sharded table: every rank groups its own rows (qe_groupby_run), the
# per-group records -- fixed-width arrays: key bits, null flags, global first row, partials -- are all-gathered as
# tensors and merged BY KEY on every rank by qe_groupby_merge: partials folded in rank order (qe_agg_combine, as for
# the global aggregates), groups ordered by the smallest global row index at which they appear -- the order a single
# scan of the whole table would have produced (LinkedHashMap insertion order,
# .../operator/GroupByAggregationOperator.kt:22,29-36).  STRING keys travel as dictionary codes, which are private to
# a process: the ranks' dictionaries are merged (in rank order) first and the codes re-coded.

def canonical_key_bits(v, dtype) -> int:
    """64-bit identity of one key value with the reference's equality: Double.equals semantics (all NaNs equal,
    -0.0 != 0.0), codes and booleans as integers."""
    from .ast import DataType
    if dtype == DataType.DOUBLE:
        f = float(v)
        if f != f:
            return 0x7ff8000000000000
        return int(np.array([f], dtype=np.float64).view(np.uint64)[0])
    return int(v) & 0xFFFFFFFFFFFFFFFF


def group_arrays(result, nkeys: int, nagg: int, row_offset: int = 0):
    """This rank's groups out of a qe_groupby_run result as fixed-width arrays:
    keys uint64[g, nkeys], key_null uint8[g, nkeys], first_row int64[g] (global), partials float64[g, nagg, 4], key types."""
    import ctypes as C
    from .ast import DataType
    n = result.num_rows
    keys = np.zeros((n, max(nkeys, 1)), np.uint64)
    nulls = np.zeros((n, max(nkeys, 1)), np.uint8)
    types = []
    for k in range(nkeys):
        vals, valid = result.column(k)
        t = result.column_type(k)[0]
        types.append(t)
        nulls[:, k] = ~np.asarray(valid, dtype=bool)
        if t == DataType.DOUBLE:
            bits = np.ascontiguousarray(vals, np.float64).view(np.uint64).copy()
            bits[np.isnan(vals)] = 0x7ff8000000000000
        elif t == DataType.BOOLEAN:
            bits = np.asarray(vals, dtype=np.uint64)
        else:                                                  # INT64 / DATE32 / STRING codes: the raw bits, zero-extended
            v = np.ascontiguousarray(vals)
            bits = v.view(np.uint32).astype(np.uint64) if v.dtype.itemsize == 4 else v.view(np.uint64)
        keys[:, k] = np.where(valid, bits, 0)
    first = np.zeros(max(n, 1), np.int64)
    capi.check(capi.lib().qe_result_group_first_rows(result.handle, first.ctypes.data))
    first = first[:n] + np.int64(row_offset)
    parts = np.zeros((n, max(nagg, 1), 4), np.float64)
    p = capi.AggPartial()
    for g in range(n):
        for j in range(nagg):
            capi.check(capi.lib().qe_result_group_partial(result.handle, g, j, C.byref(p)))
            parts[g, j] = (p.sum, np.array([p.count], np.int64).view(np.float64)[0], p.min, p.max)
    return keys, nulls, first, parts, types


def merge_group_arrays(fns: Sequence[AggregationFunction], nkeys: int, parts):
    """qe_groupby_merge over [(keys, key_null, first_row, partials)] in partition order ->
    (keys uint64[G, nkeys], key_null uint8[G, nkeys], first_row int64[G], values float64[G, nagg], value_null uint8[G, nagg])."""
    import ctypes as C
    nagg = len(fns)
    arr = (capi.GroupPart * max(1, len(parts)))()
    hold = []
    total = 0
    for i, (keys, nulls, first, partials) in enumerate(parts):
        keys = np.ascontiguousarray(keys, np.uint64)
        nulls = np.ascontiguousarray(nulls, np.uint8)
        first = np.ascontiguousarray(first, np.int64)
        partials = np.ascontiguousarray(partials, np.float64)
        hold.append((keys, nulls, first, partials))
        arr[i].ngroups = first.shape[0]
        arr[i].keys, arr[i].key_null = keys.ctypes.data, nulls.ctypes.data
        arr[i].first_row, arr[i].partials = first.ctypes.data, partials.ctypes.data
        total += first.shape[0]
    ok = np.zeros((max(total, 1), max(nkeys, 1)), np.uint64)
    on = np.zeros((max(total, 1), max(nkeys, 1)), np.uint8)
    of = np.zeros(max(total, 1), np.int64)
    ov = np.zeros((max(total, 1), max(nagg, 1)), np.float64)
    vn = np.zeros((max(total, 1), max(nagg, 1)), np.uint8)
    fnarr = (C.c_int32 * max(1, nagg))(*[int(f) for f in fns])
    ng = C.c_int64(0)
    capi.check(capi.lib().qe_groupby_merge(arr, len(parts), nkeys, fnarr, nagg, total, ok.ctypes.data, on.ctypes.data,
                                           of.ctypes.data, ov.ctypes.data, vn.ctypes.data, C.byref(ng)))
    g = ng.value
    return ok[:g, :nkeys], on[:g, :nkeys], of[:g], ov[:g, :nagg], vn[:g, :nagg]


def merged_rows(fns: Sequence[AggregationFunction], key_types, merged, decode=None) -> List[list]:
    """Reference-shaped rows [key..., finish()...] of a merge_group_arrays result; decode(code) -> str for STRING keys."""
    from .ast import DataType
    keys, knull, _first, vals, vnull = merged
    rows = []
    for g in range(keys.shape[0]):
        row = []
        for k, t in enumerate(key_types):
            if knull[g, k]:
                row.append(None)
            elif t == DataType.DOUBLE:
                row.append(float(np.array([keys[g, k]], np.uint64).view(np.float64)[0]))
            elif t == DataType.BOOLEAN:
                row.append(bool(keys[g, k]))
            elif t == DataType.STRING:
                row.append(decode(int(keys[g, k])) if decode else int(keys[g, k]))
            elif t == DataType.DATE32:
                row.append(float(np.array([keys[g, k]], np.uint64).astype(np.uint32).view(np.int32)[0]))
            else:
                row.append(float(np.array([keys[g, k]], np.uint64).view(np.int64)[0]))
        for j, fn in enumerate(fns):
            if vnull[g, j]:
                row.append(None)
            elif fn == AggregationFunction.COUNT:
                row.append(int(vals[g, j]))                    # CountAccumulator.finish() is an Int
            else:
                row.append(float(vals[g, j]))
        rows.append(row)
    return rows


def _all_gather_rows(mine: np.ndarray, device=None, group=None) -> List[np.ndarray]:
    """uint8[n_local, width] records of this rank -> the records of every rank, in rank order."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    n = torch.tensor([mine.shape[0]], dtype=torch.int64, device=device)
    ns = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(ns, n, group=group)
    ns = [int(x.item()) for x in ns]
    pad = np.zeros((max(max(ns), 1), mine.shape[1]), np.uint8)
    pad[:mine.shape[0]] = mine
    t = torch.from_numpy(pad)
    if device is not None:
        t = t.to(device)
    out = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(out, t, group=group)
    return [o.cpu().numpy()[:k] for o, k in zip(out, ns)]


def sharded_groupby(result, nkeys: int, fns: Sequence[AggregationFunction], dictionary, row_offset: int, device=None, group=None) -> List[list]:
    """This rank's group-by result -> the rows of the whole table, on every rank."""
    import torch.distributed as dist
    from .ast import DataType
    nagg = len(fns)
    keys, nulls, first, parts, types = group_arrays(result, nkeys, nagg, row_offset)
    decode = dictionary.decode
    if not (dist.is_available() and dist.is_initialized()):
        return merged_rows(fns, types, merge_group_arrays(fns, nkeys, [(keys, nulls, first, parts)]), decode)
    if any(t == DataType.STRING for t in types):
        # dictionaries are per process: merge them in rank order and re-code this rank's keys
        dicts = [None] * dist.get_world_size(group)
        dist.all_gather_object(dicts, list(dictionary._strings), group=group)       # the strings, not the groups
        merged_dict = {}
        for d in dicts:
            for s_ in d:
                merged_dict.setdefault(s_, len(merged_dict))
        strings = list(merged_dict)
        for k, t in enumerate(types):
            if t == DataType.STRING:
                keys[:, k] = [merged_dict[dictionary.decode(int(c))] if not nulls[g, k] else 0 for g, c in enumerate(keys[:, k])]
        decode = lambda c: strings[c]                                                # noqa: E731
    # one fixed-width record per group: keys | nulls | first row | partials
    n = first.shape[0]
    rec = np.concatenate([keys.reshape(n, -1).view(np.uint8).reshape(n, -1), nulls.reshape(n, -1),
                          first.reshape(n, 1).view(np.uint8).reshape(n, -1), parts.reshape(n, -1).view(np.uint8).reshape(n, -1)], axis=1)
    kw, nw = keys.shape[1] * 8, nulls.shape[1]
    gathered = []
    for r in _all_gather_rows(np.ascontiguousarray(rec), device=device, group=group):
        m = r.shape[0]
        gathered.append((np.ascontiguousarray(r[:, :kw]).view(np.uint64).reshape(m, -1),
                         np.ascontiguousarray(r[:, kw:kw + nw]),
                         np.ascontiguousarray(r[:, kw + nw:kw + nw + 8]).view(np.int64).reshape(m),
                         np.ascontiguousarray(r[:, kw + nw + 8:]).view(np.float64).reshape(m, -1, 4)))
    return merged_rows(fns, types, merge_group_arrays(fns, nkeys, gathered), decode)
