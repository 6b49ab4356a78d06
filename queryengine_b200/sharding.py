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

ALIGN = 128


def partition(nrows: int, world: int, rank: int) -> Tuple[int, int]:
    """Rows [begin, end) of rank `rank`: ceil(nrows / world) rounded up to a multiple of 128 per rank."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError("bad rank / world size")
    per = -(-nrows // world)
    per = -(-per // ALIGN) * ALIGN
    begin = min(nrows, rank * per)
    end = min(nrows, begin + per)
    return begin, end


def partitions(nrows: int, world: int) -> List[Tuple[int, int]]:
    return [partition(nrows, world, r) for r in range(world)]


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


def sharded_aggregate(result, fns: Sequence[AggregationFunction], device=None, group=None) -> List[Optional[float]]:
    """This rank's aggregation result (PipelineResult of its partition) -> the global aggregates on every rank."""
    mine = [result.agg_partial(k) for k in range(len(fns))]
    return combine(fns, all_gather_partials(mine, device=device, group=group))


# ---------------------------------------------------------------- group-by across partitions
# GroupByAggregationOperator over a row-range sharded table: every rank groups its own rows (qe_groupby_run),
# the per-group partials (a handful of rows) are all-gathered as Python objects and merged BY KEY on every
# rank: partials folded in rank order (qe_agg_combine, as for the global aggregates), groups ordered by the
# smallest global row index at which they appear -- the order a single scan of the whole table would have
# produced (LinkedHashMap insertion order, .../operator/GroupByAggregationOperator.kt:22,29-36).

def _key_id(v):
    """Hashable identity of one key value with the reference's equality: Double.equals semantics (all NaNs
    equal, -0.0 != 0.0), null a value of its own."""
    if isinstance(v, float):
        if v != v:
            return ("f", "nan")
        return ("f", np.array([v], dtype=np.float64).view(np.uint64)[0].item())
    return (type(v).__name__, v)


def group_partials(result, nkeys: int, fns: Sequence[AggregationFunction], dictionary, row_offset: int = 0):
    """This rank's groups out of a qe_groupby_run result: [(keys, first_row_global, [(sum, count, min, max)])].
    STRING keys are decoded (dictionary codes are private to a process)."""
    import ctypes as C
    from .ast import DataType
    n = result.num_rows
    cols = [result.column(k) for k in range(nkeys)]
    types = [result.column_type(k)[0] for k in range(nkeys)]
    first = np.zeros(max(n, 1), np.int64)
    capi.check(capi.lib().qe_result_group_first_rows(result.handle, first.ctypes.data))
    out = []
    for g in range(n):
        key = []
        for k in range(nkeys):
            vals, valid = cols[k]
            if not valid[g]:
                key.append(None)
            elif types[k] == DataType.STRING:
                key.append(dictionary.decode(vals[g]))
            elif types[k] == DataType.BOOLEAN:
                key.append(bool(vals[g]))
            else:
                key.append(float(vals[g]))
        parts = []
        for j in range(len(fns)):
            p = capi.AggPartial()
            capi.check(capi.lib().qe_result_group_partial(result.handle, g, j, C.byref(p)))
            parts.append((p.sum, p.count, p.min, p.max))
        out.append((tuple(key), int(first[g]) + int(row_offset), parts))
    return out


def combine_groups(fns: Sequence[AggregationFunction], by_rank) -> List[list]:
    """Merge per-rank group lists (rank order) into the rows of the whole table: [key..., finish()...]."""
    merged = {}
    for rank_groups in by_rank:
        for key, first, parts in rank_groups:
            kid = tuple(_key_id(v) for v in key)
            e = merged.get(kid)
            if e is None:
                merged[kid] = [key, first, [[p] for p in parts]]
            else:
                e[1] = min(e[1], first)
                for j, p in enumerate(parts):
                    e[2][j].append(p)
    rows = []
    for key, first, parts in sorted(merged.values(), key=lambda e: e[1]):
        finished = combine(fns, [[capi.AggPartial(*p[r]) for p in parts] for r in range(len(parts[0]))]) if parts else []
        rows.append(list(key) + [int(v) if (v is not None and fn == AggregationFunction.COUNT) else v
                                 for v, fn in zip(finished, fns)])
    return rows


def sharded_groupby(result, nkeys: int, fns: Sequence[AggregationFunction], dictionary, row_offset: int, group=None) -> List[list]:
    """This rank's group-by result -> the rows of the whole table, on every rank."""
    import torch.distributed as dist
    mine = group_partials(result, nkeys, fns, dictionary, row_offset)
    if not (dist.is_available() and dist.is_initialized()):
        return combine_groups(fns, [mine])
    gathered = [None] * dist.get_world_size(group)
    dist.all_gather_object(gathered, mine, group=group)
    return combine_groups(fns, gathered)
