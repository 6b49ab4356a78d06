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
