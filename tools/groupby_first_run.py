#!/usr/bin/env python
"""Wall time of the FIRST group-by of a query (no remembered variant: every level that overflows is tried once) and of
a repeat into the same result (development tool).  python tools/groupby_first_run.py"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from queryengine_b200 import capi
from queryengine_b200.ast import AggregationFunction as A, DataType, col
from queryengine_b200.engine import ColumnarTable, Context, Field, GroupByPipeline
ctx = Context(0)
D = DataType.DOUBLE
n = 600_000_000
t = ColumnarTable(ctx, [Field("k", D), Field("x", D)], n)
t.generate(1, capi.QE_GEN_UNIFORM_REAL, 5, 0.0, 1.0)
# compile every kernel once on a small table so that the walls below hold no JIT time
for card in (8, 100, 1000, 100000):
    t.generate(0, capi.QE_GEN_UNIFORM_INT, 6, 0, card - 1)
    ctx.synchronize()
    pipe = GroupByPipeline(t, None, [col(0)], [col(1), col(1)], [A.SUM, A.COUNT], max_groups=max(65536, 2 * card))
    for label in ("jit+first", "first", "repeat"):
        t0 = time.perf_counter()
        res = pipe.run() if label != "repeat" else pipe.run(res)
        g = res.num_rows
        wall = (time.perf_counter() - t0) * 1e3
        print(card, label, "groups", g, "variant", capi.lib().qe_result_group_variant(res.handle), "wall_ms", round(wall, 2), flush=True)
        if label == "jit+first":
            res.close()
    res.close()
