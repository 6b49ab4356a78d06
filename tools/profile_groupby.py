#!/usr/bin/env python
"""One group-by launch for ncu (development tool):  ncu --set full -k qe_groupby_private_kernel ... python tools/profile_groupby.py --card 32"""
import argparse, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from queryengine_b200 import capi
from queryengine_b200.ast import AggregationFunction as A, DataType, col
from queryengine_b200.engine import ColumnarTable, Context, Field, GroupByPipeline

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=100_000_000)
ap.add_argument("--card", type=int, default=32)
ap.add_argument("--q1", action="store_true")
ap.add_argument("--atomic", action="store_true")
ap.add_argument("--runs", type=int, default=3)
args = ap.parse_args()
ctx = Context(0)
if args.q1:
    t, pipe = bench.q1_pipeline(ctx, args.rows, atomic=args.atomic)
else:
    D = DataType.DOUBLE
    t = ColumnarTable(ctx, [Field("k", D), Field("x", D)], args.rows)
    t.generate(1, capi.QE_GEN_UNIFORM_REAL, 5, 0.0, 1.0)
    t.generate(0, capi.QE_GEN_UNIFORM_INT, 6, 0, args.card - 1)
    pipe = GroupByPipeline(t, None, [col(0)], [col(1), col(1)], [A.SUM, A.COUNT], max_groups=max(65536, 2 * args.card), atomic=args.atomic)
res = pipe.run()
for _ in range(args.runs):
    pipe.run(res)
print("groups", res.num_rows, "kernel_ms", res.kernel_stats())
