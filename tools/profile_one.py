#!/usr/bin/env python
"""Run one pipeline a few times (for ncu captures; development tool).

  python tools/profile_one.py --what filter --sel 0.5 --rows 1000000000 --reps 3 [--opt key=value ...]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from queryengine_b200 import capi  # noqa: E402
from queryengine_b200.ast import DataType, Function, col, fn  # noqa: E402
from queryengine_b200.engine import ColumnarTable, Context, Field, Pipeline  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--what", default="filter", choices=["filter", "tree", "q6", "sf", "q1"])
ap.add_argument("--sel", type=float, default=0.5)
ap.add_argument("--rows", type=int, default=1_000_000_000)
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--opt", action="append", default=[])
args = ap.parse_args()
ctx = Context(0)
for kv in args.opt:
    k, v = kv.split("=")
    ctx.set_option(k, int(v))
I, D = DataType.INT64, DataType.DOUBLE
if args.what == "filter":
    t = ColumnarTable(ctx, [Field("col", I, False)], args.rows)
    t.generate(0, capi.QE_GEN_UNIFORM_U40, 42)
    pipe = Pipeline(t, fn(Function.CMP_LT, col(0, I), float(int(args.sel * 2 ** 40))), [col(0, I)])
elif args.what == "tree":
    import bench
    t = ColumnarTable(ctx, [Field(n, D, True) for n in "abcd"], args.rows)
    for c, rate in enumerate((0.0, 0.01, 0.10, 0.50)):
        t.generate(c, capi.QE_GEN_UNIFORM_REAL, 52 + c, -1000.0, 1000.0, null_rate=rate)
    pipe = Pipeline(t, None, [bench.tree_expr()])
elif args.what == "sf":
    import bench
    t, pipe = bench.sf_filter_pipeline(ctx, args.rows)
elif args.what == "q1":
    import bench
    t, pipe = bench.q1_pipeline(ctx, args.rows)
else:
    import bench
    t, pipe = bench.q6_pipeline(ctx, args.rows)
res = pipe.run()
for _ in range(args.reps - 1):
    res = pipe.run(res)
print(args.what, "rows out", res.num_rows, "kernel ms", res.kernel_ms)
res.close()
t.close()
ctx.close()
