#!/usr/bin/env python
"""Group-by kernel timing on one GPU (development tool): TPC-H Q1 shape, private and atomic variants,
and a cardinality sweep of `SELECT k, SUM(x), COUNT(x) GROUP BY k`.

  python tools/bench_groupby.py --rows 600000000 --cards 4,64,1024,65536
"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench
from queryengine_b200 import capi
from queryengine_b200.ast import AggregationFunction as A, DataType, Function, col, fn
from queryengine_b200.engine import ColumnarTable, Context, Field, GroupByPipeline

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=600_000_000)
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--cards", default="4,32,1024,65536")
ap.add_argument("--opts", default="")
ap.add_argument("--skip-q1", action="store_true")
ap.add_argument("--key-scale", type=float, default=0.0, help="group by k * scale instead of k: keys that are not small integers")
args = ap.parse_args()
peak, _ = bench.measured_peak_gbs()
ctx = Context(0)
for kv in (x for x in args.opts.split(",") if x):
    k, v = kv.split("=")
    ctx.set_option(k, int(v))


def timed(pipe, reps):
    import time
    t0 = time.perf_counter()
    res = pipe.run()
    n = res.num_rows
    first_ms = (time.perf_counter() - t0) * 1e3
    pipe.run(res)
    m0, n0 = res.kernel_stats()
    t0 = time.perf_counter()
    for _ in range(reps):
        pipe.run(res)
        n = res.num_rows
    wall_ms = (time.perf_counter() - t0) * 1e3 / max(1, reps)
    m1, n1 = res.kernel_stats()
    ms = (m1 - m0) / max(1, n1 - n0)
    v = capi.lib().qe_result_group_variant(res.handle)
    g = (res.num_rows, capi.lib().qe_result_group_threads(res.handle), capi.lib().qe_result_group_copies(res.handle), round(wall_ms, 3), round(first_ms, 1))
    res.close()
    return ms, v, g


for atomic in (() if args.skip_q1 else (False, True)):
    t, pipe = bench.q1_pipeline(ctx, args.rows, atomic=atomic)
    ms, v, g = timed(pipe, args.reps)
    print(json.dumps({"what": "q1", "rows": args.rows, "variant": ["private", "atomic", "shared", "registers", "lanes", "lanes8", "warps", "twopass"][v], "groups": g[0], "threads": g[1], "kernel_ms": round(ms, 3),
                      "gb_per_s": round(args.rows * 48 / ms / 1e6, 1), "frac": round(args.rows * 48 / ms / 1e6 / peak, 3)}), flush=True)
    t.close()
D = DataType.DOUBLE
t = ColumnarTable(ctx, [Field("k", D), Field("x", D)], args.rows)
t.generate(1, capi.QE_GEN_UNIFORM_REAL, 5, 0.0, 1.0)
for card in (int(c) for c in args.cards.split(",")):
    t.generate(0, capi.QE_GEN_UNIFORM_INT, 6, 0, card - 1)
    key = fn(Function.MUL, col(0), args.key_scale) if args.key_scale else col(0)
    pipe = GroupByPipeline(t, None, [key], [col(1), col(1)], [A.SUM, A.COUNT], max_groups=max(65536, 2 * card))
    ms, v, g = timed(pipe, args.reps)
    print(json.dumps({"what": "sum_count_by_k", "cardinality": card, "variant": ["private", "atomic", "shared", "registers", "lanes", "lanes8", "warps", "twopass"][v], "groups": g[0], "threads": g[1], "copies": g[2],
                      "wall_ms": g[3], "first_run_ms": g[4], "kernel_ms": round(ms, 3), "gb_per_s": round(args.rows * 16 / ms / 1e6, 1),
                      "frac": round(args.rows * 16 / ms / 1e6 / peak, 3)}), flush=True)
