#!/usr/bin/env python
"""Cold-query latency and the no-NVRTC path, measured in a fresh process (development tool; bench.py reports the same
figures as `first_query_ms` and `interpreter`).

  python tools/cold_start.py [--rows N]

Prints one JSON line: wall time from qe_ctx_create to the first result of configs[1]'s 50 % filter (NVRTC or the
cubin caches included), the same query again, a second literal, and the interpreter kernels' device time on the sweep.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from queryengine_b200 import capi  # noqa: E402
from queryengine_b200.ast import DataType, Function, col, fn  # noqa: E402
from queryengine_b200.engine import ColumnarTable, Context, Field, Pipeline  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=1_000_000_000)
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--skip-interpreter", action="store_true")
args = ap.parse_args()
I = DataType.INT64
out = {"tool": "cold_start", "rows": args.rows}
t0 = time.perf_counter()
ctx = Context(0)
out["ctx_create_ms"] = (time.perf_counter() - t0) * 1e3
t = ColumnarTable(ctx, [Field("col", I, False)], args.rows)
t.generate(0, capi.QE_GEN_UNIFORM_U40, 42)
ctx.synchronize()


def filt(s):
    return fn(Function.CMP_LT, col(0, I), float(int(s * 2 ** 40)))


def first(s, **kw):
    t1 = time.perf_counter()
    p = Pipeline(t, filt(s), [col(0, I)], **kw)
    r = p.run()
    n = r.num_rows
    ms = (time.perf_counter() - t1) * 1e3
    return p, r, n, ms


p, r, n, ms = first(0.5)
out["first_query_ms"] = ms
out["first_query_kernel_ms"] = r.kernel_ms
t1 = time.perf_counter()
r = p.run(r)
n = r.num_rows
out["same_query_again_ms"] = (time.perf_counter() - t1) * 1e3
t1 = time.perf_counter()
r = p.run(r)
n = r.num_rows
out["same_query_third_ms"] = (time.perf_counter() - t1) * 1e3
r.close()
p2, r2, n2, ms2 = first(0.25)
out["other_literal_first_ms"] = ms2
r2.close()
# a pipeline that is not in the ahead-of-time table (the one bench.py's cold_start leg compiled before): disk cache or NVRTC
from queryengine_b200.engine import jit_stats  # noqa: E402
s0 = jit_stats()
t1 = time.perf_counter()
odd = fn(Function.CMP_LT, fn(Function.SUB, col(0, I), 12345.0), float(int(0.25 * 2 ** 40)))
r4 = Pipeline(t, odd, [col(0, I)]).run()
n4 = r4.num_rows
out["uncached_pipeline_first_result_ms"] = (time.perf_counter() - t1) * 1e3
s1 = jit_stats()
out["uncached_pipeline_origin"] = {k: s1[k] - s0[k] for k in ("aot", "disk", "nvrtc")}
out["jit_stats"] = s1
r4.close()
# interpreter kernels (what a deployment without libnvrtc runs)
interp = []
for s in (() if args.skip_interpreter else (0.01, 0.5, 0.99)):
    p3, r3, n3, ms3 = first(s, interpreter=True)
    for _ in range(args.reps):
        r3 = p3.run(r3)
    tot, runs = r3.kernel_stats()
    kms = tot / max(1, runs)
    nbytes = 8.0 * args.rows + 12.0 * n3
    interp.append({"selectivity": s, "kept": n3, "kernel_ms": kms, "gb_per_s": nbytes / (kms * 1e-3) / 1e9, "used_jit": r3.used_jit})
    r3.close()
out["interpreter"] = interp
t.close()
ctx.close()
print(json.dumps(out))
