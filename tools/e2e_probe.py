#!/usr/bin/env python
"""End-to-end (host buffers in, host buffers out) step of bench.py's e2e leg under different context options
(development tool):   python tools/e2e_probe.py --configs "staging=2;list_cap=-1" [--rows 1000000000]"""
import argparse
import ctypes as C
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
from queryengine_b200 import capi  # noqa: E402
from queryengine_b200.ast import DataType, col  # noqa: E402
from queryengine_b200.engine import ColumnarTable, Context, Field, Pipeline, stream_pipelines  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=1_000_000_000)
ap.add_argument("--chunk", type=int, default=1 << 25)
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--configs", default="staging=2")
args = ap.parse_args()
import torch  # noqa: E402
I64 = DataType.INT64
ctx = Context(0)
n = args.rows
caps = [int(s * n * 1.02) + 4096 for s in bench.SELECTIVITIES]
h_in = bench.PinnedBuffer(n * 8)
h_sel = [bench.PinnedBuffer(c * 4) for c in caps]
h_val = [bench.PinnedBuffer(c * 8) for c in caps]
gen_t = ColumnarTable(ctx, [Field("col", I64, False)], min(n, 1 << 28))
done = 0
while done < n:
    m = min(gen_t.nrows, n - done)
    gen_t.generate(0, capi.QE_GEN_UNIFORM_U40, bench.SEED, row_offset=done)
    dense = Pipeline(gen_t, None, [col(0, I64, "col")]).run(None, 0, m)
    capi.check(capi.lib().qe_result_column(dense.handle, 0, C.c_void_p(h_in.ptr.value + 8 * done), None))
    dense.close()
    done += m
gen_t.close()
ctx.trim()
for cfg in args.configs.split(";"):
    opts = dict(kv.split("=") for kv in cfg.split(",") if kv)
    for k in ("list_cap", "nsw", "tile_rows", "ncw", "stages", "stage_kb", "defer"):
        ctx.set_option(k, 0)
    for k, v in opts.items():
        ctx.set_option(k, int(v))
    t = ColumnarTable(ctx, [Field("col", I64, False)], 3 * args.chunk)
    row = dict(opts)
    for which, name in ((0, "full"), (1, "sel_only")):
        pipes = [Pipeline(t, bench.filter_expr(s), [col(0, I64, "col")], out_capacity_rows=c, values=(which == 0))
                 for s, c in zip(bench.SELECTIVITIES, caps)]
        results = [None] * 3

        def one_step():
            res = stream_pipelines(t, pipes, [h_in.ptr.value], chunk_rows=args.chunk, results=results,
                                   host_selection=[b.ptr.value for b in h_sel],
                                   host_out_values=[[b.ptr.value] for b in h_val] if which == 0 else None, total_rows=n)
            for i, r in enumerate(res):
                results[i] = r
            return [r.num_rows for r in res]
        one_step()
        per = []
        for _ in range(args.steps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            one_step()
            torch.cuda.synchronize()
            per.append(round((time.perf_counter() - t0) * 1e3, 1))
        row[name + "_ms"] = per
        row[name + "_kernel_ms"] = [round(r.kernel_stats()[0], 2) for r in results]
        for r in results:
            r.close()
    t.close()
    print(json.dumps(row), flush=True)
