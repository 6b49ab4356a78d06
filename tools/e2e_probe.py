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
ap.add_argument("--window", type=int, default=3, help="chunk slots of the device window")
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
    t = ColumnarTable(ctx, [Field("col", I64, False)], args.window * args.chunk)
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

# ---- the same bytes as plain copies INTO / OUT OF THE SAME host buffers (distinct memory, not a recycled 1 GiB buffer)
import glob  # noqa: E402
cands = glob.glob(os.path.join(os.path.dirname(torch.__file__), "lib", "libcudart*.so*")) + ["libcudart.so.12", "libcudart.so"]
rt = None
for cnd in cands:
    try:
        rt = C.CDLL(cnd)
        break
    except OSError:
        pass
if rt is not None:
    rt.cudaMemcpyAsync.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
    s_up, s_down = torch.cuda.Stream(), torch.cuda.Stream()
    dbuf_up = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")
    dbuf_down = torch.ones(1 << 30, dtype=torch.uint8, device="cuda")
    piece = 64 << 20
    for mode in ("both", "d2h_only", "h2d_only"):
        per = []
        for _ in range(3):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            if mode != "d2h_only":
                off = 0
                while off < n * 8:
                    m = min(piece, n * 8 - off)
                    rt.cudaMemcpyAsync(C.c_void_p(dbuf_up.data_ptr() + off % (1 << 30)), C.c_void_p(h_in.ptr.value + off), m, 1, C.c_void_p(s_up.cuda_stream))
                    off += m
            if mode != "h2d_only":
                for hb, width in [(b, 4) for b in h_sel] + [(b, 8) for b in h_val]:
                    total = (hb.nbytes if hasattr(hb, "nbytes") else 0)
                for hb, cap_, width in [(h_sel[i], caps[i], 4) for i in range(3)] + [(h_val[i], caps[i], 8) for i in range(3)]:
                    total = int(bench.SELECTIVITIES[[h_sel, h_val][width == 8].index(hb)] * n) * width
                    off = 0
                    while off < total:
                        m = min(piece, total - off)
                        rt.cudaMemcpyAsync(C.c_void_p(hb.ptr.value + off), C.c_void_p(dbuf_down.data_ptr() + off % ((1 << 30) - piece)), m, 2, C.c_void_p(s_down.cuda_stream))
                        off += m
            torch.cuda.synchronize()
            per.append(round((time.perf_counter() - t0) * 1e3, 1))
        print(json.dumps({"plain_copies_same_host_buffers": mode, "ms": per}), flush=True)
