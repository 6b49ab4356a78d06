#!/usr/bin/env python
"""Per-phase cycle breakdown of the register-sourced filter kernel (development tool)."""
import argparse, ctypes as C, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from queryengine_b200 import capi
from queryengine_b200.ast import DataType, Function, col, fn
from queryengine_b200.engine import ColumnarTable, Context, Field, Pipeline

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=1_000_000_000)
ap.add_argument("--shapes", default="2x4")
ap.add_argument("--flags", default="16,24")
ap.add_argument("--sel", type=float, default=0.5)
ap.add_argument("--grids", default="0")
args = ap.parse_args()
names = ["ticket", "load+eval", "barrier", "lookback", "scatter"]
for shape in args.shapes.split(","):
    steps, vec = (int(x) for x in shape.split("x"))
    for flags in (int(x) for x in args.flags.split(",")):
      for g in (int(x) for x in args.grids.split(",")):
        ctx = Context(0)
        for k, v in (("steps", steps), ("vec", vec), ("staging", 0), ("debug_flags", flags), ("grid_per_sm", g)):
            ctx.set_option(k, v)
        t = ColumnarTable(ctx, [Field("col", DataType.INT64, False)], args.rows)
        t.generate(0, capi.QE_GEN_UNIFORM_U40, 42)
        pipe = Pipeline(t, fn(Function.CMP_LT, col(0, DataType.INT64), float(int(args.sel * 2 ** 40))), [col(0, DataType.INT64)])
        res = pipe.run(); res = pipe.run(res); _ = res.num_rows
        buf = np.zeros(4096 * 8, np.uint64)
        capi.check(capi.lib().qe_result_debug_counters(res.handle, buf.ctypes.data, buf.size))
        d = buf.reshape(4096, 8)
        d = d[d[:, 5] > 0]
        tiles = d[:, 5].astype(np.float64)
        per = d[:, :5].astype(np.float64) / tiles[:, None]
        print(f"shape {shape} flags {flags} grid/sm {g}: kernel {res.kernel_ms:.3f} ms, {len(d)} CTAs, tiles/CTA {tiles.mean():.0f}; "
              "cycles per tile: " + ", ".join(f"{n} {per[:, i].mean():.0f}" for i, n in enumerate(names)) +
              f"; total {per.sum(axis=1).mean():.0f}")
        res.close(); t.close(); ctx.close()
