#!/usr/bin/env python
"""Per-role cycle breakdown of the ring filter kernel (development tool; QE_DEBUG_FLAGS bit 16).

  python tools/phase_timing.py --sel 0.5 --configs "stage_kb=16,defer=6;stage_kb=32,defer=3,nlb=1"
"""
import argparse, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from queryengine_b200 import capi
from queryengine_b200.ast import DataType, Function, col, fn
from queryengine_b200.engine import ColumnarTable, Context, Field, Pipeline

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=1_000_000_000)
ap.add_argument("--sel", type=float, default=0.5)
ap.add_argument("--configs", default="staging=2")
args = ap.parse_args()
names = {0: "prod.wait_empty", 1: "prod.wait_loads", 2: "prod.tiles", 3: "lb.wait_eval", 4: "lb.first_read", 5: "lb.repoll",
         6: "lb.repolls", 7: "lb.wait_prev", 8: "lb.gap_len", 9: "cons.wait_full", 10: "cons.phaseA", 11: "cons.wait_pref",
         12: "cons.phaseC", 14: "lb.early_reads"}
ctx = Context(0)
I = DataType.INT64
t = ColumnarTable(ctx, [Field("col", I, False)], args.rows)
t.generate(0, capi.QE_GEN_UNIFORM_U40, 42)
for cfg in args.configs.split(";"):
    for k in ("steps", "vec", "grid_per_sm", "stages", "stage_kb", "tile_rows", "ncw", "defer", "nlb", "loads"):
        ctx.set_option(k, 0)
    ctx.set_option("staging", 2)
    flags = 16
    for kv in cfg.split(","):
        if not kv:
            continue
        k, v = kv.split("=")
        if k == "debug_flags":
            flags |= int(v)
        else:
            ctx.set_option(k, int(v))
    ctx.set_option("debug_flags", flags)
    pipe = Pipeline(t, fn(Function.CMP_LT, col(0, I), float(int(args.sel * 2 ** 40))), [col(0, I)])
    res = pipe.run(); res = pipe.run(res); _ = res.num_rows
    buf = np.zeros(4096 * 8, np.uint64)
    capi.check(capi.lib().qe_result_debug_counters(res.handle, buf.ctypes.data, buf.size))
    d = buf.reshape(2048, 16).astype(np.float64)
    d = d[d[:, 2] > 0]
    tiles = d[:, 2] - 1          # the producer also counts its end marker
    nlb = max(1, int(dict(kv.split("=") for kv in cfg.split(",") if kv).get("nlb", 1)))
    out = [f"cfg {cfg} sel {args.sel}: kernel {res.kernel_ms:.3f} ms, {len(d)} CTAs, tiles/CTA {tiles.mean():.0f} "
           f"(min {tiles.min():.0f} max {tiles.max():.0f})"]
    for i, n in names.items():
        if i == 2:
            continue
        per = d[:, i] / np.maximum(tiles, 1)
        if n.startswith("lb."):
            per = per * nlb      # look-back warp 0 sees every nlb-th tile
        out.append(f"{n} {per.mean():.0f}")
    print("; ".join(out) + "   [cycles or counts per tile]", flush=True)
    res.close()
t.close(); ctx.close()
