#!/usr/bin/env python
"""Option sweep of the multi-column filter + two projections (BASELINE.json configs[4] shape) on one GPU (development tool).

  python tools/sweep_lineitem.py --rows 500000000 --configs "stage_kb=32;stage_kb=40,ncw=8"
"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench
from queryengine_b200.engine import Context

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=500_000_000)
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--configs", default="staging=2")
args = ap.parse_args()
peak, _ = bench.measured_peak_gbs()
for cfg in args.configs.split(";"):
    ctx = Context(0)
    opts = dict(kv.split("=") for kv in cfg.split(",") if kv)
    for k, v in opts.items():
        ctx.set_option(k, int(v))
    try:
        r = bench.extra_sf_filter(ctx, args.rows, args.reps, 2, peak)
        row = dict(opts, kernel_ms=round(r["kernel_ms"], 3), gb_per_s=round(r["gb_per_s"], 1), frac=round(r["frac_of_peak"], 3))
    except Exception as e:
        row = dict(opts, error=repr(e))
    print(json.dumps(row), flush=True)
    ctx.close()
