#!/usr/bin/env python
"""Option sweep of the filter+compaction kernel on one GPU (development tool).

  python tools/sweep_filter.py --configs "staging=2,stage_kb=32,ncw=8,defer=2;staging=0,steps=4,vec=4"

Each config is a ';'-separated list of qe_ctx_set_option key=value pairs.  Prints, per config, the device
time of the generated kernel for the selectivities of BASELINE.json configs[1].
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from bench import SEED, measured_peak_gbs  # noqa: E402
from queryengine_b200 import capi  # noqa: E402
from queryengine_b200.ast import DataType, Function, col, fn  # noqa: E402
from queryengine_b200.engine import ColumnarTable, Context, Field, Pipeline  # noqa: E402


def time_pipeline(pipe, reps):
    res = pipe.run()
    for _ in range(2):
        pipe.run(res)
    m0, n0 = res.kernel_stats()
    for _ in range(reps):
        pipe.run(res)
    m1, n1 = res.kernel_stats()
    kept = res.num_rows
    res.close()
    return (m1 - m0) / max(1, n1 - n0), kept


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=1_000_000_000)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--type", default="i64", choices=["i64", "f64"])
    ap.add_argument("--configs", default="staging=2")
    ap.add_argument("--sels", default="0.01,0.5,0.99")
    ap.add_argument("--null-rate", type=float, default=-1.0, help=">= 0: the column is nullable (its compacted copy then has a validity bitmap)")
    ap.add_argument("--payload", action="store_true", help="a second, nullable f64 column (10 %% nulls) the filter does not read is returned too: "
                                                           "its validity is a compacted bit column that has to be assembled row by row")
    args = ap.parse_args()
    peak, _ = measured_peak_gbs()
    dt = DataType.INT64 if args.type == "i64" else DataType.DOUBLE
    ctx0 = Context(0)
    fields = [Field("col", dt, args.null_rate >= 0)] + ([Field("y", DataType.DOUBLE, True)] if args.payload else [])
    t = ColumnarTable(ctx0, fields, args.rows)
    if args.payload:
        t.generate(1, capi.QE_GEN_UNIFORM_REAL, SEED + 1, 0.0, 1.0, null_rate=0.1)
    if args.null_rate >= 0:
        t.generate(0, capi.QE_GEN_UNIFORM_U40, SEED, null_rate=args.null_rate)
    else:
        t.generate(0, capi.QE_GEN_UNIFORM_U40, SEED)
    for cfg in args.configs.split(";"):
        opts = dict(kv.split("=") for kv in cfg.split(",") if kv)
        # options are per context; the table stays on ctx0, so reset every known option first
        for k in ("steps", "vec", "grid_per_sm", "stages", "stage_kb", "tile_rows", "ncw", "defer", "debug_flags", "nlb", "loads", "list_cap", "nsw"):
            ctx0.set_option(k, 0)
        ctx0.set_option("staging", -1)
        ctx0.set_option("pipeline", 1)
        for k, v in opts.items():
            ctx0.set_option(k, int(v))
        row = dict(opts)
        try:
            for s in [float(x) for x in args.sels.split(",")]:
                k = float(int(s * 2 ** 40))
                outs = [col(0, dt, "col")] + ([col(1, DataType.DOUBLE, "y")] if args.payload else [])
                pipe = Pipeline(t, fn(Function.CMP_LT, col(0, dt, "col"), k), outs)
                ms, kept = time_pipeline(pipe, args.reps)
                gbs = ((16.125 if args.payload else 8.0) * args.rows + (20.125 if args.payload else 12.0) * kept) / (ms * 1e-3) / 1e9
                row[f"ms_{s}"] = round(ms, 3)
                row[f"gbs_{s}"] = round(gbs, 1)
                row[f"frac_{s}"] = round(gbs / peak, 3)
        except Exception as e:
            row["error"] = repr(e)
        print(json.dumps(row), flush=True)
    t.close()
    ctx0.close()


if __name__ == "__main__":
    main()
