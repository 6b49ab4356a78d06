#!/usr/bin/env python
"""Tile-shape sweep of the filter+compaction kernel on one GPU (development tool).

Prints, per (steps, vec, CTAs/SM cap), the device time of the generated kernel for
the three selectivities of BASELINE.json configs[1], and the Q6 / tree kernels.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from bench import SEED, SELECTIVITIES, filter_expr, measured_peak_gbs  # noqa: E402
from queryengine_b200 import capi  # noqa: E402
from queryengine_b200.ast import DataType, col  # noqa: E402
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
    ap.add_argument("--shapes", default="1x4,2x4,4x4,8x4,2x2,4x2,8x2")
    ap.add_argument("--grids", default="0")
    ap.add_argument("--staging", default="1", help="comma list of 0/1")
    ap.add_argument("--stages", default="0", help="comma list")
    ap.add_argument("--debug-flags", default="0", help="comma list of QE_DEBUG_FLAGS values (timing experiments)")
    ap.add_argument("--sels", default="0.01,0.5,0.99")
    ap.add_argument("--pipeline", default="1", help="comma list of 0/1")
    args = ap.parse_args()
    peak, _ = measured_peak_gbs()
    dt = DataType.INT64 if args.type == "i64" else DataType.DOUBLE
    out = []
    for shape in args.shapes.split(","):
        steps, vec = (int(x) for x in shape.split("x"))
        for g, stg, nst, dbg, pl in [(int(g), int(a), int(b), int(d), int(e)) for g in args.grids.split(",")
                                     for a in args.staging.split(",") for b in (args.stages.split(",") if a == "1" else ["0"])
                                     for d in args.debug_flags.split(",") for e in args.pipeline.split(",")]:
            if stg and vec != 2:
                continue
            ctx = Context(0)
            ctx.set_option("steps", steps)
            ctx.set_option("vec", vec)
            ctx.set_option("grid_per_sm", g)
            ctx.set_option("staging", stg)
            ctx.set_option("stages", nst)
            ctx.set_option("debug_flags", dbg)
            ctx.set_option("pipeline", pl)
            t = ColumnarTable(ctx, [Field("col", dt, False)], args.rows)
            t.generate(0, capi.QE_GEN_UNIFORM_U40, SEED)
            row = {"steps": steps, "vec": vec, "grid_per_sm": g, "staging": stg, "stages": nst, "dbg": dbg, "pipeline": pl}
            for s in [float(x) for x in args.sels.split(",")]:
                from queryengine_b200.ast import Function, fn
                k = float(int(s * 2 ** 40))
                pipe = Pipeline(t, fn(Function.CMP_LT, col(0, dt, "col"), k), [col(0, dt, "col")])
                ms, kept = time_pipeline(pipe, args.reps)
                gbs = (8.0 * args.rows + 12.0 * kept) / (ms * 1e-3) / 1e9
                row[f"ms_{s}"] = round(ms, 3)
                row[f"gbs_{s}"] = round(gbs, 1)
                row[f"frac_{s}"] = round(gbs / peak, 3)
            print(json.dumps(row), flush=True)
            out.append(row)
            t.close()
            ctx.close()


if __name__ == "__main__":
    main()
