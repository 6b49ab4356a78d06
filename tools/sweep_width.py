#!/usr/bin/env python
"""Row-width sweep of the ring filter kernel on one GPU (development tool).

`SELECT c0 .. c{W-1} FROM t WHERE c0 < k` over W f64 columns, W = 1..10: the filter reads 8*W bytes per row and
writes 8*W + 4 bytes per kept row (compacted columns + selection vector).  Prints device time and the
fraction of the measured copy peak, to check the tile-size rule of choose_shape for every row width.

  python tools/sweep_width.py --bytes 8000000000 --sel 0.5 --widths 1,2,3,4,5,6,8,10 --opts "stage_kb=40"
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bytes", type=float, default=8e9, help="input bytes per run (rows = bytes / (8*W))")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--sel", type=float, default=0.5)
    ap.add_argument("--widths", default="1,2,3,4,5,6,8,10")
    ap.add_argument("--opts", default="")
    args = ap.parse_args()
    peak, _ = measured_peak_gbs()
    for w in (int(x) for x in args.widths.split(",")):
        ctx = Context(0)
        for kv in (x for x in args.opts.split(",") if x):
            k, v = kv.split("=")
            ctx.set_option(k, int(v))
        rows = int(args.bytes / (8 * w)) // 128 * 128
        t = ColumnarTable(ctx, [Field("c%d" % i, DataType.DOUBLE, False) for i in range(w)], rows)
        for i in range(w):
            t.generate(i, capi.QE_GEN_UNIFORM_INT, SEED + i, 0, 999)
        pipe = Pipeline(t, fn(Function.CMP_LT, col(0), float(int(args.sel * 1000))), [col(i) for i in range(w)])
        row = {"width": w, "rows": rows}
        try:
            res = pipe.run()
            for _ in range(2):
                pipe.run(res)
            m0, n0 = res.kernel_stats()
            for _ in range(args.reps):
                pipe.run(res)
            m1, n1 = res.kernel_stats()
            ms = (m1 - m0) / max(1, n1 - n0)
            kept = res.num_rows
            nbytes = rows * 8.0 * w + kept * (8.0 * w + 4.0)
            row.update(kept=kept, ms=round(ms, 3), gb_per_s=round(nbytes / ms / 1e6, 1), frac=round(nbytes / ms / 1e6 / peak, 3))
            res.close()
        except Exception as e:
            row["error"] = repr(e)
        print(json.dumps(row), flush=True)
        t.close()
        ctx.close()


if __name__ == "__main__":
    main()
