#!/usr/bin/env python
"""Generate the CUDA source and sm_100a cubin of a pipeline without a GPU (development tool).

  python tools/dump_kernel.py filter_i64 [/tmp/out_prefix] [key=value ...]   then: cuobjdump -sass <prefix>.cubin
"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from queryengine_b200 import capi  # noqa: E402
from queryengine_b200.ast import AggregationFunction, DataType, Function, col, fn  # noqa: E402
from queryengine_b200.program import serialize  # noqa: E402

I, D = DataType.INT64, DataType.DOUBLE


def pipelines():
    a, b, c_, d = col(0), col(1), col(2), col(3)
    inner = fn(Function.DIV, fn(Function.MUL, fn(Function.ADD, a, b), fn(Function.SUB, c_, d)),
               fn(Function.ADD, fn(Function.MUL, a, 0.5), fn(Function.SUB, fn(Function.MOD, c_, 7.0), fn(Function.UNARY_MINUS, b))))
    tree = fn(Function.IF, fn(Function.CMP_GT, inner, 0.0),
              fn(Function.ADD, fn(Function.MUL, inner, fn(Function.SUB, a, d)), fn(Function.DIV, c_, fn(Function.ADD, b, 1.5))),
              fn(Function.SUB, fn(Function.UNARY_MINUS, inner), fn(Function.MUL, d, 0.25)))
    q6f = fn(Function.AND, fn(Function.AND, fn(Function.AND, fn(Function.AND,
             fn(Function.CMP_GE, col(3), 8766.0), fn(Function.CMP_LT, col(3), 9131.0)),
             fn(Function.CMP_GE, col(2), 0.05)), fn(Function.CMP_LE, col(2), 0.07)), fn(Function.CMP_LT, col(0), 24.0))
    return {
        "filter_i64": ([(I, False)], fn(Function.CMP_LT, col(0, I), float(2 ** 39)), [col(0, I)], []),
        "filter_f64": ([(D, False)], fn(Function.CMP_LT, col(0), float(2 ** 39)), [col(0)], []),
        "filter_f64_nullable": ([(D, True)], fn(Function.CMP_LT, col(0), float(2 ** 39)), [col(0)], []),
        "q6": ([(D, False)] * 4, q6f, [fn(Function.MUL, col(1), col(2))], [AggregationFunction.SUM]),
        "q6_filter": ([(D, False)] * 4, q6f, [fn(Function.MUL, col(1), col(2))], []),
        "tree": ([(D, True)] * 4, None, [tree], []),
    }


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "filter_i64"
    prefix = sys.argv[2] if len(sys.argv) > 2 else "/tmp/qe_" + name
    opts = dict(kv.split("=") for kv in sys.argv[3:])
    schema, filt, projs, aggs = pipelines()[name]
    L = capi.lib()
    descs = (capi.ColDesc * len(schema))(*[capi.ColDesc(int(t), 1 if n else 0) for t, n in schema])
    exprs = ([filt] if filt is not None else []) + list(projs)
    words = [serialize(e) for e in exprs]
    ptrs = (C.c_void_p * len(words))(*[w.ctypes.data for w in words])
    lens = (C.c_size_t * len(words))(*[w.size for w in words])
    agg = (C.c_int32 * max(1, len(aggs)))(*[int(x) for x in aggs])
    src = C.create_string_buffer(1 << 17)
    cub = C.create_string_buffer(1 << 21)
    nb = C.c_size_t(0)
    rc = L.qe_debug_generate(descs, len(schema), ptrs, lens, 1 if filt is not None else 0, len(projs), agg, len(aggs),
                             int(opts.get("steps", 0)), int(opts.get("vec", 0)), int(opts.get("staging", -1)), 1,
                             src, len(src), cub, len(cub), C.byref(nb))
    if rc:
        raise SystemExit(L.qe_last_error().decode())
    open(prefix + ".cu", "w").write(src.value.decode())
    open(prefix + ".cubin", "wb").write(cub.raw[:nb.value])
    print(prefix + ".cu", prefix + ".cubin", nb.value)


if __name__ == "__main__":
    main()
