"""Ahead-of-time kernels: compiles the kernels of the BASELINE.json pipelines (queryengine_b200/workloads.py) through the
library's own code generator and caches (qe_aot_compile, no GPU needed).  Run by queryengine_b200/build.py with
QE_CACHE_DIR pointing at a staging directory; prints one line `aot <key> <what>` per kernel, and build.py bakes the
cubins named there into libqegpu.so (csrc/qe_aot_table.cpp).

The reference's analogue is compile-once-per-expression (src/main/java/net/jhorstmann/queryengine/evaluator/
BytecodeCompiler.kt:134-175), which costs it milliseconds; NVRTC costs 0.2-0.7 s per kernel, so the kernels of the
pipelines the benchmarks run are made when the library is built.  Literal values are kernel parameters, so e.g. ONE
kernel serves `col < k` for every k.
"""
from __future__ import annotations

import ctypes as C
import sys

from . import capi, workloads as W
from .ast import DataType
from .program import serialize


def compile_one(what: str, schema, filt, projs, aggs=(), group_keys: int = -1, sel_hint: float = -1.0, min_blocks: int = 0,
                seen_groups: int = 0, run_flags: int = 0) -> str:
    L = capi.lib()
    cols = (capi.ColDesc * len(schema))(*[capi.ColDesc(int(t), 1 if nullable else 0) for _, t, nullable in schema])
    exprs = ([filt] if filt is not None else []) + list(projs)
    words = [serialize(e) for e in exprs]
    progs = (C.c_void_p * len(words))(*[w.ctypes.data for w in words])
    lens = (C.c_size_t * len(words))(*[w.size for w in words])
    fns = (C.c_int32 * max(1, len(aggs)))(*[int(a) for a in aggs])
    d = capi.AotDesc()
    d.cols, d.ncols = cols, len(schema)
    d.programs, d.nwords = progs, lens
    d.has_filter, d.nproj = (1 if filt is not None else 0), len(projs)
    d.agg_fns, d.nagg = fns, len(aggs)
    d.group_keys = group_keys
    d.sel_hint, d.min_blocks, d.seen_groups, d.run_flags = sel_hint, min_blocks, seen_groups, run_flags
    key = C.create_string_buffer(33)
    origin = C.c_int(0)
    capi.check(L.qe_aot_compile(C.byref(d), key, C.byref(origin)))
    print(f"aot {key.value.decode()} {what} (origin {origin.value})")
    return key.value.decode()


def main() -> None:
    from .ast import col
    for dtype in (DataType.INT64, DataType.DOUBLE):
        schema = W.single_column_schema(dtype)
        filt = W.single_column_filter(0.5, dtype)
        # unknown / sparse: 20 consumer warps + row lists + sweeper warps; 11-40 % kept last time: 20 warps, masks only;
        # above 40 %: 16 warps
        # (known to keep at most 8 %: 24 consumer warps, 42 KB tiles and a third sweeper warp)
        for hint in (-1.0, 0.01, 0.2, 0.99):
            for flags in (0, capi.QE_RUN_NO_VALUES):
                compile_one(f"configs[1] {dtype.name} filter hint={hint} flags={flags}", schema, filt, [col(0, dtype, "col")],
                            sel_hint=hint, run_flags=flags)
    compile_one("configs[2] Q6 filter + SUM", W.q6_schema(), W.q6_filter(), [W.q6_projection()], aggs=[2])
    for mb in (4, 0):
        compile_one(f"configs[3] nullable tree min_blocks={mb}", W.tree_schema(), None, [W.tree_expr()], min_blocks=mb)
    compile_one("configs[4] lineitem filter + 2 projections", W.lineitem_schema(), W.lineitem_filter(), W.lineitem_projections())
    filt, keys, inputs, fns = W.q1_groupby()
    for seen in (0, 6):
        compile_one(f"Q1 group-by, register variant, seen_groups={seen}", W.q1_schema(), filt, list(keys) + list(inputs),
                    aggs=fns, group_keys=len(keys), seen_groups=seen)


if __name__ == "__main__":
    main()
    sys.stdout.flush()
