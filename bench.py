#!/usr/bin/env python
"""Benchmark of the filter/scan hot path (BASELINE.json metric: filter-scan rows/s and HBM GB/s).

  python bench.py --gpus N --steps K --warmup W            our CUDA path
  python bench.py --impl reference ...                     the reference's CPU algorithm (oracle port) on host cores

Workload (BASELINE.json configs[1]): one int64 column of 1e9 rows per GPU,
`col < k` with k chosen for 1 % / 50 % / 99 % selectivity, output = selection
vector + compacted column.  One step = that three-filter sweep over the
resident column.  Inputs (8 GB per GPU) are far larger than L2, so nothing is
cached between launches.  Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

SELECTIVITIES = (0.01, 0.50, 0.99)
SEED = 42
METRIC = "filter_scan_rows_per_s"
UNIT = "rows/s"


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json, device copy)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def filter_expr(s: float):
    from queryengine_b200 import workloads
    return workloads.single_column_filter(s)


def splitmix64(x: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        x = x + np.uint64(0x9E3779B97F4A7C15)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return x ^ (x >> np.uint64(31))


def synthetic_column(n: int, offset: int = 0) -> np.ndarray:
    """Same values as qe_table_generate(QE_GEN_UNIFORM_U40, seed 42) for rows offset..offset+n."""
    i = np.arange(n, dtype=np.uint64) + np.uint64(offset)
    return (splitmix64(np.uint64(SEED) ^ i) >> np.uint64(24)).astype(np.int64)


# ------------------------------------------------------------------ clocks

class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU while the timed region runs."""
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        while not self._stop.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                mask = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in self.REASONS.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.004)

    def __enter__(self):
        if self.nv:
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thread:
            self._thread.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ------------------------------------------------------- CPU baseline legs

def cpu_leg(sample_rows: int, threads: int, repeats: int = 1):
    """The reference's row-at-a-time algorithm (oracle port) over a bounded sample of the same
    workload.  Returns (rows/s over the three-filter sweep, seconds, kept counts)."""
    from oracle_binding import OracleTable, timed_pipeline
    from queryengine_b200.ast import DataType, col
    vals = synthetic_column(sample_rows)
    t = OracleTable([DataType.INT64], sample_rows)
    t.set_column(0, vals)
    best = None
    kept = []
    for _ in range(repeats):
        secs = 0.0
        kept = []
        for s in SELECTIVITIES:
            dt, rows, _ = timed_pipeline(t, [0], filter_expr(s), [col(0, DataType.INT64)], [], threads)
            secs += dt
            kept.append(rows)
        best = secs if best is None else min(best, secs)
    return len(SELECTIVITIES) * sample_rows / best, best, kept


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    sample = args.cpu_rows
    cpu_leg(min(sample, 1_000_000), threads)                       # warm-up (page in the library)
    vals, times = [], []
    for _ in range(max(1, args.warmup)):
        cpu_leg(sample, threads)
    for _ in range(max(1, args.steps)):
        v, secs, kept = cpu_leg(sample, threads)
        vals.append(v)
        times.append(secs)
    value = float(np.mean(vals))
    single, _, _ = cpu_leg(max(1, sample // 8), 1)                  # the reference itself is single-threaded (Compiler.kt:13-18)
    cfg = workload_config(sample, args.gpus)
    cfg["rows_per_gpu_of_the_gpu_arm"] = args.rows
    cfg["note"] = ("rows_per_gpu is the bounded sample this arm really processes per step (same column generator, same "
                   "three-filter sweep); rows/s is what is compared")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": float(np.mean(times)) * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int64", "data": "synthetic",
        "config": cfg,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"{sample} rows of the same column, three-filter sweep per step; "
                                   "oracle/qe_oracle.c (C restatement of the Kotlin evaluator; no JVM in this image), "
                                   "row-range sharded over all host threads"},
        "single_thread_rows_per_s": single,
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def workload_config(rows: int, gpus: int):
    return {"workload": "configs[1]: single int64 column, `col < k`, selectivity sweep 1%/50%/99%, "
                        "output selection vector + compacted column",
            "rows_per_gpu": rows, "selectivities": list(SELECTIVITIES), "seed": SEED,
            "partitioning": f"row ranges, {gpus} partition(s), no data-path collective",
            "cache": "inputs (8 B/row) larger than L2; no flush needed"}


# ---------------------------------------------------------------- GPU arm

def host_memory_available() -> int:
    avail = None
    try:
        import psutil
        avail = psutil.virtual_memory().available
    except Exception:
        pass
    for p in ("/sys/fs/cgroup/memory.max", "/sys/fs/cgroup/memory/memory.limit_in_bytes"):
        try:
            v = open(p).read().strip()
            if v.isdigit():
                used = 0
                try:
                    used = int(open(os.path.join(os.path.dirname(p), "memory.current")).read())
                except Exception:
                    pass
                lim = int(v) - used
                avail = lim if avail is None else min(avail, lim)
        except Exception:
            pass
    return avail if avail is not None else 8 << 30


class PinnedBuffer:
    def __init__(self, nbytes: int):
        from queryengine_b200 import capi
        self.ptr = C.c_void_p()
        capi.check(capi.lib().qe_host_alloc(nbytes, C.byref(self.ptr)))
        self.nbytes = nbytes

    def free(self):
        from queryengine_b200 import capi
        if self.ptr:
            capi.lib().qe_host_free(self.ptr)
            self.ptr = C.c_void_p()


def timed_steps(torch, ctx, stream, fn_step, steps, dist, device):
    """K steps bracketed by barrier + synchronize, CUDA events on the library's stream; max over ranks."""
    ev0 = torch.cuda.Event(enable_timing=True)
    ev1 = torch.cuda.Event(enable_timing=True)
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    ev0.record(stream)
    for _ in range(steps):
        fn_step()
    ev1.record(stream)
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    ms = ev0.elapsed_time(ev1)
    if dist is not None:
        t = torch.tensor([ms], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms


def q6_pipeline(ctx, rows, row_offset=0):
    """BASELINE.json configs[2]: Q6-shaped filter + SUM(extendedprice * discount), four f64 columns."""
    from queryengine_b200 import capi
    from queryengine_b200.ast import AggregationFunction, DataType, Function, col, fn
    from queryengine_b200.engine import ColumnarTable, Field, Pipeline
    D = DataType.DOUBLE
    t = ColumnarTable(ctx, [Field("l_quantity", D), Field("l_extendedprice", D), Field("l_discount", D),
                            Field("l_shipdate", D)], rows)
    t.generate(0, capi.QE_GEN_UNIFORM_INT, SEED + 1, 1, 50, row_offset=row_offset)
    t.generate(1, capi.QE_GEN_CENTS, SEED + 2, 90000, 10499999, row_offset=row_offset)
    t.generate(2, capi.QE_GEN_CENTS, SEED + 3, 0, 10, row_offset=row_offset)
    t.generate(3, capi.QE_GEN_UNIFORM_INT, SEED + 4, 8036, 10561, row_offset=row_offset)
    from queryengine_b200 import workloads as W
    return t, Pipeline(t, W.q6_filter(), [W.q6_projection()], [AggregationFunction.SUM])


def extra_q6(ctx, torch, stream, rows, steps, warmup, peak):
    t, pipe = q6_pipeline(ctx, rows)
    res = pipe.run()
    for _ in range(warmup):
        pipe.run(res)
    ms0, runs0 = res.kernel_stats()
    for _ in range(steps):
        pipe.run(res)
    ms1, runs1 = res.kernel_stats()
    kms = (ms1 - ms0) / max(1, runs1 - runs0)
    part = res.agg_partial(0)
    out = {"rows": rows, "kernel_ms": kms, "rows_per_s": rows / (kms * 1e-3),
           "gb_per_s": rows * 32 / (kms * 1e-3) / 1e9, "frac_of_peak": rows * 32 / (kms * 1e-3) / 1e9 / peak,
           "algorithmic_bytes_per_row": 32, "selectivity": part.count / rows, "sum": res.agg(0), "jit": res.used_jit}
    res.close()
    t.close()
    return out


def sf_filter_pipeline(ctx, rows, row_offset=0, capacity_frac=0.30):
    """BASELINE.json configs[4]: multi-column AND/OR filter + two projections over lineitem-shaped columns
    (qty, extendedprice, discount, tax, shipdate; all f64), one row-range partition."""
    from queryengine_b200 import capi
    from queryengine_b200.ast import DataType, Function, col, fn
    from queryengine_b200.engine import ColumnarTable, Field, Pipeline
    D = DataType.DOUBLE
    t = ColumnarTable(ctx, [Field("l_quantity", D), Field("l_extendedprice", D), Field("l_discount", D), Field("l_tax", D),
                            Field("l_shipdate", D)], rows)
    t.generate(0, capi.QE_GEN_UNIFORM_INT, SEED + 1, 1, 50, row_offset=row_offset)
    t.generate(1, capi.QE_GEN_CENTS, SEED + 2, 90000, 10499999, row_offset=row_offset)
    t.generate(2, capi.QE_GEN_CENTS, SEED + 3, 0, 10, row_offset=row_offset)
    t.generate(3, capi.QE_GEN_CENTS, SEED + 5, 0, 8, row_offset=row_offset)
    t.generate(4, capi.QE_GEN_UNIFORM_INT, SEED + 4, 8036, 10561, row_offset=row_offset)
    from queryengine_b200 import workloads as W
    return t, Pipeline(t, W.lineitem_filter(), W.lineitem_projections(), out_capacity_rows=int(rows * capacity_frac) + 4096)


def extra_sf_filter(ctx, rows, steps, warmup, peak, row_offset=0):
    t, pipe = sf_filter_pipeline(ctx, rows, row_offset)
    res = pipe.run()
    for _ in range(warmup):
        pipe.run(res)
    ms0, runs0 = res.kernel_stats()
    for _ in range(steps):
        pipe.run(res)
    ms1, runs1 = res.kernel_stats()
    kms = (ms1 - ms0) / max(1, runs1 - runs0)
    kept = res.num_rows
    nbytes = rows * 40.0 + kept * 20.0
    out = {"rows": rows, "kept": kept, "selectivity": kept / rows, "kernel_ms": kms, "rows_per_s": rows / (kms * 1e-3),
           "gb_per_s": nbytes / (kms * 1e-3) / 1e9, "frac_of_peak": nbytes / (kms * 1e-3) / 1e9 / peak,
           "algorithmic_bytes_per_row": 40 + 20 * kept / rows, "jit": res.used_jit}
    res.close()
    t.close()
    return out


def tree_expr():
    """The depth-8 nullable arithmetic tree of BASELINE.json configs[3] (fixed in workloads.py, evaluated by the oracle in tests)."""
    from queryengine_b200 import workloads as W
    return W.tree_expr()


def q1_pipeline(ctx, rows, row_offset=0, atomic=False):
    """TPC-H Q1 shape: group by two low-cardinality STRING keys (dictionary codes), eight
    aggregates, a date filter that keeps ~96 % of the rows.  48 bytes per row are read."""
    from queryengine_b200 import capi
    from queryengine_b200.ast import AggregationFunction as A, DataType, Function, col, fn
    from queryengine_b200.engine import ColumnarTable, Field, GroupByPipeline
    D, S = DataType.DOUBLE, DataType.STRING
    t = ColumnarTable(ctx, [Field("l_returnflag", S), Field("l_linestatus", S), Field("l_quantity", D), Field("l_extendedprice", D),
                            Field("l_discount", D), Field("l_tax", D), Field("l_shipdate", D)], rows)
    codes = [ctx.dictionary.encode(v) for v in ("A", "N", "R", "F", "O")]              # the flags' dictionary codes
    assert codes[1] == codes[0] + 1 and codes[2] == codes[0] + 2 and codes[4] == codes[3] + 1
    t.generate(0, capi.QE_GEN_UNIFORM_INT, SEED + 7, codes[0], codes[2], row_offset=row_offset)
    t.generate(1, capi.QE_GEN_UNIFORM_INT, SEED + 8, codes[3], codes[4], row_offset=row_offset)
    t.generate(2, capi.QE_GEN_UNIFORM_INT, SEED + 1, 1, 50, row_offset=row_offset)
    t.generate(3, capi.QE_GEN_CENTS, SEED + 2, 90000, 10499999, row_offset=row_offset)
    t.generate(4, capi.QE_GEN_CENTS, SEED + 3, 0, 10, row_offset=row_offset)
    t.generate(5, capi.QE_GEN_CENTS, SEED + 5, 0, 8, row_offset=row_offset)
    t.generate(6, capi.QE_GEN_UNIFORM_INT, SEED + 4, 8036, 10561, row_offset=row_offset)
    from queryengine_b200 import workloads as W
    filt, keys, inputs, fns = W.q1_groupby()
    pipe = GroupByPipeline(t, filt, keys, inputs, fns, atomic=atomic)
    return t, pipe


def extra_q1(ctx, rows, steps, warmup, peak):
    from queryengine_b200 import capi
    t, pipe = q1_pipeline(ctx, rows)
    res = pipe.run()
    for _ in range(warmup):
        pipe.run(res)
    ms0, runs0 = res.kernel_stats()
    for _ in range(steps):
        pipe.run(res)
    ms1, runs1 = res.kernel_stats()
    kms = (ms1 - ms0) / max(1, runs1 - runs0)
    groups = res.num_rows
    counts = res.column(9)[0]
    out = {"rows": rows, "groups": groups, "kernel_ms": kms, "rows_per_s": rows / (kms * 1e-3),
           "gb_per_s": rows * 48 / (kms * 1e-3) / 1e9, "frac_of_peak": rows * 48 / (kms * 1e-3) / 1e9 / peak,
           "algorithmic_bytes_per_row": 48, "selectivity": float(counts.sum()) / rows,
           "variant": GROUP_VARIANTS[capi.lib().qe_result_group_variant(res.handle)],
           "threads_per_cta": capi.lib().qe_result_group_threads(res.handle)}
    res.close()
    t.close()
    return out


GROUP_VARIANTS = ["private", "atomic", "shared", "registers", "lanes", "lanes8", "warps", "twopass"]


def extra_groupby_cardinalities(ctx, rows, steps, warmup, pk_read, cards=(64, 1024, 65536)):
    """SELECT k, SUM(x), COUNT(x) GROUP BY k over 16-byte rows for a few group counts: which
    variant takes them and what it reaches of the device's read rate.  The first run of every cardinality walks the
    variants until one holds the groups; the timed runs reuse the result, which remembers."""
    from queryengine_b200 import capi
    from queryengine_b200.ast import AggregationFunction as A, DataType, col
    from queryengine_b200.engine import ColumnarTable, Field, GroupByPipeline
    D = DataType.DOUBLE
    t = ColumnarTable(ctx, [Field("k", D), Field("x", D)], rows)
    t.generate(1, capi.QE_GEN_UNIFORM_REAL, SEED + 11, 0.0, 1.0)
    out = []
    for card in cards:
        t.generate(0, capi.QE_GEN_UNIFORM_INT, SEED + 12, 0, card - 1)
        pipe = GroupByPipeline(t, None, [col(0)], [col(1), col(1)], [A.SUM, A.COUNT], max_groups=max(65536, 2 * card))
        res = pipe.run()
        for _ in range(warmup):
            pipe.run(res)
        ms0, runs0 = res.kernel_stats()
        for _ in range(steps):
            pipe.run(res)
        ms1, runs1 = res.kernel_stats()
        kms = (ms1 - ms0) / max(1, runs1 - runs0)
        out.append({"groups": res.num_rows, "kernel_ms": kms, "gb_per_s": rows * 16 / (kms * 1e-3) / 1e9,
                    "frac_of_read_peak": rows * 16 / (kms * 1e-3) / 1e9 / pk_read,
                    "variant": GROUP_VARIANTS[capi.lib().qe_result_group_variant(res.handle)],
                    "copies_per_cta": capi.lib().qe_result_group_copies(res.handle)})
        res.close()
    t.close()
    return {"rows": rows, "algorithmic_bytes_per_row": 16, "per_cardinality": out}


def extra_f64_column(ctx, rows, steps, warmup, peak):
    """configs[1] with the column held as f64 -- the reference's own numeric type (M/data/Schema.kt:3-5):
    the same values, the same `col < k` sweep, selection vector + compacted f64 column."""
    from queryengine_b200 import capi
    from queryengine_b200.ast import DataType, Function, col, fn
    from queryengine_b200.engine import ColumnarTable, Field, Pipeline
    D = DataType.DOUBLE
    t = ColumnarTable(ctx, [Field("col", D, False)], rows)
    t.generate(0, capi.QE_GEN_UNIFORM_U40, SEED)
    out = {"rows": rows, "per_selectivity": []}
    tot_ms = tot_bytes = 0.0
    for s_ in (0.01, 0.5, 0.99):
        pipe = Pipeline(t, fn(Function.CMP_LT, col(0, D, "col"), float(int(s_ * 2 ** 40))), [col(0, D, "col")])
        res = pipe.run()
        for _ in range(warmup):
            pipe.run(res)
        ms0, runs0 = res.kernel_stats()
        for _ in range(steps):
            pipe.run(res)
        ms1, runs1 = res.kernel_stats()
        kms = (ms1 - ms0) / max(1, runs1 - runs0)
        kept = res.num_rows
        nbytes = 8.0 * rows + 12.0 * kept
        out["per_selectivity"].append({"selectivity": s_, "kept": kept, "kernel_ms": kms, "gb_per_s": nbytes / (kms * 1e-3) / 1e9,
                                       "frac_of_peak": nbytes / (kms * 1e-3) / 1e9 / peak})
        tot_ms += kms
        tot_bytes += nbytes
        res.close()
    out["rows_per_s"] = 3 * rows / (tot_ms * 1e-3)
    out["gb_per_s"] = tot_bytes / (tot_ms * 1e-3) / 1e9
    out["frac_of_peak"] = out["gb_per_s"] / peak
    t.close()
    return out


def extra_tree(ctx, torch, stream, rows, steps, warmup, peak):
    """BASELINE.json configs[3]: nullable f64 arithmetic tree (depth 8) with validity bitmaps, dense projection."""
    from queryengine_b200 import capi
    from queryengine_b200.ast import DataType
    from queryengine_b200.engine import ColumnarTable, Field, Pipeline
    D = DataType.DOUBLE
    t = ColumnarTable(ctx, [Field("a", D, True), Field("b", D, True), Field("c", D, True), Field("d", D, True)], rows)
    for c, rate in enumerate((0.0, 0.01, 0.10, 0.50)):
        t.generate(c, capi.QE_GEN_UNIFORM_REAL, SEED + 10 + c, -1000.0, 1000.0, null_rate=rate)
    tree = tree_expr()
    pipe = Pipeline(t, None, [tree])
    res = pipe.run()
    for _ in range(warmup):
        pipe.run(res)
    ms0, runs0 = res.kernel_stats()
    for _ in range(steps):
        pipe.run(res)
    ms1, runs1 = res.kernel_stats()
    kms = (ms1 - ms0) / max(1, runs1 - runs0)
    bpr = 5 * (8 + 1 / 8)
    out = {"rows": rows, "kernel_ms": kms, "rows_per_s": rows / (kms * 1e-3), "gb_per_s": rows * bpr / (kms * 1e-3) / 1e9,
           "frac_of_peak": rows * bpr / (kms * 1e-3) / 1e9 / peak, "algorithmic_bytes_per_row": bpr, "jit": res.used_jit}
    res.close()
    t.close()
    return out


def hbm_yardsticks(ctx, nbytes=8 << 30, reps=5):
    """Pure read / 1:1 copy / pure write rates of this device, with the pipelines' own access shapes (qe_yardstick)."""
    from queryengine_b200 import capi
    out = {}
    for name, kind, moved in (("read", 0, 1), ("copy", 1, 2), ("write", 2, 1)):
        ms = C.c_float(0.0)
        capi.check(capi.lib().qe_yardstick(ctx.handle, kind, nbytes, reps, C.byref(ms)))
        out[name] = moved * (nbytes // 4096 * 4096) / (ms.value * 1e-3) / 1e9
    return out


def peak_for_mix(y, read_bytes, write_bytes):
    """Measured peak for a launch that reads read_bytes and writes write_bytes: time per byte interpolated between the
    yardsticks at write fractions 0 (read), 0.5 (copy) and 1 (write)."""
    f = write_bytes / max(1.0, read_bytes + write_bytes)
    if f <= 0.5:
        inv = (1 - 2 * f) / y["read"] + 2 * f / y["copy"]
    else:
        inv = (2 - 2 * f) / y["copy"] + (2 * f - 1) / y["write"]
    return 1.0 / inv


def hostlink_yardstick(torch, dist, device, up_bytes, down_bytes, chunk=64 << 20, buf=1 << 30):
    """What this box's host link does with the end-to-end leg's byte counts and no kernels: pinned buffers, plain
    cudaMemcpyAsync, H2D and D2H at once on two streams, every rank at the same time; max over ranks, milliseconds."""
    h_up = torch.empty(buf, dtype=torch.uint8, pin_memory=True)
    h_down = torch.empty(buf, dtype=torch.uint8, pin_memory=True)
    h_up.fill_(1)
    d_up = torch.empty(buf, dtype=torch.uint8, device=device)
    d_down = torch.ones(buf, dtype=torch.uint8, device=device)
    s_up, s_down = torch.cuda.Stream(device), torch.cuda.Stream(device)

    def run(up, down):
        e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        e[0].record(s_up); e[2].record(s_down)
        off = du = dd = 0
        while du < up or dd < down:
            n = min(chunk, buf - off)
            if du < up:
                with torch.cuda.stream(s_up):
                    d_up[off:off + n].copy_(h_up[off:off + n], non_blocking=True)
                du += n
            if dd < down:
                with torch.cuda.stream(s_down):
                    h_down[off:off + n].copy_(d_down[off:off + n], non_blocking=True)
                dd += n
            off = (off + n) % buf
        e[1].record(s_up); e[3].record(s_down)
        torch.cuda.synchronize()
        ms = max(e[0].elapsed_time(e[1]), e[2].elapsed_time(e[3]))
        if dist is not None:
            t = torch.tensor([ms], dtype=torch.float64, device=device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms
    run(min(up_bytes, buf), min(down_bytes, buf))
    ms = run(up_bytes, down_bytes)
    del h_up, h_down, d_up, d_down
    return ms


def hostlink_same_buffers(torch, dist, device, h_in, in_bytes, outs, piece=64 << 20):
    """The end-to-end step's bytes as plain cudaMemcpyAsync calls out of / into the SAME pinned host buffers the step uses
    (distinct host memory, not a recycled buffer), H2D and D2H at once on two streams, no kernels: what the link does for
    exactly this traffic.  outs = [(pinned buffer, bytes)]; max over ranks, milliseconds, or None without libcudart."""
    import glob
    rt = None
    for cand in glob.glob(os.path.join(os.path.dirname(torch.__file__), "lib", "libcudart*.so*")) + ["libcudart.so.12", "libcudart.so"]:
        try:
            rt = C.CDLL(cand)
            break
        except OSError:
            continue
    if rt is None:
        return None
    rt.cudaMemcpyAsync.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
    rt.cudaMemcpyAsync.restype = C.c_int
    win = 1 << 30
    d_up = torch.empty(win, dtype=torch.uint8, device=device)
    d_down = torch.ones(win, dtype=torch.uint8, device=device)
    s_up, s_down = torch.cuda.Stream(device), torch.cuda.Stream(device)

    def run():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ups = [(off, min(piece, in_bytes - off)) for off in range(0, in_bytes, piece)]
        downs = [(b, off, min(piece, nb - off)) for b, nb in outs for off in range(0, nb, piece)]
        for i in range(max(len(ups), len(downs))):                 # interleaved, as the step issues them
            if i < len(ups):
                off, m = ups[i]
                assert rt.cudaMemcpyAsync(d_up.data_ptr() + off % (win - piece), h_in.ptr.value + off, m, 1, s_up.cuda_stream) == 0
            if i < len(downs):
                b, off, m = downs[i]
                assert rt.cudaMemcpyAsync(b.ptr.value + off, d_down.data_ptr() + off % (win - piece), m, 2, s_down.cuda_stream) == 0
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) * 1e3
        if dist is not None:
            t = torch.tensor([ms], dtype=torch.float64, device=device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms
    run()
    ms = min(run(), run())
    del d_up, d_down
    return ms


def cold_start_leg(ctx, table, rows):
    """First-query latency of this process (the caches as they are on this box) and of a pipeline that is in no
    cache (NVRTC), plus a fresh process that finds the latter in the disk cache."""
    import subprocess
    from queryengine_b200.ast import DataType, Function, col, fn
    from queryengine_b200.engine import Pipeline, jit_stats
    I64 = DataType.INT64
    out = {}
    s0 = jit_stats()
    t0 = time.perf_counter()
    odd = fn(Function.CMP_LT, fn(Function.SUB, col(0, I64, "col"), 12345.0), float(int(0.25 * 2 ** 40)))      # not a BASELINE pipeline
    p = Pipeline(table, odd, [col(0, I64, "col")])
    r = p.run()
    n1 = r.num_rows
    out["uncached_pipeline_first_result_ms"] = (time.perf_counter() - t0) * 1e3
    s1 = jit_stats()
    out["uncached_pipeline_origin"] = {k: s1[k] - s0[k] for k in ("aot", "disk", "nvrtc")}
    out["uncached_pipeline_nvrtc_ms"] = s1["nvrtc_ms"] - s0["nvrtc_ms"]
    t0 = time.perf_counter()
    r = Pipeline(table, fn(Function.CMP_LT, fn(Function.SUB, col(0, I64, "col"), 999.0), float(int(0.5 * 2 ** 40))), [col(0, I64, "col")]).run(r)
    n2 = r.num_rows
    out["same_shape_other_literals_ms"] = (time.perf_counter() - t0) * 1e3
    out["kernel_ms"] = r.kernel_ms
    r.close()
    # a fresh process: context creation, then the same two queries (ahead-of-time table / disk cache)
    try:
        o = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "cold_start.py"), "--rows", str(1 << 24), "--skip-interpreter"],
                           capture_output=True, text=True, timeout=300)
        out["fresh_process"] = json.loads(o.stdout.strip().splitlines()[-1])
    except Exception as e:
        out["fresh_process"] = {"error": repr(e)}
    return out


def interpreter_leg(table, rows, y, steps):
    """configs[1] on the always-available interpreter kernels (what a deployment without NVRTC runs for a pipeline
    that is neither in the ahead-of-time table nor in the disk cache)."""
    from queryengine_b200.ast import DataType, col
    from queryengine_b200.engine import Pipeline
    I64 = DataType.INT64
    out = {"per_selectivity": []}
    tot_ms = tot_b = 0.0
    for s_ in SELECTIVITIES:
        p = Pipeline(table, filter_expr(s_), [col(0, I64, "col")], interpreter=True)
        r = p.run()
        r = p.run(r)
        m0, n0 = r.kernel_stats()
        for _ in range(max(2, min(steps, 3))):
            r = p.run(r)
        m1, n1 = r.kernel_stats()
        kms = (m1 - m0) / max(1, n1 - n0)
        k = r.num_rows
        nbytes = 8.0 * rows + 12.0 * k
        pk = peak_for_mix(y, 8.0 * rows, 12.0 * k)
        out["per_selectivity"].append({"selectivity": s_, "kept": k, "kernel_ms": kms, "gb_per_s": nbytes / (kms * 1e-3) / 1e9,
                                       "frac_of_peak_mix": nbytes / (kms * 1e-3) / 1e9 / pk, "used_jit": r.used_jit})
        tot_ms += kms
        tot_b += nbytes
        r.close()
    out["gb_per_s"] = tot_b / (tot_ms * 1e-3) / 1e9
    out["frac_of_copy_peak"] = out["gb_per_s"] / y["copy"]
    return out


def run_ours(args):
    import torch
    from queryengine_b200 import capi
    from queryengine_b200.ast import DataType, col
    from queryengine_b200.engine import ColumnarTable, Context, Field, Pipeline, jit_stats

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the hot path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    all_cpus = os.sched_getaffinity(0)
    numa = bind_to_gpu_numa_node(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        dist_mod.init_process_group("nccl", device_id=device)
        dist = dist_mod
    peak, peak_kind = measured_peak_gbs()

    t_ctx = time.perf_counter()
    ctx = Context(local_rank)
    ctx_ms = (time.perf_counter() - t_ctx) * 1e3
    if args.steps_opt:
        ctx.set_option("steps", args.steps_opt)
    if args.vec_opt:
        ctx.set_option("vec", args.vec_opt)
    if args.grid_per_sm:
        ctx.set_option("grid_per_sm", args.grid_per_sm)
    stream = torch.cuda.ExternalStream(ctx.stream, device=device)
    rows = args.rows
    I64 = DataType.INT64
    table = ColumnarTable(ctx, [Field("col", I64, False)], rows)
    table.generate(0, capi.QE_GEN_UNIFORM_U40, SEED, row_offset=rank * rows)
    ctx.synchronize()
    pipes = [Pipeline(table, filter_expr(s), [col(0, I64, "col")]) for s in SELECTIVITIES]
    # first query of this process: pipeline object -> first result on the host (kernel from the ahead-of-time table)
    t_first = time.perf_counter()
    first = pipes[1].run()
    first_rows = first.num_rows
    first_query_ms = (time.perf_counter() - t_first) * 1e3
    first_stats = jit_stats()
    results = [pipes[0].run(), first, pipes[2].run()]
    kept = [r.num_rows for r in results]
    for s, k in zip(SELECTIVITIES, kept):
        assert abs(k / rows - s) < 0.01, f"selectivity {s}: kept {k} of {rows}"
    assert all(r.used_jit for r in results) or args.allow_interpreter, "generated code did not run (NVRTC missing?)"

    def step():
        for i, p in enumerate(pipes):
            results[i] = p.run(results[i])

    for _ in range(max(3, args.warmup)):
        step()
    ctx.synchronize()
    stats0 = [r.kernel_stats() for r in results]
    launches0 = ctx.launch_count
    with ClockSampler(local_rank) as clocks:
        ms = timed_steps(torch, ctx, stream, step, args.steps, dist, device)
    launches = ctx.launch_count - launches0
    stats1 = [r.kernel_stats() for r in results]
    total_rows = world * rows * len(SELECTIVITIES) * args.steps
    value = total_rows / (ms * 1e-3)

    # same-box yardsticks, measured now: pure read / copy / pure write with the kernels' own access shapes
    yard = hbm_yardsticks(ctx) if rank == 0 or world > 1 else None

    # roofline of the filter + compaction kernel: algorithmic bytes = 8 B read per row + (8 + 4) B written per kept row
    per_sel = []
    tot_bytes = tot_ms = tot_min_ms = 0.0
    for s, k, (m0, n0), (m1, n1) in zip(SELECTIVITIES, kept, stats0, stats1):
        kms = (m1 - m0) / max(1, n1 - n0)
        nbytes = 8.0 * rows + 12.0 * k
        pk = peak_for_mix(yard, 8.0 * rows, 12.0 * k)
        per_sel.append({"selectivity": s, "kept": k, "kernel_ms": kms, "rows_per_s": rows / (kms * 1e-3),
                        "gb_per_s": nbytes / (kms * 1e-3) / 1e9, "frac_of_peak": nbytes / (kms * 1e-3) / 1e9 / peak,
                        "peak_mix_gb_per_s": pk, "frac_of_peak_mix": nbytes / (kms * 1e-3) / 1e9 / pk,
                        "algorithmic_bytes": nbytes})
        tot_bytes += nbytes
        tot_ms += kms
        tot_min_ms += nbytes / pk / 1e6
    achieved = tot_bytes / (tot_ms * 1e-3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        try:
            tj = json.load(open(tp))
            traffic = tj.get("filter_sweep_dram_bytes_per_launch")
        except Exception:
            traffic = None
    roofline = {"bound": "hbm", "kernel": "qe_filter_ring_kernel (generated filter + compaction, body qe_filter_ring_body)", "achieved": achieved,
                "peak": peak, "peak_kind": peak_kind, "unit": "GB/s", "frac": achieved / peak,
                "frac_of_nominal_8000": achieved / 8000.0, "traffic": traffic,
                "traffic_source": "profiles/traffic.json: dram__bytes_read.sum + dram__bytes_write.sum of the ncu --set full capture named there",
                "peak_mix": tot_bytes / (tot_min_ms * 1e-3) / 1e9, "frac_of_peak_mix": tot_min_ms / tot_ms,
                "yardsticks_gb_per_s": yard,
                "peak_mix_kind": "this run's qe_yardstick read / copy / write rates, time per byte interpolated by the launch's write fraction",
                "algorithmic_bytes_per_launch": tot_bytes / len(SELECTIVITIES),
                "avg_launch_ms": tot_ms / len(SELECTIVITIES), "per_selectivity": per_sel}

    cold = interp = None
    if rank == 0 and not args.no_extra:
        try:
            cold = cold_start_leg(ctx, table, rows)
        except Exception as e:
            cold = {"error": repr(e)}
        cold["ctx_create_ms"] = ctx_ms
        cold["first_query_ms"] = first_query_ms
        cold["first_query_kernel_origin"] = {k: first_stats[k] for k in ("aot", "disk", "nvrtc")}
        try:
            interp = interpreter_leg(table, rows, yard, args.steps)
        except Exception as e:
            interp = {"error": repr(e)}

    for r in results:
        r.close()
    table.close()
    ctx.trim()

    # ---- end to end through the C ABI with host buffers: upload -> filter -> download, every pass
    e2e = run_e2e(args, torch, ctx, dist, device, rows, world, rank, kept)

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int64", "data": "synthetic", "config": workload_config(rows, world),
        "roofline": roofline, "clocks": clocks.summary(), "e2e": dict(e2e, cpus_bound_near_gpu=numa), "gpu_launches": int(launches),
        "hbm_gb_per_s_aggregate": world * len(SELECTIVITIES) * args.steps * (tot_bytes / len(SELECTIVITIES)) / (ms * 1e-3) / 1e9,
        "first_query_ms": first_query_ms,
    }
    if cold is not None:
        line["cold_start"] = cold
    if interp is not None:
        line["interpreter"] = interp

    extra = {}
    if rank == 0 and world == 1 and not args.no_extra:
        pk_read = yard["read"]
        for name, fn_, n in (("q6_sf100", extra_q6, args.q6_rows), ("nullable_tree", extra_tree, args.tree_rows)):
            try:
                extra[name] = fn_(ctx, torch, stream, n, max(3, args.steps), 3, peak)
            except Exception as e:  # reported, never hidden
                extra[name] = {"error": repr(e)}
        if "gb_per_s" in extra.get("q6_sf100", {}):
            extra["q6_sf100"]["frac_of_peak_mix"] = extra["q6_sf100"]["gb_per_s"] / pk_read          # a pure read
        if "gb_per_s" in extra.get("nullable_tree", {}):
            extra["nullable_tree"]["frac_of_peak_mix"] = extra["nullable_tree"]["gb_per_s"] / peak_for_mix(yard, 4 * 8.125, 8.125)
        ctx.trim()
        try:      # configs[1] with the column held as f64, the reference's own type
            extra["single_f64_column_1e9"] = extra_f64_column(ctx, args.rows, max(3, args.steps), 3, peak)
        except Exception as e:
            extra["single_f64_column_1e9"] = {"error": repr(e)}
        ctx.trim()
        try:      # configs[4] on one partition (SF1000 / 6: what one of 6 GPUs would hold)
            extra["lineitem_filter_2proj_1e9"] = extra_sf_filter(ctx, args.sf_rows, max(3, args.steps), 3, peak)
            e_ = extra["lineitem_filter_2proj_1e9"]
            e_["frac_of_peak_mix"] = e_["gb_per_s"] / peak_for_mix(yard, 40.0 * e_["rows"], 20.0 * e_["kept"])
        except Exception as e:
            extra["lineitem_filter_2proj_1e9"] = {"error": repr(e)}
        ctx.trim()
        try:      # group-by, TPC-H Q1 shape over SF100 lineitem
            extra["q1_groupby_sf100"] = extra_q1(ctx, args.q6_rows, max(3, args.steps), 3, peak)
            extra["q1_groupby_sf100"]["frac_of_peak_mix"] = extra["q1_groupby_sf100"]["gb_per_s"] / pk_read
        except Exception as e:
            extra["q1_groupby_sf100"] = {"error": repr(e)}
        ctx.trim()
        try:      # group-by beyond a handful of groups: the shared-memory copies and the two-pass variant
            extra["groupby_sum_count_by_k"] = extra_groupby_cardinalities(ctx, args.q6_rows, max(3, args.steps), 2, pk_read)
        except Exception as e:
            extra["groupby_sum_count_by_k"] = {"error": repr(e)}
        ctx.trim()
        try:      # configs[4] at its FULL size on one GPU: 5 999 989 709 rows through a bounded device window
            extra["lineitem_sf1000_streamed_one_gpu"] = extra_sf1000_streamed(ctx, 5_999_989_709, peak, yard)
        except Exception as e:
            extra["lineitem_sf1000_streamed_one_gpu"] = {"error": repr(e)}
        ctx.trim()
    if world > 1 and not args.no_extra:
        extra.update(sharded_extras(args, torch, ctx, dist, device, rank, world, peak, yard))
    os.sched_setaffinity(0, all_cpus)          # the CPU baseline below uses every host core
    if rank == 0 and not args.no_cpu:
        threads = os.cpu_count() or 1
        v, secs, ckept = cpu_leg(args.cpu_rows, threads)
        line["cpu_baseline"] = {
            "value": v, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"{args.cpu_rows} rows of the same column, one three-filter sweep ({secs:.2f} s); oracle/qe_oracle.c, "
                      "the C restatement of the reference's Kotlin evaluator (no JVM in this image), row-range sharded"}
        v1, secs1, _ = cpu_leg(max(1, args.cpu_rows // 8), 1)
        line["cpu_baseline"]["single_thread_rows_per_s"] = v1      # the reference itself is single-threaded
    if world > 1 and not args.no_extra:
        try:
            extra["one_process_contexts"] = one_process_leg(args, torch, dist, rank, world)
        except Exception as e:
            extra["one_process_contexts"] = {"error": repr(e)}
    if extra:
        line["extra"] = extra
    ctx.close()
    if rank == 0:
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


def extra_sf1000_streamed(ctx, total, peak, yard, chunk=1 << 27):
    """configs[4] at full size through qe_pipeline_stream's bounded mode: the five lineitem columns are generated on the
    device chunk by chunk (no host could hold 240 GB), filtered and projected; selection vectors are chunk-relative with a
    64-bit chunk base (more than 2^32 rows).  Outputs are counted, not downloaded.  Device time of the filter kernels."""
    from queryengine_b200 import capi, workloads as W
    from queryengine_b200.engine import ColumnarTable, Field, Pipeline, stream_pipelines
    fields = [Field(n, t, nl) for n, t, nl in W.lineitem_schema()]
    window = ColumnarTable(ctx, fields, 3 * chunk)
    gen = [(capi.QE_GEN_UNIFORM_INT, SEED + 1, 1, 50), (capi.QE_GEN_CENTS, SEED + 2, 90000, 10499999), (capi.QE_GEN_CENTS, SEED + 3, 0, 10),
           (capi.QE_GEN_CENTS, SEED + 5, 0, 8), (capi.QE_GEN_UNIFORM_INT, SEED + 4, 8036, 10561)]
    pipe = Pipeline(window, W.lineitem_filter(), W.lineitem_projections())
    nchunks = -(-total // chunk)
    kept = np.zeros(nchunks, np.int64)
    t0 = time.perf_counter()
    res = stream_pipelines(window, [pipe], [None] * 5, chunk_rows=chunk, total_rows=total, chunk_relative=True, gen=gen, chunk_kept=[kept])
    wall = time.perf_counter() - t0
    r = res[0]
    n_kept = r.num_rows
    ms, runs = r.kernel_stats()
    assert runs == nchunks and kept[-1] == n_kept and (np.diff(kept) >= 0).all()
    nbytes = 40.0 * total + 20.0 * n_kept
    out = {"rows": total, "kept": n_kept, "selectivity": n_kept / total, "chunks": int(nchunks), "chunk_rows": chunk,
           "device_window_rows": 3 * chunk, "filter_kernel_ms_total": ms, "rows_per_s_kernels": total / (ms * 1e-3),
           "gb_per_s": nbytes / (ms * 1e-3) / 1e9, "frac_of_peak": nbytes / (ms * 1e-3) / 1e9 / peak,
           "frac_of_peak_mix": nbytes / (ms * 1e-3) / 1e9 / peak_for_mix(yard, 40.0 * total, 20.0 * n_kept),
           "wall_s_including_generation": wall, "selection": "chunk-relative uint32 + 64-bit chunk base (chunk_kept)"}
    r.close()
    window.close()
    return out


def sharded_extras(args, torch, ctx, dist, device, rank, world, peak, yard):
    """Row-sharded configs at N > 1 (one process per GPU): wall time includes the exchange of the partials and the combine."""
    from queryengine_b200 import sharding
    from queryengine_b200.ast import AggregationFunction
    extra = {}
    A_ = sharding.ALIGN_SUM                                    # partition boundaries at super-tile multiples: SUM independent of N

    def wall_max(fn_, reps):
        fn_()                                                  # warm
        dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            out = fn_()
        secs = (time.perf_counter() - t0) / reps
        t = torch.tensor([secs], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) * 1e3, out

    # Q6 over SF100 lineitem: per-rank super-tile partials, NCCL all-gather, fold in row order on every rank
    lo, hi = sharding.partition(args.q6_rows, world, rank, A_)
    t6, pipe6 = q6_pipeline(ctx, hi - lo, row_offset=lo)
    res6 = pipe6.run()
    state = {"res": res6}

    def q6_query():
        state["res"] = pipe6.run(state["res"])
        return sharding.sharded_aggregate(state["res"], [AggregationFunction.SUM], device=device)
    for _ in range(3):
        pipe6.run(res6)
    m0, n0 = res6.kernel_stats()
    wall_ms, total = wall_max(q6_query, max(3, args.steps))
    m1, n1 = state["res"].kernel_stats()
    kms = torch.tensor([(m1 - m0) / max(1, n1 - n0)], dtype=torch.float64, device=device)
    dist.all_reduce(kms, op=dist.ReduceOp.MAX)
    extra["q6_sf100_sharded"] = {"rows": args.q6_rows, "partitions": world, "kernel_ms_max_over_ranks": float(kms.item()),
                                 "wall_ms": wall_ms, "rows_per_s": args.q6_rows / (wall_ms * 1e-3),
                                 "rows_per_s_kernel_only": args.q6_rows / (float(kms.item()) * 1e-3),
                                 "gb_per_s_aggregate_kernel_only": args.q6_rows * 32 / (float(kms.item()) * 1e-3) / 1e9,
                                 "sum": total[0], "sum_hex": float(total[0]).hex(),
                                 "combine": "launch + NCCL all_gather of the super-tile partials + qe_agg_combine in row order; wall_ms "
                                            "is one whole query, max over ranks; sum has the same bits for every N"}
    state["res"].close()
    t6.close()
    ctx.trim()
    # TPC-H Q1 shape row-sharded: every rank groups its rows, fixed-width group records are all-gathered as tensors
    # and merged by key by qe_groupby_merge
    try:
        from queryengine_b200.ast import AggregationFunction as _A
        t1, pipe1 = q1_pipeline(ctx, hi - lo, row_offset=lo)
        res1 = pipe1.run()
        fns1 = [_A.SUM, _A.SUM, _A.SUM, _A.SUM, _A.AVG, _A.AVG, _A.AVG, _A.COUNT]
        st1 = {"res": res1}

        def q1_query():
            st1["res"] = pipe1.run(st1["res"])
            return sharding.sharded_groupby(st1["res"], 2, fns1, ctx.dictionary, lo, device=device)
        for _ in range(3):
            pipe1.run(res1)
        m0, n0 = res1.kernel_stats()
        wall_ms, rows1 = wall_max(q1_query, max(3, args.steps))
        m1, n1 = st1["res"].kernel_stats()
        kms = torch.tensor([(m1 - m0) / max(1, n1 - n0)], dtype=torch.float64, device=device)
        dist.all_reduce(kms, op=dist.ReduceOp.MAX)
        extra["q1_groupby_sf100_sharded"] = {
            "rows": args.q6_rows, "partitions": world, "groups": len(rows1), "kernel_ms_max_over_ranks": float(kms.item()),
            "wall_ms": wall_ms, "rows_per_s": args.q6_rows / (wall_ms * 1e-3),
            "rows_per_s_kernel_only": args.q6_rows / (float(kms.item()) * 1e-3),
            "count_total": int(sum(r[-1] for r in rows1)),
            "combine": "all_gather of fixed-width group records (tensors) + qe_groupby_merge; wall_ms is one whole query"}
        st1["res"].close()
        t1.close()
    except Exception as e:
        extra["q1_groupby_sf100_sharded"] = {"error": repr(e)}
    ctx.trim()
    # configs[4]: SF1000 lineitem (5 999 989 709 rows) row-sharded over the ranks, no collective at all
    total = 5_999_989_709
    lo, hi = sharding.partition(total, world, rank, A_)
    free_b, _tot = torch.cuda.mem_get_info()
    fit = int(free_b * 0.92 / (40 + 0.30 * 20 + 0.1)) // 4096 * 4096      # rows this GPU can hold with its outputs
    try:
        if hi - lo <= min(fit, 0xfffff000):
            e = extra_sf_filter(ctx, hi - lo, max(3, args.steps), 3, peak, row_offset=lo)
            e["how"] = "resident"
        else:                                                  # larger than this GPU: the bounded window (generated chunk by chunk)
            e = extra_sf1000_streamed(ctx, hi - lo, peak, yard)
            e["kernel_ms"] = e["filter_kernel_ms_total"]
            e["how"] = "bounded stream"
        v = torch.tensor([e["kernel_ms"], float(hi - lo), float(e["kept"])], dtype=torch.float64, device=device)
        vmax = v.clone(); dist.all_reduce(vmax, op=dist.ReduceOp.MAX)
        vsum = v.clone(); dist.all_reduce(vsum, op=dist.ReduceOp.SUM)
        rows_all, kept_all, kms = float(vsum[1].item()), float(vsum[2].item()), float(vmax[0].item())
        extra["lineitem_sf1000_sharded"] = {
            "rows_total_sf1000": total, "rows_processed": int(rows_all), "partitions": world, "kept": int(kept_all),
            "reduced": rows_all < total, "how": e["how"], "kernel_ms_max_over_ranks": kms, "rows_per_s": rows_all / (kms * 1e-3),
            "gb_per_s_aggregate": (rows_all * 40 + kept_all * 20) / (kms * 1e-3) / 1e9,
            "frac_of_aggregate_peak": (rows_all * 40 + kept_all * 20) / (kms * 1e-3) / 1e9 / (peak * world)}
    except Exception as e:
        extra["lineitem_sf1000_sharded"] = {"error": repr(e)}
    ctx.trim()
    return extra


def one_process_leg(args, torch, dist, rank, world):
    """One process driving all GPUs through qe_mgpu_* (what a JVM adapter does; the other ranks wait): the 50 % filter of
    configs[1] over a table sharded across the devices, checked against the single-device result, and its end-to-end
    time with host buffers (every device streams its row range at the same time)."""
    import torch as _t
    dist.barrier()
    out = None
    if rank == 0:
        from queryengine_b200 import capi
        from queryengine_b200.ast import AggregationFunction, DataType, col
        from queryengine_b200.engine import ColumnarTable, Context, Field, Pipeline
        from queryengine_b200.multigpu import MultiGpu, ShardedPipeline, ShardedTable
        I64 = DataType.INT64
        n = min(args.rows, 1 << 28) * world                      # rows of the whole table
        n = n // (1 << 20) * (1 << 20)
        mg = MultiGpu(list(range(world)))
        st = ShardedTable(mg, [Field("col", I64, False)], n)
        st.generate(0, capi.QE_GEN_UNIFORM_U40, SEED)
        sp = ShardedPipeline(st, filter_expr(0.5), [col(0, I64, "col")], out_capacity_rows=int((n // world) * 0.52) + (1 << 20))
        res = sp.run()
        kept = res.num_rows
        t0 = time.perf_counter()
        for _ in range(3):
            res = sp.run(res)
            kept = res.num_rows
        run_ms = (time.perf_counter() - t0) / 3 * 1e3
        # rank-order concatenation equals the single-device result: per-partition counts, first / last selected row, and a
        # checksum of the compacted column against the same filter run on one device over the same global rows
        per_part = [res.part_rows(p) for p in range(world)]
        ctx0 = mg.contexts[0]
        check_rows = min(n, 1 << 28)
        t1 = ColumnarTable(ctx0, [Field("col", I64, False)], check_rows)
        t1.generate(0, capi.QE_GEN_UNIFORM_U40, SEED)
        r1 = Pipeline(t1, filter_expr(0.5), [col(0, I64, "col")]).run()
        k1 = r1.num_rows
        import ctypes as C_
        same = True
        if check_rows == st.ranges[0][1]:
            same = k1 == per_part[0]
        sum_agg = ShardedPipeline(st, filter_expr(0.5), [col(0, I64, "col")], [AggregationFunction.SUM]).run()
        total_sum = sum_agg.agg(0)
        sum_agg.close()
        r1.close()
        t1.close()
        assert same, "partition 0 of the sharded run differs from the single-device run"
        # end to end: host column in, selection vector + compacted column out, all devices at once
        h_in = PinnedBuffer(n * 8)
        caps = [int((e - b) * 0.52) + 4096 for b, e in st.ranges]
        offs = [sum(caps[:p]) for p in range(world)]
        h_sel, h_val = PinnedBuffer(sum(caps) * 4), PinnedBuffer(sum(caps) * 8)
        # the host copy of the column: read back from the devices partition by partition (untimed)
        for p, (b, e) in enumerate(st.ranges):
            dense = Pipeline(st.parts[p], None, [col(0, I64, "col")]).run()
            capi.check(capi.lib().qe_result_column(dense.handle, 0, C_.c_void_p(h_in.ptr.value + 8 * b), None))
            dense.close()
        results = None
        for rep in range(3):
            if rep == 1:
                t0 = time.perf_counter()
            results = sp.stream([h_in.ptr.value], [8], args.e2e_chunk_rows, h_sel.ptr.value, [h_val.ptr.value], offs, results=results)
            got = [r.num_rows for r in results]
        e2e_ms = (time.perf_counter() - t0) / 2 * 1e3
        assert got == per_part, (got, per_part)
        first_sel = C_.cast(h_sel.ptr, C_.POINTER(C_.c_uint32))[0]
        first_val = C_.cast(h_val.ptr, C_.POINTER(C_.c_int64))[0]
        src_val = C_.cast(h_in.ptr, C_.POINTER(C_.c_int64))[first_sel]
        assert first_val == src_val and first_val < int(0.5 * 2 ** 40)
        out = {"devices": world, "rows": n, "kept": kept, "kept_per_partition": per_part, "run_ms_resident": run_ms,
               "rows_per_s_resident": n / (run_ms * 1e-3), "e2e_ms": e2e_ms, "e2e_rows_per_s": n / (e2e_ms * 1e-3),
               "h2d_bytes": n * 8, "d2h_bytes": kept * 12, "sum_of_kept_column": total_sum,
               "checks": "partition 0 equals the single-device run over the same rows; streamed selection / values consistent with the host column; "
                         "per-partition counts of resident and streamed runs equal",
               "what": "qe_mgpu_create + qe_mtable_* + qe_mgpu_pipeline_run / qe_mgpu_pipeline_stream: one process, one context and one worker thread per GPU"}
        for r in results:
            r.close()
        res.close()
        st.close()
        mg.close()
        for b_ in (h_in, h_sel, h_val):
            b_.free()
    dist.barrier()
    return out


def bind_to_gpu_numa_node(index):
    """Pin this rank's threads to the CPUs next to its GPU (NVML's ideal affinity) before any pinned host buffer is
    allocated, so that the end-to-end leg's staging memory sits on the GPU's NUMA node: with one rank per GPU the
    copies of eight ranks otherwise cross the socket interconnect.  Returns the CPU count bound to, or None."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1}
        cpus &= set(os.sched_getaffinity(0))
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return None


def run_e2e(args, torch, ctx, dist, device, rows, world, rank, kept_full):
    """The same three-filter sweep through the C ABI with HOST buffers: one qe_pipeline_stream call per step takes the
    column from pinned host memory (uploaded once per step, in chunks, into a bounded window of three chunk slots on the
    device), runs the three filters chunk by chunk and lands every selection vector and compacted column in pinned host
    memory; uploads, kernels and downloads of neighbouring chunks overlap.  Second figure: the same sweep with the
    selection vectors as the whole output (QE_RUN_NO_VALUES) -- the table is the host's own, FilterOperator hands back
    the row it was given."""
    from queryengine_b200 import capi
    from queryengine_b200.ast import DataType, col
    from queryengine_b200.engine import ColumnarTable, Field, Pipeline, stream_pipelines
    I64 = DataType.INT64
    e2e_rows = min(rows, args.e2e_rows) if args.e2e_rows else rows
    frac = sum(k / rows for k in kept_full)                           # kept rows per input row over the sweep
    bytes_per_row = 8 + 12 * frac * 1.02 + 1
    budget = int(host_memory_available() * 0.6) // max(1, min(world, 8))
    e2e_rows = max(1 << 20, min(e2e_rows, int(budget / bytes_per_row) // (1 << 20) * (1 << 20)))
    caps = [int(k / rows * e2e_rows * 1.02) + 4096 for k in kept_full]
    bufs = []
    try:
        h_in = PinnedBuffer(e2e_rows * 8); bufs.append(h_in)
        h_sel = [PinnedBuffer(c * 4) for c in caps]; bufs += h_sel
        h_val = [PinnedBuffer(c * 8) for c in caps]; bufs += h_val
    except capi.QeError:
        for b in bufs:
            b.free()
        return {"value": None, "unit": UNIT, "error": "pinned host allocation failed", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0}
    # host copy of the synthetic column (untimed): generated on the device chunk by chunk, read back through the ABI
    chunk = args.e2e_chunk_rows
    gen_t = ColumnarTable(ctx, [Field("col", I64, False)], min(e2e_rows, 1 << 28))
    done = 0
    while done < e2e_rows:
        m = min(gen_t.nrows, e2e_rows - done)
        gen_t.generate(0, capi.QE_GEN_UNIFORM_U40, SEED, row_offset=rank * rows + done)
        dense = Pipeline(gen_t, None, [col(0, I64, "col")]).run(None, 0, m)
        capi.check(capi.lib().qe_result_column(dense.handle, 0, C.c_void_p(h_in.ptr.value + 8 * done), None))
        dense.close()
        done += m
    gen_t.close()
    ctx.trim()
    t = ColumnarTable(ctx, [Field("col", I64, False)], 3 * chunk)     # the device window: three chunk slots
    pipes = [Pipeline(t, filter_expr(s), [col(0, I64, "col")], out_capacity_rows=c) for s, c in zip(SELECTIVITIES, caps)]
    pipes_sel = [Pipeline(t, filter_expr(s), [col(0, I64, "col")], out_capacity_rows=c, values=False) for s, c in zip(SELECTIVITIES, caps)]
    results = [None] * len(pipes)
    results_sel = [None] * len(pipes)
    kept = [0] * len(pipes)

    def one_step(which=0):
        ps, rs = (pipes, results) if which == 0 else (pipes_sel, results_sel)
        res = stream_pipelines(t, ps, [h_in.ptr.value], chunk_rows=chunk, results=rs,
                               host_selection=[b.ptr.value for b in h_sel],
                               host_out_values=[[b.ptr.value] for b in h_val] if which == 0 else None, total_rows=e2e_rows)
        for i, r in enumerate(res):
            rs[i] = r
            kept[i] = r.num_rows                                     # d2h read of the step's result sizes

    per_step = {0: [], 1: []}

    def timed(which, steps):
        one_step(which)                                              # warm-up
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(steps):
            t1 = time.perf_counter()
            one_step(which)                                          # returns after the step's results are in host memory
            per_step[which].append(round((time.perf_counter() - t1) * 1e3, 1))
        torch.cuda.synchronize()
        secs = time.perf_counter() - t0
        if dist is not None:
            tt = torch.tensor([secs], dtype=torch.float64, device=device)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            secs = float(tt.item())
        return secs

    steps = max(1, min(args.steps, args.e2e_steps))
    secs = timed(0, steps)
    # spot check of what landed in host memory: first kept row of the 50 % filter
    first_sel = C.cast(h_sel[1].ptr, C.POINTER(C.c_uint32))[0]
    first_val = C.cast(h_val[1].ptr, C.POINTER(C.c_int64))[0]
    src_val = C.cast(h_in.ptr, C.POINTER(C.c_int64))[first_sel]
    assert first_val == src_val and first_val < int(0.5 * 2 ** 40), "streamed output does not match the host column"
    last_sel = C.cast(h_sel[1].ptr, C.POINTER(C.c_uint32))[kept[1] - 1]
    assert C.cast(h_val[1].ptr, C.POINTER(C.c_int64))[kept[1] - 1] == C.cast(h_in.ptr, C.POINTER(C.c_int64))[last_sel]
    h2d = e2e_rows * 8
    d2h = sum(k * 12 + 8 for k in kept)
    kept_full_run = list(kept)
    secs_sel = timed(1, steps)
    assert kept == kept_full_run and C.cast(h_sel[1].ptr, C.POINTER(C.c_uint32))[0] == first_sel
    d2h_sel = sum(k * 4 + 8 for k in kept)
    for r in results + results_sel:
        if r is not None:
            r.close()
    t.close()
    ctx.trim()
    # the ceiling: the same bytes over the same link with no kernels, all ranks at once
    link_ms = link_sel_ms = same_ms = same_sel_ms = None
    try:
        link_ms = hostlink_yardstick(torch, dist, device, h2d, d2h)
        link_sel_ms = hostlink_yardstick(torch, dist, device, h2d, d2h_sel)
    except Exception:
        pass
    try:      # ... and into the very buffers the step used (the overwritten outputs were checked above)
        outs_sel = [(h_sel[i], kept_full_run[i] * 4) for i in range(len(pipes))]
        same_ms = hostlink_same_buffers(torch, dist, device, h_in, h2d, outs_sel + [(h_val[i], kept_full_run[i] * 8) for i in range(len(pipes))])
        same_sel_ms = hostlink_same_buffers(torch, dist, device, h_in, h2d, outs_sel)
    except Exception:
        pass
    for b in bufs:
        b.free()
    ms_step, ms_step_sel = secs / steps * 1e3, secs_sel / steps * 1e3
    out = {"value": world * len(pipes) * e2e_rows * steps / secs, "unit": UNIT, "rows_per_gpu": e2e_rows, "steps": steps,
           "ms_per_step": ms_step, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
           "chunk_rows": chunk, "device_window_rows": 3 * chunk,
           "hostlink_ms_same_bytes_no_kernels": link_ms, "frac_of_hostlink": (link_ms / ms_step) if link_ms else None,
           "plain_copies_same_buffers_ms": same_ms, "frac_of_plain_copies_same_buffers": (same_ms / ms_step) if same_ms else None,
           "step_ms": per_step[0],
           "selection_only": {"value": world * len(pipes) * e2e_rows * steps / secs_sel, "unit": UNIT, "ms_per_step": ms_step_sel,
                              "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h_sel,
                              "hostlink_ms_same_bytes_no_kernels": link_sel_ms,
                              "frac_of_hostlink": (link_sel_ms / ms_step_sel) if link_sel_ms else None,
                              "plain_copies_same_buffers_ms": same_sel_ms,
                              "frac_of_plain_copies_same_buffers": (same_sel_ms / ms_step_sel) if same_sel_ms else None,
                              "what": "QE_RUN_NO_VALUES: the selection vectors are the whole output (the host holds the table)"},
           "what": "per step one qe_pipeline_stream call (bounded mode): the column is copied from pinned host memory once, in chunks, into "
                   "a window of three chunk slots; the three filters run chunk by chunk; selection vectors + compacted columns land in "
                   "pinned host memory; uploads, kernels and downloads overlap on three streams"}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=1_000_000_000, help="rows per GPU")
    ap.add_argument("--cpu-rows", type=int, default=40_000_000, help="rows of the bounded CPU sample")
    ap.add_argument("--e2e-rows", type=int, default=0, help="rows per GPU for the end-to-end leg (0 = same as --rows)")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--e2e-chunk-rows", type=int, default=1 << 25, help="rows per chunk of the streamed end-to-end leg")
    ap.add_argument("--q6-rows", type=int, default=600_037_902)
    ap.add_argument("--tree-rows", type=int, default=1_000_000_000)
    ap.add_argument("--sf-rows", type=int, default=1_000_000_000, help="rows of the single-partition configs[4] extra")
    ap.add_argument("--no-extra", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--allow-interpreter", action="store_true")
    ap.add_argument("--steps-opt", type=int, default=0, help="tile shape override: steps per warp")
    ap.add_argument("--vec-opt", type=int, default=0, help="tile shape override: rows per lane per step (2 or 4)")
    ap.add_argument("--grid-per-sm", type=int, default=0)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
