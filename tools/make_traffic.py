#!/usr/bin/env python
"""profiles/traffic.json (what bench.py reports as roofline.traffic) from an ncu summary made by tools/ncu_summarize.py.

  python tools/make_traffic.py profiles/r02_ncu_all_kernels_summary.json
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src = sys.argv[1]
s = json.load(open(src))


def num(v):
    x, unit = float(v[0].replace(",", "")), v[1]
    return x * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1.0, "us": 1e-3, "ns": 1e-6, "s": 1e3, "msecond": 1.0,
                "usecond": 1e-3, "nsecond": 1e-6, "second": 1e3}.get(unit, 1.0)


per = {}
for key, sel in (("filter_s001", "0.01"), ("filter_s050", "0.5"), ("filter_s099", "0.99")):
    k = s[key]
    per[sel] = {"dram_bytes": num(k["dram__bytes_read.sum"]) + num(k["dram__bytes_write.sum"]), "duration_ms": num(k["gpu__time_duration.sum"])}
out = {"filter_sweep_dram_bytes_per_launch": sum(v["dram_bytes"] for v in per.values()) / 3, "per_selectivity": per,
       "source": f"ncu --set full --clock-control none, one launch per selectivity, tools/ncu_all_kernels.sh (tools/profile_one.py --sel S, 1e9 int64 rows); {os.path.relpath(src, ROOT)}",
       "kernel": "qe_filter_ring_kernel = qe_filter_ring_body<P>, 4480-row (35 KB) stages, 6 stages, 20 consumer warps, defer 3"}
json.dump(out, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
print(json.dumps(out, indent=1))
