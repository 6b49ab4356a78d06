#!/usr/bin/env python
"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list (profiles/*_launches*.csv).

  python tools/ncu_launch_summary.py launches.csv summary.json "what the command was"
"""
import csv
import json
import sys


def main():
    src, dst = sys.argv[1], sys.argv[2]
    what = sys.argv[3] if len(sys.argv) > 3 else ""
    rows = [r for r in csv.reader(open(src, newline="", errors="replace")) if r and r[0].isdigit()]
    per = {}
    for r in rows:
        name, ns = r[4], float(r[-1])
        k = per.setdefault(name[:60], {"launches": 0, "total_ms": 0.0})
        k["launches"] += 1
        k["total_ms"] += ns / 1e6
    total = sum(v["total_ms"] for v in per.values()) or 1.0
    out = {"_what": what + "; per-launch times under ncu are cold-cache and serialised",
           "kernels": [dict(kernel=k, share=v["total_ms"] / total, **v) for k, v in sorted(per.items(), key=lambda kv: -kv[1]["total_ms"])]}
    json.dump(out, open(dst, "w"), indent=1)
    for k in out["kernels"]:
        print(f"{k['kernel'][:50]:50s} {k['launches']:5d} {k['total_ms']:10.3f} ms {k['share']:.3f}")


if __name__ == "__main__":
    main()
