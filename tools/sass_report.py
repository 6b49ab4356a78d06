#!/usr/bin/env python
"""SASS evidence for the hot kernels, without a GPU: the cubins NVRTC makes for the BASELINE.json pipelines are
disassembled (cuobjdump -sass) and summarised -- which instructions prove the design (UBLKCP = cp.async.bulk on the TMA
unit, SYNCS = mbarrier arrive / try_wait, VOTE + POPC = ballot compaction, LDG.E.*256* / STG, REDUX, SHFL, ATOMS), how
many there are, registers and shared memory.  Writes profiles/r02_sass_<kernel>.txt (excerpt) and
profiles/r02_sass_summary.json.

  python tools/sass_report.py
"""
import collections
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

KERNELS = {"filter_i64": "qe_filter_ring_kernel", "q6": "qe_agg_kernel", "tree": "qe_project_kernel", "q6_filter": "qe_filter_ring_kernel",
           "filter_f64_nullable": "qe_filter_ring_kernel"}
MARKERS = ["UBLKCP", "SYNCS", "VOTE", "POPC", "REDUX", "SHFL", "LDG", "STG", "LDS", "STS", "ATOMS", "ATOMG", "ATOM", "RED", "DSETP", "ISETP",
           "DFMA", "DMUL", "DADD", "MUFU", "BAR", "NANOSLEEP", "S2R", "S2UR", "CCTL", "UTMALDG", "UTMAPF", "BSSY", "BSYNC", "LDC", "LDCU"]


def main():
    out_dir = os.path.join(ROOT, "profiles")
    summary = {}
    for name, entry in KERNELS.items():
        prefix = f"/tmp/qe_sass_{name}"
        subprocess.run([sys.executable, os.path.join(ROOT, "tools", "dump_kernel.py"), name, prefix], check=True, capture_output=True)
        sass = subprocess.run(["cuobjdump", "-sass", prefix + ".cubin"], check=True, capture_output=True, text=True).stdout
        res = subprocess.run(["cuobjdump", "-res-usage", prefix + ".cubin"], capture_output=True, text=True).stdout
        ops = collections.Counter()
        lines = []
        for ln in sass.splitlines():
            m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
            if m:
                ops[m.group(2)] += 1
                lines.append(ln.rstrip())
        total = sum(ops.values())
        by_marker = {mk: sum(v for k, v in ops.items() if k.split(".")[0] == mk) for mk in MARKERS}
        by_marker = {k: v for k, v in by_marker.items() if v}
        wide = {k: v for k, v in ops.items() if ("256" in k or "128" in k) and k.split(".")[0] in ("LDG", "STG", "LDS", "STS")}
        usage = re.search(r"REG:(\d+).*?SHARED:(\d+)", res.replace("\n", " "))
        summary[name] = {"entry": entry, "instructions": total, "by_mnemonic": by_marker, "wide_memory_ops": wide,
                         "registers": int(usage.group(1)) if usage else None, "static_shared_bytes": int(usage.group(2)) if usage else None,
                         "top": ops.most_common(12)}
        keep = [ln for ln in lines if re.search(r"UBLKCP|SYNCS|VOTE|POPC|REDUX|LDG|STG|ATOMS|ATOMG|RED\.|STS\.U8|STS\.U16|LDS\.U8|LDS\.U16|NANOSLEEP|UTMA", ln)]
        with open(os.path.join(out_dir, f"r02_sass_{name}.txt"), "w") as f:
            f.write(f"# {entry} of the '{name}' pipeline (tools/dump_kernel.py {name}; cuobjdump -sass): {total} instructions, "
                    f"{summary[name]['registers']} registers\n# the lines that carry the design (bulk copies, mbarriers, ballot / popc, wide loads and stores)\n")
            f.write("\n".join(keep[:400]) + "\n")
    with open(os.path.join(out_dir, "r02_sass_summary.json"), "w") as f:
        json.dump(summary, f, indent=1)
    for k, v in summary.items():
        print(k, v["entry"], v["instructions"], "instr;", v["by_mnemonic"])


if __name__ == "__main__":
    main()
