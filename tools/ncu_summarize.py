#!/usr/bin/env python
"""Summarise `ncu --page raw --csv` exports (tools/ncu_all_kernels.sh) into one JSON for profiles/.

  python tools/ncu_summarize.py gpurun_out/ncu_r01 profiles/r01_ncu_all_kernels_summary.json
"""
import csv
import glob
import json
import os
import sys

KEEP = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__bytes.sum.per_second", "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "launch__registers_per_thread", "launch__block_size", "launch__grid_size",
    "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static", "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts.sum.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "smsp__average_warp_latency_issue_stalled_long_scoreboard_per_warp_active.pct",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_drain_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_tex_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_imc_miss_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_gmma_per_issue_active.ratio",
]


def read(path):
    rows = list(csv.reader(open(path, newline="")))
    rows = [r for r in rows if r]
    head = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    names, units, vals = rows[head], rows[head + 1], rows[head + 2]
    out = {"kernel": vals[names.index("Kernel Name")] if "Kernel Name" in names else None}
    for m in KEEP:
        if m in names:
            i = names.index(m)
            out[m] = [vals[i], units[i]]
    return out


def main():
    src, dst = sys.argv[1], sys.argv[2]
    out = {"_what": "one `ncu --set full --clock-control none --import-source on` capture per hot kernel, third launch of "
                    "tools/profile_one.py at the size bench.py times it (tools/ncu_all_kernels.sh); durations are under the "
                    "profiler (serialised, counters on) and are not bench values"}
    for f in sorted(glob.glob(os.path.join(src, "*.raw.csv"))):
        name = os.path.basename(f)[:-len(".raw.csv")]
        try:
            out[name] = read(f)
        except Exception as e:   # noqa: BLE001
            out[name] = {"error": str(e)}
    json.dump(out, open(dst, "w"), indent=1)
    for k, v in out.items():
        if isinstance(v, dict) and "gpu__time_duration.sum" in v:
            rd, wr = v.get("dram__bytes_read.sum", ["?", ""]), v.get("dram__bytes_write.sum", ["?", ""])
            print(f"{k:18s} {v['kernel'][:28]:28s} {v['gpu__time_duration.sum'][0]:>10s} {v['gpu__time_duration.sum'][1]:3s}"
                  f" rd {rd[0]} {rd[1]} wr {wr[0]} {wr[1]} dram% {v.get("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", ["?"])[0]}"
                  f" issue% {v.get('smsp__issue_active.avg.pct_of_peak_sustained_active', ['?'])[0]}"
                  f" regs {v.get('launch__registers_per_thread', ['?'])[0]}")


if __name__ == "__main__":
    main()
