#!/bin/bash
# One `ncu --set full` capture per hot kernel of DESIGN.md section 4, at the size bench.py times it
# (run on the GPU box; each program has already exited 0 without ncu in the bench / tests).
#   bash tools/ncu_all_kernels.sh [outdir]      -> <outdir>/<name>.ncu-rep + <name>.raw.csv
out=${1:-gpurun_out/ncu}
mkdir -p "$out"
run() {   # name kernel-regex args...
    name=$1; kern=$2; shift 2
    python tools/profile_one.py "$@" --reps 3 > "$out/$name.plain.log" 2>&1 || { echo "$name: plain run failed"; return; }
    ncu --set full --clock-control none --import-source on -k "regex:$kern" --launch-skip 2 --launch-count 1 \
        -f -o "$out/$name" python tools/profile_one.py "$@" --reps 3 > "$out/$name.ncu.log" 2>&1
    ncu -i "$out/$name.ncu-rep" --page raw --csv > "$out/$name.raw.csv" 2>/dev/null
    echo "$name: $(tail -1 "$out/$name.plain.log")"
}
run filter_s001 qe_filter_ring_kernel --what filter --sel 0.01
run filter_s050 qe_filter_ring_kernel --what filter --sel 0.5
run filter_s099 qe_filter_ring_kernel --what filter --sel 0.99
run q6_agg qe_agg_kernel --what q6 --rows 600037902
run tree_project qe_project_kernel --what tree
run sf_filter_2proj qe_filter_ring_kernel --what sf
run q1_groupby qe_groupby_reg_kernel --what q1 --rows 600037902
