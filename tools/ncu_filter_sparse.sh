#!/bin/bash
# `ncu --set full` captures of the ring filter's sparse path (row lists + sweeper warps) at 1 % and 5 % kept, and of the
# masks-only kernel at 1 % beside it (run on the GPU box).   bash tools/ncu_filter_sparse.sh [outdir]
out=${1:-gpurun_out/ncu_sparse}
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
run filter_s005 qe_filter_ring_kernel --what filter --sel 0.05
run filter_s001_masks_only qe_filter_ring_kernel --what filter --sel 0.01 --opt list_cap=-1
