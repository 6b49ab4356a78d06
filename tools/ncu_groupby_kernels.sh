set -x
cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_fullsize.py -x -q -m gpu -k groupby > gpurun_out/r02_t31_full_groupby.log 2>&1; echo t=$?; tail -5 gpurun_out/r02_t31_full_groupby.log
mkdir -p gpurun_out/ncu_gb
ncu --set full --clock-control none --import-source on -k regex:qe_groupby_lane8_kernel -s 1 -c 1 -f -o gpurun_out/ncu_gb/groupby_copies8_1024 python tools/bench_groupby.py --skip-q1 --cards 1024 --reps 1 > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:qe_groupby_lane_kernel -s 1 -c 1 -f -o gpurun_out/ncu_gb/groupby_copies32_256 python tools/bench_groupby.py --skip-q1 --cards 256 --reps 1 > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:qe_groupby_scatter_kernel -s 1 -c 1 -f -o gpurun_out/ncu_gb/groupby_scatter_65536 python tools/bench_groupby.py --skip-q1 --cards 65536 --reps 1 > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:qe_groupby_gather_kernel -s 1 -c 1 -f -o gpurun_out/ncu_gb/groupby_gather_65536 python tools/bench_groupby.py --skip-q1 --cards 65536 --reps 1 > /dev/null 2>&1
for f in gpurun_out/ncu_gb/*.ncu-rep; do ncu -i $f --page raw --csv > ${f%.ncu-rep}.raw.csv 2>/dev/null; done
ls -la gpurun_out/ncu_gb
