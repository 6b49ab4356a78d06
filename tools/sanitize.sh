#!/bin/bash
# compute-sanitizer over tools/sanitizer_subset.py (memcheck, racecheck, synccheck); logs under gpurun_out/, summaries
# are copied to profiles/ by hand.  Run on the GPU box:  bash tools/sanitize.sh [tools...]
set -u
tools=${@:-"memcheck racecheck synccheck"}
mkdir -p gpurun_out
for tool in $tools; do
  log=gpurun_out/r02_sanitizer_${tool}.log
  extra=""
  [ "$tool" == "memcheck" ] && extra="--leak-check no"
  timeout 1500 compute-sanitizer --tool $tool $extra --print-limit 40 python tools/sanitizer_subset.py > $log 2>&1
  echo "$tool: exit $?; $(grep -c 'ERROR SUMMARY\|RACECHECK SUMMARY' $log) summary line(s): $(grep 'ERROR SUMMARY\|RACECHECK SUMMARY' $log | tail -1)"
  tail -3 $log
done
