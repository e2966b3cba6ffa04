#!/bin/bash
# compute-sanitizer pass over the device code (the reference keeps its inner loop
# allocation-free and NaN-poisons stale iterates in Debug, alpaqa/cmake/Sanitizers.cmake:1-15,
# pantr.tpp:242,460-467; the CUDA analogue is memcheck / racecheck / initcheck / synccheck).
# usage: bash scripts/gpu_sanitize.sh <tag> ["case1;case2;..."]   logs: gpurun_out/<tag>_sanitize_<tool>.log
TAG=${1:-r02}
CASES=${2:-"hanging_chain 7 96 warp face;quadcopter 7 96 warp face;realistic_quadcopter 7 64 warp face;quadcopter 33 40 warp"}
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
for tool in memcheck racecheck initcheck synccheck; do
  log=gpurun_out/${TAG}_sanitize_${tool}.log
  : > $log
  IFS=';' read -ra LIST <<< "$CASES"
  for c in "${LIST[@]}"; do
    echo "=== $tool: $c" >> $log
    timeout 900 $CS --tool $tool --print-limit 20 --error-exitcode 99 python scripts/sanitize_case.py $c >> $log 2>&1
    echo "exit code $?" >> $log
  done
  echo "$tool: $(grep -c 'ERROR SUMMARY: 0 errors' $log) clean of $(grep -c '^===' $log) cases; $(grep 'ERROR SUMMARY' $log | grep -vc ': 0 errors') with errors"
done
