#!/bin/bash
# usage: gpu_prof.sh <tag> <lib under build_variants or empty>
TAG=$1; LIB=$2
[ -n "$LIB" ] && export PANTR_B200_LIB=$PWD/build_variants/$LIB
timeout 300 python tests/quick_bench.py 60 2368 > gpurun_out/quick_small_$TAG.log 2>&1 && \
timeout 900 ncu --section SourceCounters --section WarpStateStats --section SchedulerStats --section InstructionStats --section SpeedOfLight --section ComputeWorkloadAnalysis --section MemoryWorkloadAnalysis --section Occupancy --section LaunchStats --clock-control none --import-source on -k regex:solve_kernel -c 1 -o gpurun_out/src_$TAG python tests/quick_bench.py 60 2368 > gpurun_out/ncu_src_$TAG.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/quick_small_$TAG.log
