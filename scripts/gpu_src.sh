#!/bin/bash
# Source-level ncu capture of the warm solve kernel (B = one wave).  usage: gpu_src.sh <tag> [lib under build_variants]
TAG=${1:-src}; LIB=$2
[ -n "$LIB" ] && export PANTR_B200_LIB=$PWD/build_variants/$LIB
mkdir -p gpurun_out
timeout 300 python tests/quick_bench.py 60 2368 > gpurun_out/quick_small_$TAG.log 2>&1 && \
timeout 900 ncu --section SourceCounters --section WarpStateStats --section SchedulerStats --section InstructionStats --section MemoryWorkloadAnalysis --section SpeedOfLight --clock-control none --import-source on -k regex:solve_kernel -s 2 -c 1 -o gpurun_out/src_$TAG python tests/quick_bench.py 60 2368 > gpurun_out/ncu_src_$TAG.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/quick_small_$TAG.log
