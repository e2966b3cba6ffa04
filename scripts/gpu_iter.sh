#!/bin/bash
# Iteration loop on the GPU: parity tests, quick throughput, light ncu capture of one solve kernel.
# usage: bash scripts/gpu_iter.sh <tag> [ncu]
TAG=${1:-iter}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_$TAG.log
timeout 600 python tests/quick_bench.py 60 16384 > gpurun_out/quick_$TAG.log 2>&1; echo "quick rc=$?"; cat gpurun_out/quick_$TAG.log
if [ "$2" == "ncu" ]; then
timeout 300 python tests/quick_bench.py 60 1776 > gpurun_out/quick_small_$TAG.log 2>&1 && \
timeout 900 ncu --section SpeedOfLight --section WarpStateStats --section SchedulerStats --section InstructionStats --section Occupancy --section LaunchStats --section MemoryWorkloadAnalysis --section ComputeWorkloadAnalysis --clock-control none --import-source on -k regex:solve_kernel -c 1 -o gpurun_out/prof_$TAG python tests/quick_bench.py 60 1776 > gpurun_out/ncu_$TAG.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/quick_small_$TAG.log
fi
