#!/bin/bash
# ncu --set full capture of the condensed quadcopter solve kernel on a short, iteration-capped launch (one resident wave),
# plus the source-level sections; usage: gpu_ncu_quad.sh <tag> [problem] [B]
TAG=${1:-quadc}; NAME=${2:-quadcopter}; B=${3:-2368}
mkdir -p gpurun_out
ARGS="$NAME 60 $B warp 0.1 alm.max_iter=3 solver.max_iter=25 alm.initial_tolerance=1e-6"
timeout 200 python tests/quick_bench_quad.py $ARGS > gpurun_out/quick_${TAG}.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:solve_kernel -c 1 -o gpurun_out/${TAG} python tests/quick_bench_quad.py $ARGS > gpurun_out/ncu_${TAG}.log 2>&1
echo "rc=$?"; cut -c1-300 gpurun_out/quick_${TAG}.log; tail -3 gpurun_out/ncu_${TAG}.log
