#!/bin/bash
# usage: gpu_quad.sh <tag> : quadcopter HVP throughput in both execution modes with bounded iteration counts
TAG=${1:-quad}
mkdir -p gpurun_out
CAP="alm.max_iter=10 solver.max_iter=100"
for args in "quadcopter 60 4736 warp 0.1 $CAP" "quadcopter 60 37888 thread 0.1 $CAP" "quadcopter 60 75776 thread 0.1 $CAP" "realistic_quadcopter 60 37888 thread 0.1 $CAP" "quadcopter 20 37888 thread 0.1"; do
  timeout 240 python tests/quick_bench_quad.py $args 2>&1 | tee -a gpurun_out/quick_quad_$TAG.log
done
