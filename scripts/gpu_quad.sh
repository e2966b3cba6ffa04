#!/bin/bash
# usage: gpu_quad.sh <tag> : all GPU tests, then quadcopter quick benches in both execution modes
TAG=${1:-quad}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_$TAG.log
for args in "quadcopter 60 1024 warp 0.1" "quadcopter 60 4096 thread 0.1" "quadcopter 60 37888 thread 0.1" "realistic_quadcopter 60 4096 thread 0.1"; do
  timeout 900 python tests/quick_bench_quad.py $args 2>&1 | tee -a gpurun_out/quick_quad_$TAG.log
done
