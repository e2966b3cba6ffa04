#!/bin/bash
# thread-kernel variants: GPU tests with the in-tree library, then the capped quadcopter throughput run per variant
TAG=$1; shift
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_$TAG.log
for lib in "$@"; do
  echo "=== $lib"
  PANTR_B200_LIB=$PWD/build_variants/$lib timeout 200 python tests/quick_bench_quad.py quadcopter 60 37888 thread 0.1 alm.max_iter=4 solver.max_iter=100 2>&1 | tail -1 | tee -a gpurun_out/tvar_$TAG.log
done
