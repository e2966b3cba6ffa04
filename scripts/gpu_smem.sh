#!/bin/bash
# usage: gpu_smem.sh <tag> <lib>:<smem_set> ...   parity tests with the in-tree library, then quick bench of each variant
TAG=$1; shift
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_$TAG.log
for v in "$@"; do
  lib=${v%%:*}; set=${v##*:}
  echo "=== $lib SMEM_SET=$set"
  PANTR_B200_SMEM_SET=$set PANTR_B200_LIB=$PWD/build_variants/$lib timeout 600 python tests/quick_bench.py 60 16384 2>&1 | tee -a gpurun_out/variants_$TAG.log
done
