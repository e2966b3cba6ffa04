#!/bin/bash
# usage: gpu_variants.sh <tag> <lib1> <lib2> ...   (libs under build_variants/)
TAG=$1; shift
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_$TAG.log
for lib in "$@"; do
  echo "=== $lib"
  PANTR_B200_LIB=$PWD/build_variants/$lib timeout 600 python tests/quick_bench.py 60 16384 2>&1 | tee -a gpurun_out/variants_$TAG.log
done
