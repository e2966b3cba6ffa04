#!/bin/bash
# Hanging-chain Hessian-vector-product variants (PANTR_HC_VARIANT 0/1/2): parity subset + quick bench each.
# usage: gpu_hcvar.sh <tag> <lib1> <lib2> ...   (libs under build_variants/)
TAG=$1; shift
mkdir -p gpurun_out
for lib in "$@"; do
  echo "=== $lib" | tee -a gpurun_out/hcvar_$TAG.log
  PANTR_B200_LIB=$PWD/build_variants/$lib timeout 300 python -m pytest tests -m gpu -x -q -k "hanging_chain or iteration_sliced or order_hint or closed_loop_driver" 2>&1 | tail -3 | tee -a gpurun_out/hcvar_$TAG.log
  PANTR_B200_LIB=$PWD/build_variants/$lib timeout 200 python tests/quick_bench.py 60 16384 2>&1 | tee -a gpurun_out/hcvar_$TAG.log
done
for set in ${HCVAR_SETS:-}; do
  echo "=== libv2.so SMEM_SET=$set" | tee -a gpurun_out/hcvar_$TAG.log
  PANTR_B200_SMEM_SET=$set PANTR_B200_LIB=$PWD/build_variants/libv2.so timeout 200 python tests/quick_bench.py 60 16384 2>&1 | tee -a gpurun_out/hcvar_$TAG.log
done
