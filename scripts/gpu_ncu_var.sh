#!/bin/bash
# one --set full capture of the warm solve (B = 2368, one wave) for a variant library
# usage: gpu_ncu_var.sh <tag> <lib under build_variants/>
TAG=$1; LIB=$2
mkdir -p gpurun_out
export PANTR_B200_LIB=$PWD/build_variants/$LIB
timeout 200 python tests/quick_bench.py 60 2368 > gpurun_out/quick_small_$TAG.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:solve_kernel -s 2 -c 1 -o gpurun_out/full_$TAG python tests/quick_bench.py 60 2368 > gpurun_out/ncu_full_$TAG.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/quick_small_$TAG.log; ls -la gpurun_out/ | tail -5
