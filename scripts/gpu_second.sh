#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json; tail -3 gpurun_out/bench.err
timeout 300 python tests/quick_bench.py 60 1776 > gpurun_out/quick.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:solve_kernel -c 1 -o gpurun_out/prof_v0 python tests/quick_bench.py 60 1776 > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/quick.log; tail -5 gpurun_out/ncu_full.log
