#!/bin/bash
# ncu --set full capture of the thread-per-instance quadcopter solve kernel (short, iteration-capped launch)
TAG=${1:-thr}
mkdir -p gpurun_out
ARGS="quadcopter 60 37888 thread 0.1 alm.max_iter=2 solver.max_iter=20 alm.initial_tolerance=1e-6"
timeout 200 python tests/quick_bench_quad.py $ARGS > gpurun_out/quick_thread_$TAG.log 2>&1 && \
timeout 700 ncu --set full --clock-control none --import-source on -k regex:solve_kernel_thread -c 1 -o gpurun_out/thread_$TAG python tests/quick_bench_quad.py $ARGS > gpurun_out/ncu_thread_$TAG.log 2>&1
echo "rc=$?"; cat gpurun_out/quick_thread_$TAG.log | cut -c1-300; tail -3 gpurun_out/ncu_thread_$TAG.log
