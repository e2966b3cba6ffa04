#!/bin/bash
# round-2 call U: configs[4] (mixed) at 3 x 16384 instances: large enough that the quadcopter launches are not one wave
mkdir -p gpurun_out
timeout 520 python bench.py --config mixed --batch 49152 --steps 1 --warmup 1 --cpu-seconds 15 > gpurun_out/r02u_bench_mixed_49152.json 2> gpurun_out/r02u_bench_mixed.err; echo "mixed rc=$?"; cut -c1-400 gpurun_out/r02u_bench_mixed_49152.json; tail -3 gpurun_out/r02u_bench_mixed.err
