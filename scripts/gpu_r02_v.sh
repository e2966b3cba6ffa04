#!/bin/bash
# round-2 call V: ncu launch list of a configs[2] bench command (kernel share of the step)
mkdir -p gpurun_out
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02v_launches_quadcopter.csv python bench.py --config quadcopter --batch 4096 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r02v_ncu.log 2>&1; echo "rc=$?"; wc -l gpurun_out/r02v_launches_quadcopter.csv; tail -2 gpurun_out/r02v_ncu.log | cut -c1-300
