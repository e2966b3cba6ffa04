#!/bin/bash
# round-2 call D: condensed quadcopter sweeps v3 (locals, rolled RK4): parity, the GPU
# test suite, speed, ncu, quadcopter bench line at B = 16384
mkdir -p gpurun_out
timeout 500 python tests/quick_quad_condensed.py parity > gpurun_out/r02d_quad_parity.log 2>&1; echo "parity rc=$?"; grep -v bit-exact gpurun_out/r02d_quad_parity.log | tail -8
timeout 400 python tests/quick_quad_condensed.py speed2 > gpurun_out/r02d_quad_speed.log 2>&1; echo "speed rc=$?"; cat gpurun_out/r02d_quad_speed.log
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/r02d_pytest.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/r02d_pytest.log | cut -c1-250
timeout 600 bash scripts/gpu_ncu_quad.sh r02d_quadc quadcopter 2368
ncu -i gpurun_out/r02d_quadc.ncu-rep --page raw --csv 2>/dev/null | python tools/ncu_summary.py > gpurun_out/r02d_quadc_ncu_full_summary.csv
ncu -i gpurun_out/r02d_quadc.ncu-rep --page source --csv --print-source sass 2>/dev/null > gpurun_out/r02d_quadc_source_sass.csv; rm -f gpurun_out/r02d_quadc.ncu-rep
timeout 900 python bench.py --config quadcopter --batch 16384 --steps 1 --warmup 1 --cpu-seconds 20 > gpurun_out/r02d_bench_quadcopter_16384.json 2> gpurun_out/r02d_bench_quadcopter.err; echo "quad bench rc=$?"; cut -c1-400 gpurun_out/r02d_bench_quadcopter_16384.json; tail -3 gpurun_out/r02d_bench_quadcopter.err
