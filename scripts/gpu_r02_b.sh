#!/bin/bash
# round-2 call B: condensed quadcopter path with the TMA ring (parity, speed), full GPU test suite, default bench (both arms),
# ncu capture of the quadcopter solve kernel, quadcopter bench line at a reduced batch
mkdir -p gpurun_out
timeout 500 python tests/quick_quad_condensed.py parity > gpurun_out/r02b_quad_parity.log 2>&1; echo "parity rc=$?"; tail -4 gpurun_out/r02b_quad_parity.log
timeout 400 python tests/quick_quad_condensed.py speed2 > gpurun_out/r02b_quad_speed.log 2>&1; echo "speed rc=$?"; cat gpurun_out/r02b_quad_speed.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02b_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r02b_pytest.log
timeout 600 python bench.py --impl reference > gpurun_out/r02b_bench_ref.json 2> gpurun_out/r02b_bench_ref.err; echo "ref rc=$?"; cut -c1-300 gpurun_out/r02b_bench_ref.json
timeout 900 python bench.py > gpurun_out/r02b_bench.json 2> gpurun_out/r02b_bench.err; echo "bench rc=$?"; cut -c1-1500 gpurun_out/r02b_bench.json; tail -3 gpurun_out/r02b_bench.err
timeout 900 bash scripts/gpu_ncu_quad.sh r02b_quadc quadcopter 2368
ncu -i gpurun_out/r02b_quadc.ncu-rep --page raw --csv 2>/dev/null | python tools/ncu_summary.py > gpurun_out/r02b_quadc_ncu_full_summary.csv; rm -f gpurun_out/r02b_quadc.ncu-rep.tmp
timeout 900 python bench.py --config quadcopter --batch 16384 --steps 1 --warmup 1 --cpu-seconds 20 > gpurun_out/r02b_bench_quadcopter_16384.json 2> gpurun_out/r02b_bench_quadcopter.err; echo "quad bench rc=$?"; cut -c1-1200 gpurun_out/r02b_bench_quadcopter_16384.json; tail -3 gpurun_out/r02b_bench_quadcopter.err
