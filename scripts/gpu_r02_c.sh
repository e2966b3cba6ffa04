#!/bin/bash
# round-2 call C: condensed quadcopter recurrences v2 (TMA for AB, cp.async rows, no per-step fences): parity, the GPU
# test suite, speed, ncu; hanging-chain TMA transport A/B
mkdir -p gpurun_out
timeout 500 python tests/quick_quad_condensed.py parity > gpurun_out/r02c_quad_parity.log 2>&1; echo "parity rc=$?"; grep -v bit-exact gpurun_out/r02c_quad_parity.log | tail -8
timeout 400 python tests/quick_quad_condensed.py speed2 > gpurun_out/r02c_quad_speed.log 2>&1; echo "speed rc=$?"; cat gpurun_out/r02c_quad_speed.log
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/r02c_pytest.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/r02c_pytest.log | cut -c1-250
timeout 600 bash scripts/gpu_ncu_quad.sh r02c_quadc quadcopter 2368
ncu -i gpurun_out/r02c_quadc.ncu-rep --page raw --csv 2>/dev/null | python tools/ncu_summary.py > gpurun_out/r02c_quadc_ncu_full_summary.csv
ncu -i gpurun_out/r02c_quadc.ncu-rep --page source --csv --print-source sass 2>/dev/null > gpurun_out/r02c_quadc_source_sass.csv; rm -f gpurun_out/r02c_quadc.ncu-rep
for lib in default libhc_tma.so; do
  echo "=== hanging chain, $lib" | tee -a gpurun_out/r02c_hc_tma_ab.log
  if [ $lib = default ]; then unset PANTR_B200_LIB; else export PANTR_B200_LIB=$PWD/build_variants/$lib; fi
  timeout 300 python -m pytest tests -m gpu -x -q -k "hanging_chain and not plugin" 2>&1 | tail -2 | tee -a gpurun_out/r02c_hc_tma_ab.log
  timeout 200 python tests/quick_bench.py 60 16384 2>&1 | tee -a gpurun_out/r02c_hc_tma_ab.log
done
unset PANTR_B200_LIB
