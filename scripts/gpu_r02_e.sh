#!/bin/bash
# round-2 call E: condensed quadcopter v4 (variable columns only): parity, trace diff of a failing ALM-backtrack case,
# speed of the default build and of the ring-depth / occupancy variants, full GPU test suite, ncu
mkdir -p gpurun_out
timeout 500 python tests/quick_quad_condensed.py parity > gpurun_out/r02e_quad_parity.log 2>&1; echo "parity rc=$?"; grep -v bit-exact gpurun_out/r02e_quad_parity.log | tail -8
timeout 300 python tests/debug_branch_trace.py quadcopter 12 1 alm.max_num_retries=3 alm.max_total_num_retries=3 solver.max_iter=4 alm.penalty_update_factor_lower=0.5 alm.initial_tolerance=1e-2 > gpurun_out/r02e_trace_diff.log 2>&1; echo "trace rc=$?"; head -60 gpurun_out/r02e_trace_diff.log
for lib in default libqc_d8b16.so libqc_d8b12.so libqc_d4b20.so libqc_d2b24.so; do
  echo "=== $lib" | tee -a gpurun_out/r02e_quad_variants.log
  if [ $lib = default ]; then unset PANTR_B200_LIB; else export PANTR_B200_LIB=$PWD/build_variants/$lib; fi
  timeout 200 python tests/quick_quad_condensed.py speed3 2>&1 | grep -v "^$" | awk 'NR%2==0' | tee -a gpurun_out/r02e_quad_variants.log
done
unset PANTR_B200_LIB
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/r02e_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/r02e_pytest.log | cut -c1-250
timeout 600 bash scripts/gpu_ncu_quad.sh r02e_quadc quadcopter 2368
ncu -i gpurun_out/r02e_quadc.ncu-rep --page raw --csv 2>/dev/null | python tools/ncu_summary.py > gpurun_out/r02e_quadc_ncu_full_summary.csv
ncu -i gpurun_out/r02e_quadc.ncu-rep --page source --csv --print-source sass 2>/dev/null > gpurun_out/r02e_quadc_source_sass.csv; rm -f gpurun_out/r02e_quadc.ncu-rep
