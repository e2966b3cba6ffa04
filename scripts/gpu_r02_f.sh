#!/bin/bash
# round-2 call F: after the ALM NaN-initialisation fix: parity, the branch tests, speed of the unrolled-RK4 build
mkdir -p gpurun_out
timeout 500 python tests/quick_quad_condensed.py parity > gpurun_out/r02f_quad_parity.log 2>&1; echo "parity rc=$?"; grep -v bit-exact gpurun_out/r02f_quad_parity.log | tail -4
timeout 200 python tests/quick_quad_condensed.py speed3 2>&1 | grep -v "^$" | awk 'NR%2==0' | tee gpurun_out/r02f_quad_speed.log
timeout 1200 python -m pytest tests/test_gpu_branches.py tests/test_gpu_parity.py -m gpu -q > gpurun_out/r02f_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r02f_pytest.log | cut -c1-250
