#!/bin/bash
# round-2 call L: v8 = rolled RK4 stage loop over a running weighted sum + batched (load-before-store) solver vector loops
mkdir -p gpurun_out
timeout 500 python tests/quick_quad_condensed.py parity > gpurun_out/r02l_quad_parity.log 2>&1; echo "parity rc=$?"; grep -v bit-exact gpurun_out/r02l_quad_parity.log | tail -4
timeout 200 python tests/quick_quad_condensed.py speed3 2>&1 | grep -v "^$" | awk 'NR%2==0' | tee gpurun_out/r02l_quad_speed.log
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r02l_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r02l_pytest.log | cut -c1-250
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r02l_bench.json 2> gpurun_out/r02l_bench.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/r02l_bench.json
