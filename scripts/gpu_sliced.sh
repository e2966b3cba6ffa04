#!/bin/bash
TAG=${1:-sl}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
echo "=== warp"; PANTR_EXEC_MODE=warp timeout 200 python tests/quick_bench.py 60 16384 2>&1 | tail -2 | tee -a gpurun_out/sliced_$TAG.log
echo "=== sliced rr"; PANTR_B200_URGENT_CG=-1 PANTR_EXEC_MODE=sliced timeout 200 python tests/quick_bench.py 60 16384 2>&1 | tail -1 | tee -a gpurun_out/sliced_$TAG.log
