#!/bin/bash
TAG=${1:-sl}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "sliced" 2>&1 | tail -15
for mode in warp sliced; do
  echo "=== $mode"; PANTR_EXEC_MODE=$mode timeout 200 python tests/quick_bench.py 60 16384 2>&1 | tee -a gpurun_out/sliced_$TAG.log
done
