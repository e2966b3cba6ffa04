#!/bin/bash
# Parity tests + quick quadcopter throughput.   usage: bash scripts/gpu_quad.sh <tag>
TAG=${1:-quad}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_$TAG.log
timeout 600 python tests/quick_bench_quad.py quadcopter 60 4096 > gpurun_out/quick_quad_$TAG.log 2>&1; cat gpurun_out/quick_quad_$TAG.log
timeout 600 python tests/quick_bench_quad.py realistic_quadcopter 60 4096 > gpurun_out/quick_rquad_$TAG.log 2>&1; cat gpurun_out/quick_rquad_$TAG.log
timeout 600 python tests/quick_bench.py 60 16384 > gpurun_out/quick_$TAG.log 2>&1; cat gpurun_out/quick_$TAG.log
