#!/bin/bash
# Lean round checkpoint: parity tests, smoke, both bench arms, ncu launch list of the bench command, one
# `--set full` capture of the bench's own launch shape.   usage: bash scripts/gpu_checkpoint2.sh <tag>
TAG=${1:-r01}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_$TAG.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke_$TAG.log
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err; echo "ref rc=$?"
timeout 900 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_$TAG.csv python bench.py --no-cpu-baseline > gpurun_out/bench_under_ncu_$TAG.log 2>&1
echo "bench/ncu rc=$?"; cat gpurun_out/bench_$TAG.json; tail -3 gpurun_out/bench_$TAG.err; cat gpurun_out/bench_ref_$TAG.json
timeout 900 ncu --set full --clock-control none --import-source on -k regex:solve_kernel -s 4 -c 1 -o gpurun_out/fullbench_$TAG python bench.py --no-cpu-baseline --steps 1 --warmup 3 > gpurun_out/ncu_fullbench_$TAG.log 2>&1
echo "ncu fullbench rc=$?"
