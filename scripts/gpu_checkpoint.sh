#!/bin/bash
# Round checkpoint on the GPU: parity tests, smoke, both bench arms, ncu launch list of the bench
# command, one `--set full` ncu capture of the solve kernel.   usage: bash scripts/gpu_checkpoint.sh <tag>
TAG=${1:-r01}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_$TAG.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke_$TAG.log
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err; echo "ref rc=$?"
timeout 900 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_$TAG.csv python bench.py --no-cpu-baseline > gpurun_out/bench_under_ncu_$TAG.log 2>&1
echo "bench/ncu rc=$?"; cat gpurun_out/bench_$TAG.json; tail -3 gpurun_out/bench_$TAG.err; cat gpurun_out/bench_ref_$TAG.json
timeout 300 python tests/quick_bench.py 60 2368 > gpurun_out/quick_small_$TAG.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:solve_kernel -s 2 -c 1 -o gpurun_out/full_$TAG python tests/quick_bench.py 60 2368 > gpurun_out/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"; cat gpurun_out/quick_small_$TAG.log; ls -la gpurun_out/
# one --set full capture of the bench's own timed launch shape (B = 16384 warm solve): DRAM traffic per launch
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:solve_kernel -s 4 -c 1 -o gpurun_out/fullbench_$TAG python bench.py --no-cpu-baseline --steps 1 --warmup 3 > gpurun_out/ncu_fullbench_$TAG.log 2>&1
echo "ncu fullbench rc=$?"
