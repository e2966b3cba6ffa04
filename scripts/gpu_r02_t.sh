#!/bin/bash
# round-2 call T: the committed bench.py once more after its last edits (traffic lookup by source hash, parity check without
# MaxTime instances), smoke(), and the bulk throughput of the realistic quadcopter kernel: one horizon of configs[3] at 16384
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/r02t_bench.json 2> gpurun_out/r02t_bench.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/r02t_bench.json; tail -3 gpurun_out/r02t_bench.err
timeout 1200 python bench.py --config realistic_sweep --horizons 60 --batch 16384 --steps 1 --warmup 1 --cpu-seconds 20 > gpurun_out/r02t_bench_realistic_N60_16384.json 2> gpurun_out/r02t_bench_realistic.err; echo "realistic rc=$?"; cut -c1-400 gpurun_out/r02t_bench_realistic_N60_16384.json; tail -3 gpurun_out/r02t_bench_realistic.err
