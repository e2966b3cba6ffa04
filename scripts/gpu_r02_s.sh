#!/bin/bash
# round-2 final call S (2 GPUs): the 2-GPU sharded-solve tests, bench.py --gpus 2 (weak and one global batch really sharded)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sharded.py -m gpu -q > gpurun_out/r02s_pytest_sharded.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02s_pytest_sharded.log | cut -c1-250
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r02s_bench_2gpu_weak.json 2> gpurun_out/r02s_bench_2gpu_weak.err; echo "weak rc=$?"; grep '"metric"' gpurun_out/r02s_bench_2gpu_weak.json | cut -c1-400
timeout 900 $TR --master-port 29512 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline --shard global --batch 32768 > gpurun_out/r02s_bench_2gpu_global.json 2> gpurun_out/r02s_bench_2gpu_global.err; echo "global rc=$?"; grep '"metric"' gpurun_out/r02s_bench_2gpu_global.json | cut -c1-400
