#!/bin/bash
# which solver vectors live in shared memory (PANTR_B200_SMEM_SET 0/1/2, default = host rule): quick bench each
mkdir -p gpurun_out
for set in default 0 1 2; do
  echo "=== SMEM_SET=$set" | tee -a gpurun_out/smemset.log
  if [ $set = default ]; then timeout 200 python tests/quick_bench.py 60 16384 2>&1 | tee -a gpurun_out/smemset.log
  else PANTR_B200_SMEM_SET=$set timeout 200 python tests/quick_bench.py 60 16384 2>&1 | tee -a gpurun_out/smemset.log; fi
done
