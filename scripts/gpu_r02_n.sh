#!/bin/bash
# round-2 call N: v8 (plain maps): which solver vectors in shared memory (the carve-out it costs L1) -- PANTR_B200_SMEM_SET 0..3
mkdir -p gpurun_out
timeout 500 python tests/quick_quad_condensed.py parity > gpurun_out/r02n_quad_parity.log 2>&1; echo "parity rc=$?"; grep -v bit-exact gpurun_out/r02n_quad_parity.log | tail -3
for set in default 0 1 2; do
  echo "=== PANTR_B200_SMEM_SET=$set" | tee -a gpurun_out/r02n_smem_sets.log
  if [ $set = default ]; then unset PANTR_B200_SMEM_SET; else export PANTR_B200_SMEM_SET=$set; fi
  timeout 200 python tests/quick_quad_condensed.py speed3 2>&1 | grep -v "^$" | awk 'NR%2==0' | tee -a gpurun_out/r02n_smem_sets.log
done
