#!/bin/bash
# round-2 call I: v6 = recurrences with running shared-window addresses, one code path for x- and u-lanes, constant columns in shared memory
mkdir -p gpurun_out
timeout 500 python tests/quick_quad_condensed.py parity > gpurun_out/r02i_quad_parity.log 2>&1; echo "parity rc=$?"; grep -v bit-exact gpurun_out/r02i_quad_parity.log | tail -4
timeout 200 python tests/quick_quad_condensed.py speed3 2>&1 | grep -v "^$" | awk 'NR%2==0' | tee gpurun_out/r02i_quad_speed.log
for lib in $(ls build_variants/*.so 2>/dev/null); do
  echo "=== $lib" | tee -a gpurun_out/r02i_quad_speed.log
  PANTR_B200_LIB=$PWD/$lib timeout 200 python tests/quick_quad_condensed.py speed3 2>&1 | grep -v "^$" | awk 'NR%2==0' | tee -a gpurun_out/r02i_quad_speed.log
done
