#!/bin/bash
# round-2 call P: per-model occupancy: default (simplified 16, realistic 12) and realistic 8 / 10, both 14; B = 14208 (6-8 waves)
mkdir -p gpurun_out
timeout 500 python tests/quick_quad_condensed.py parity > gpurun_out/r02p_quad_parity.log 2>&1; echo "parity rc=$?"; grep -v bit-exact gpurun_out/r02p_quad_parity.log | tail -3
for lib in default build_variants/libqc_r8.so build_variants/libqc_r10.so build_variants/libqc_r14_s14.so; do
  echo "=== $lib" | tee -a gpurun_out/r02p_variants.log
  if [ $lib = default ]; then unset PANTR_B200_LIB; else export PANTR_B200_LIB=$PWD/$lib; fi
  timeout 300 python tests/quick_quad_condensed.py speed4 2>&1 | grep -v "^$" | tee -a gpurun_out/r02p_variants.log
done
