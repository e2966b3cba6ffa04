#!/bin/bash
# round-2 call O: occupancy variants of the v8 quadcopter kernel now that the L1 carve-out is larger
mkdir -p gpurun_out
for lib in default build_variants/libqc_b12.so build_variants/libqc_b20.so build_variants/libqc_b24.so; do
  echo "=== $lib" | tee -a gpurun_out/r02o_variants.log
  if [ $lib = default ]; then unset PANTR_B200_LIB; else export PANTR_B200_LIB=$PWD/$lib; fi
  timeout 200 python tests/quick_quad_condensed.py speed3 2>&1 | grep -v "^$" | awk 'NR%2==0' | tee -a gpurun_out/r02o_variants.log
done
