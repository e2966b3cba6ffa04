#!/bin/bash
# round-2 call M: A/B on one box: batched vs plain solver vector loops (hanging chain + quadcopters), occupancy / ring depth variants
mkdir -p gpurun_out
for lib in default build_variants/libmap_plain.so build_variants/libqc_b12.so build_variants/libqc_b20.so build_variants/libqc_d8.so; do
  echo "=== $lib" | tee -a gpurun_out/r02m_variants.log
  if [ $lib = default ]; then unset PANTR_B200_LIB; else export PANTR_B200_LIB=$PWD/$lib; fi
  timeout 200 python tests/quick_quad_condensed.py speed3 2>&1 | grep -v "^$" | awk 'NR%2==0' | tee -a gpurun_out/r02m_variants.log
  if [ $lib = default ] || [ $lib = build_variants/libmap_plain.so ]; then
    for rep in 1 2; do timeout 200 python tests/quick_bench.py 60 16384 2>&1 | tail -1 | tee -a gpurun_out/r02m_variants.log; done
  fi
done
