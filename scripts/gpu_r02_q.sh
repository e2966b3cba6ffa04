#!/bin/bash
# round-2 call Q: speed A/B of three variants of the v8 quadcopter kernel (rows by TMA: same bits; fma axpys in the rollout and
# split fma chains in the recurrences: different rounding, speed only -- adopted with an oracle change only if they pay)
mkdir -p gpurun_out
for lib in default build_variants/libqc_rows_tma.so build_variants/libqc_fma_axpy.so build_variants/libqc_split_chain.so; do
  echo "=== $lib" | tee -a gpurun_out/r02q_variants.log
  if [ $lib = default ]; then unset PANTR_B200_LIB; else export PANTR_B200_LIB=$PWD/$lib; fi
  timeout 300 python tests/quick_quad_condensed.py speed4 2>&1 | grep -v "^$" | tee -a gpurun_out/r02q_variants.log
done
export PANTR_B200_LIB=$PWD/build_variants/libqc_rows_tma.so
timeout 500 python tests/quick_quad_condensed.py parity > gpurun_out/r02q_rows_tma_parity.log 2>&1; echo "rows_tma parity rc=$?"; grep -v bit-exact gpurun_out/r02q_rows_tma_parity.log | tail -3
