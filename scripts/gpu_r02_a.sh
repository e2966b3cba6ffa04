#!/bin/bash
# round-2 call A: bring-up of the condensed quadcopter path (parity + speed), CPU baselines of the quadcopter configs,
# closed-loop throughput of the hanging chain (the reference's own metric), sanitizer pass
mkdir -p gpurun_out
nproc > gpurun_out/r02_nproc.txt; lscpu | head -20 >> gpurun_out/r02_nproc.txt
timeout 600 python tests/quick_quad_condensed.py parity > gpurun_out/r02a_quad_parity.log 2>&1; echo "parity rc=$?"; tail -25 gpurun_out/r02a_quad_parity.log
timeout 900 python tests/quick_quad_condensed.py speed > gpurun_out/r02a_quad_speed.log 2>&1; echo "speed rc=$?"; cat gpurun_out/r02a_quad_speed.log
timeout 600 python scripts/cpu_quad_baseline.py > gpurun_out/r02_cpu_quad_baseline.log 2>&1; echo "cpu quad rc=$?"; tail -8 gpurun_out/r02_cpu_quad_baseline.log
for warm in 1 0; do
  timeout 600 python -m pantr_cdc2023_experiments_b200.mpc_driver /tmp/hanging_chain 60 ss2p $(python -c "import pantr_cdc2023_experiments_b200 as pb; print(' '.join(pb.benchmark_options('hanging_chain')))") batch=16384 num_sim=60 warm=$warm results_name=gpurun_out/r02_closed_loop_warm$warm > gpurun_out/r02_closed_loop_warm$warm.log 2>&1
  echo "closed loop warm=$warm rc=$?"; grep -E "time:|batch:" gpurun_out/r02_closed_loop_warm$warm.log
done
rm -f gpurun_out/*.npz
( timeout 1500 bash scripts/gpu_sanitize.sh r02a ) 2>&1 | tail -8
