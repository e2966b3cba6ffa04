#!/bin/bash
# round-2 final call R2: bench lines of BASELINE configs[2..4] (quadcopter at the full 65536, realistic sweep and mixed at
# reduced batches), ncu --set full capture + per-function table of the final quadcopter kernels (capped launch, one wave)
mkdir -p gpurun_out
timeout 1500 python bench.py --config quadcopter --steps 1 --warmup 1 --cpu-seconds 20 > gpurun_out/r02r_bench_quadcopter_65536.json 2> gpurun_out/r02r_bench_quadcopter.err; echo "quad bench rc=$?"; cut -c1-700 gpurun_out/r02r_bench_quadcopter_65536.json; tail -3 gpurun_out/r02r_bench_quadcopter.err
timeout 600 bash scripts/gpu_ncu_quad.sh r02r_quadc quadcopter 2368
ncu -i gpurun_out/r02r_quadc.ncu-rep --page raw --csv 2>/dev/null | python tools/ncu_summary.py > gpurun_out/r02r_quadc_ncu_full_summary.csv
ncu -i gpurun_out/r02r_quadc.ncu-rep --page source --csv --print-source sass 2>/dev/null > gpurun_out/r02r_quadc_source_sass.csv; rm -f gpurun_out/r02r_quadc.ncu-rep
python tools/ncu_regions.py gpurun_out/r02r_quadc_source_sass.csv "v9 simplified quadcopter, B = 2368, iteration-capped launch" > gpurun_out/r02r_quadc_ncu_regions.txt; rm -f gpurun_out/r02r_quadc_source_sass.csv
timeout 600 bash scripts/gpu_ncu_quad.sh r02r_rquadc realistic_quadcopter 1776
ncu -i gpurun_out/r02r_rquadc.ncu-rep --page raw --csv 2>/dev/null | python tools/ncu_summary.py > gpurun_out/r02r_rquadc_ncu_full_summary.csv; rm -f gpurun_out/r02r_rquadc.ncu-rep
timeout 1200 python bench.py --config mixed --batch 12288 --steps 1 --warmup 1 --cpu-seconds 20 > gpurun_out/r02r_bench_mixed_12288.json 2> gpurun_out/r02r_bench_mixed.err; echo "mixed bench rc=$?"; cut -c1-500 gpurun_out/r02r_bench_mixed_12288.json; tail -3 gpurun_out/r02r_bench_mixed.err
timeout 1500 python bench.py --config realistic_sweep --batch 1024 --steps 1 --warmup 1 --cpu-seconds 20 > gpurun_out/r02r_bench_realistic_sweep_1024.json 2> gpurun_out/r02r_bench_sweep.err; echo "sweep bench rc=$?"; cut -c1-500 gpurun_out/r02r_bench_realistic_sweep_1024.json; tail -3 gpurun_out/r02r_bench_sweep.err
