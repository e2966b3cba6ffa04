#!/bin/bash
# round-2 call K: v7 with the per-model object placement: HC bench (no regression), ncu of the quadcopter kernel
mkdir -p gpurun_out
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r02k_bench.json 2> gpurun_out/r02k_bench.err; echo "bench rc=$?"; cut -c1-200 gpurun_out/r02k_bench.json
timeout 200 python tests/quick_quad_condensed.py speed3 2>&1 | grep -v "^$" | awk 'NR%2==0' | tee gpurun_out/r02k_quad_speed.log
timeout 600 bash scripts/gpu_ncu_quad.sh r02k_quadc quadcopter 2368
ncu -i gpurun_out/r02k_quadc.ncu-rep --page raw --csv 2>/dev/null | python tools/ncu_summary.py > gpurun_out/r02k_quadc_ncu_full_summary.csv
ncu -i gpurun_out/r02k_quadc.ncu-rep --page source --csv --print-source sass 2>/dev/null > gpurun_out/r02k_quadc_source_sass.csv; rm -f gpurun_out/r02k_quadc.ncu-rep
python tools/ncu_regions.py gpurun_out/r02k_quadc_source_sass.csv "v7 simple quadcopter, B = 2368, capped launch" > gpurun_out/r02k_quadc_ncu_regions.txt
