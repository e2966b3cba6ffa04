#!/bin/bash
# round-2 final call R1: full GPU test suite, default bench (reference arm first, like the driver), ncu launch list of the
# bench command, ncu --set full capture of the timed hanging-chain launch (-> profiles/traffic.json)
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r02r_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r02r_pytest.log | cut -c1-250
timeout 600 python bench.py --impl reference > gpurun_out/r02r_bench_ref.json 2> gpurun_out/r02r_bench_ref.err; echo "ref rc=$?"; cut -c1-300 gpurun_out/r02r_bench_ref.json
timeout 900 python bench.py > gpurun_out/r02r_bench.json 2> gpurun_out/r02r_bench.err; echo "bench rc=$?"; cut -c1-600 gpurun_out/r02r_bench.json; tail -3 gpurun_out/r02r_bench.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02r_launches.csv python bench.py --no-cpu-baseline > gpurun_out/r02r_ncu_launches.log 2>&1; echo "launch list rc=$?"; wc -l gpurun_out/r02r_launches.csv
timeout 900 ncu --set full --clock-control none -k regex:solve_kernel -s 4 -c 1 -o gpurun_out/r02r_hc python bench.py --no-cpu-baseline --no-closed-loop --steps 1 --warmup 3 > gpurun_out/r02r_ncu_hc.log 2>&1; echo "ncu hc rc=$?"
ncu -i gpurun_out/r02r_hc.ncu-rep --page raw --csv 2>/dev/null > gpurun_out/r02r_hc_raw.csv
python tools/ncu_summary.py < gpurun_out/r02r_hc_raw.csv > gpurun_out/r02r_hc_ncu_full_summary.csv; rm -f gpurun_out/r02r_hc.ncu-rep
head -c 600 gpurun_out/r02r_hc_ncu_full_summary.csv
