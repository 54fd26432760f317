#!/bin/bash
mkdir -p gpurun_out
python bench.py --impl reference --steps 100 --warmup 3 > gpurun_out/r1k_bench_ref.json 2>/dev/null
python bench.py > gpurun_out/r1k_bench_full.json 2> gpurun_out/r1k_bench_full.err; cut -c1-300 gpurun_out/r1k_bench_full.json
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r1k_smoke.log 2>&1; tail -1 gpurun_out/r1k_smoke.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1k_launches.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r1k_ncu_list.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:rmpc_closed_loop -s 40 -c 1 -o gpurun_out/r1k_full -f python scratch/prof_steps.py 1 42 > gpurun_out/r1k_ncu_full.log 2>&1
ncu -i gpurun_out/r1k_full.ncu-rep --page raw --csv > gpurun_out/r1k_full_raw.csv 2>/dev/null
ncu -i gpurun_out/r1k_full.ncu-rep --page source --csv > gpurun_out/r1k_full_source.csv 2>/dev/null
