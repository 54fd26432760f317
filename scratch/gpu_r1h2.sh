#!/bin/bash
mkdir -p gpurun_out
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:rmpc_closed_loop -s 40 -c 1 -o gpurun_out/r1h_full -f python scratch/prof_steps.py 1 42 > gpurun_out/r1h_ncu_full.log 2>&1
ncu -i gpurun_out/r1h_full.ncu-rep --page raw --csv > gpurun_out/r1h_full_raw.csv 2>/dev/null
ncu -i gpurun_out/r1h_full.ncu-rep --page source --csv > gpurun_out/r1h_full_source.csv 2>/dev/null
tail -3 gpurun_out/r1h_ncu_full.log
