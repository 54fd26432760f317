#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r1i_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1i_pytest.log
tail -3 gpurun_out/r1i_pytest.log
python bench.py --impl reference --steps 100 --warmup 3 > gpurun_out/r1i_bench_ref.json 2>/dev/null
python bench.py > gpurun_out/r1i_bench_full.json 2> gpurun_out/r1i_bench_full.err; cut -c1-300 gpurun_out/r1i_bench_full.json
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r1i_smoke.log 2>&1; tail -1 gpurun_out/r1i_smoke.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1i_launches.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r1i_ncu_list.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:rmpc_closed_loop -s 40 -c 1 -o gpurun_out/r1i_full -f python scratch/prof_steps.py 1 42 > gpurun_out/r1i_ncu_full.log 2>&1
ncu -i gpurun_out/r1i_full.ncu-rep --page raw --csv > gpurun_out/r1i_full_raw.csv 2>/dev/null
ncu -i gpurun_out/r1i_full.ncu-rep --page source --csv > gpurun_out/r1i_full_source.csv 2>/dev/null
timeout 900 python profiles/tools/run_configs.py config2 > gpurun_out/r1i_config2.jsonl 2>/dev/null; cut -c1-400 gpurun_out/r1i_config2.jsonl
timeout 900 python profiles/tools/run_configs.py config5 > gpurun_out/r1i_config5.jsonl 2>/dev/null; cut -c1-200 gpurun_out/r1i_config5.jsonl
timeout 900 python profiles/tools/run_configs.py config4 > gpurun_out/r1i_config4.jsonl 2>/dev/null; cut -c1-300 gpurun_out/r1i_config4.jsonl
