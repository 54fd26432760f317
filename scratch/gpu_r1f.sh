#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r1f_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1f_pytest.log
tail -3 gpurun_out/r1f_pytest.log
python bench.py --impl reference --steps 100 --warmup 3 > gpurun_out/r1f_bench_ref.json 2>/dev/null
python bench.py > gpurun_out/r1f_bench_full.json 2> gpurun_out/r1f_bench_full.err; cat gpurun_out/r1f_bench_full.json | cut -c1-700
RMPC_LIB=$PWD/robotic-mpc_b200/librmpc_w1.so python bench.py --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('warps/CTA 1','steps/s',round(d['value']),'ms/step',round(d['ms_per_step'],3),'p50',round(d['latency_ms']['p50'],3),'p99',round(d['latency_ms']['p99'],3))" | tee -a gpurun_out/r1f_cta1.txt
timeout 900 python profiles/tools/run_configs.py config2 > gpurun_out/r1f_config2.jsonl 2> gpurun_out/r1f_config2.err; cat gpurun_out/r1f_config2.jsonl; tail -3 gpurun_out/r1f_config2.err
timeout 900 python profiles/tools/run_configs.py config4 > gpurun_out/r1f_config4.jsonl 2> gpurun_out/r1f_config4.err; cut -c1-900 gpurun_out/r1f_config4.jsonl; tail -3 gpurun_out/r1f_config4.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1f_launches.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r1f_ncu_list.log 2>&1
