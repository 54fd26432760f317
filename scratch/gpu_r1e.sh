#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r1e_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1e_pytest.log
tail -5 gpurun_out/r1e_pytest.log
for w in 2 3; do RMPC_LIB=$PWD/robotic-mpc_b200/librmpc_w$w.so python bench.py --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('warps/CTA $w','steps/s',round(d['value']),'ms/step',round(d['ms_per_step'],3),'p50',round(d['latency_ms']['p50'],3),'p99',round(d['latency_ms']['p99'],3))"; done > gpurun_out/r1e_cta_shape.txt 2>&1
python bench.py --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('warps/CTA 4','steps/s',round(d['value']),'ms/step',round(d['ms_per_step'],3),'p50',round(d['latency_ms']['p50'],3),'p99',round(d['latency_ms']['p99'],3))" >> gpurun_out/r1e_cta_shape.txt 2>&1
cat gpurun_out/r1e_cta_shape.txt
timeout 600 python profiles/tools/run_configs.py config4 --batch 2048 > gpurun_out/r1e_config4_small.jsonl 2> gpurun_out/r1e_config4_small.err; cat gpurun_out/r1e_config4_small.jsonl; tail -3 gpurun_out/r1e_config4_small.err
timeout 900 python profiles/tools/run_configs.py config5 > gpurun_out/r1e_config5.jsonl 2> gpurun_out/r1e_config5.err; cat gpurun_out/r1e_config5.jsonl; tail -3 gpurun_out/r1e_config5.err
timeout 900 python profiles/tools/run_configs.py config2 > gpurun_out/r1e_config2.jsonl 2> gpurun_out/r1e_config2.err; cat gpurun_out/r1e_config2.jsonl; tail -3 gpurun_out/r1e_config2.err
