#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r1g_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1g_pytest.log
tail -3 gpurun_out/r1g_pytest.log
python bench.py --no-cpu-baseline > gpurun_out/r1g_bench.json 2> gpurun_out/r1g_bench.err; python -c "
import json; d=json.load(open('gpurun_out/r1g_bench.json')); print('steps/s',round(d['value']),'ms',round(d['ms_per_step'],3),d['latency_ms'],'e2e',round(d['e2e']['value']),'fail',d['solver_failures'])"
