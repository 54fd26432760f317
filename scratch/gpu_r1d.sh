#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r1d_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1d_pytest.log
tail -15 gpurun_out/r1d_pytest.log
for B in 2048 3072 4096 4736 8192; do python bench.py --batch $B --steps 200 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('B',d['config']['batch_per_gpu'],'steps/s',round(d['value']),'ms/step',round(d['ms_per_step'],3),'p99',round(d['latency_ms']['p99'],3))"; done > gpurun_out/r1d_batch_sweep.txt 2>&1
cat gpurun_out/r1d_batch_sweep.txt
