#!/bin/bash
mkdir -p gpurun_out
for R in 0 4 8 16 64; do
  echo "RMPC_OVERLAP=$R"; RMPC_OVERLAP=$R timeout 300 python bench.py --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d.get('latency_ms'), d['e2e']['value'])"
done
RMPC_OVERLAP=8 timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -5
