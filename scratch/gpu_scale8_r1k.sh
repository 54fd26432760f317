#!/bin/bash
mkdir -p gpurun_out
N=$1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 600 --warmup 3 --no-cpu-baseline > gpurun_out/r1k_scale_$N.json 2> gpurun_out/r1k_scale_$N.err
python -c "
import json; d=json.loads(open('gpurun_out/r1k_scale_$N.json').read().strip().splitlines()[-1]); print('gpus',d['n_gpus'],'steps/s',round(d['value']),'ms',round(d['ms_per_step'],3),'e2e',round(d['e2e']['value']),'gather_ms',d.get('summary_gather_ms'))"
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 profiles/tools/run_config5_dist.py > gpurun_out/r1k_config5_dist_$N.json 2> gpurun_out/r1k_config5_dist_$N.err
tail -1 gpurun_out/r1k_config5_dist_$N.json | cut -c1-600; tail -2 gpurun_out/r1k_config5_dist_$N.err
