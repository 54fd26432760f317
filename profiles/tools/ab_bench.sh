#!/bin/bash
# A/B of library builds on the bench workload: profiles/tools/ab_bench.sh [steps] lib1.so lib2.so ...   (run on the GPU box)
steps=$1; shift
for lib in "$@"; do
  for rep in 1 2; do
    RMPC_LIB=$PWD/$lib python bench.py --steps $steps --warmup 5 --no-cpu-baseline --no-external 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$lib', 'steps/s %.0f'%d['value'], 'ms/step %.4f'%d['ms_per_step'], 'qp_it %.3f'%d['per_rank']['qp_iters_per_mpc_step'][0], 'fail', d['solver_failures'])"
  done
done
