#!/usr/bin/env python
"""Bench workload (4096 x N = 100 x 600 steps) with RMPC_STEPS_PER_LAUNCH closed-loop steps per kernel launch (env, read by librmpc):
    for s in 1 2 4 600; do RMPC_STEPS_PER_LAUNCH=$s python profiles/tools/steps_per_launch.py; done"""
import json, os, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "profiles" / "tools"))
from run_configs import workload
import torch
import robotic_mpc_b200 as rm
B, N, K = 4096, 100, 600
surf, z0 = workload(B, 44)
eng = rm.Engine(B, N, "SQP_RTI", dt=0.01, nlog=0)
eng.set_params(rm.pack_params(B, 0.01, surf=surf))
out = []
for rep in range(2):
    eng.set_state(z0); eng.reset(); torch.cuda.synchronize()
    l0 = eng.launch_count
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); eng.closed_loop(K); b.record(); torch.cuda.synchronize()
    out.append(B * K / (a.elapsed_time(b) * 1e-3))
s = eng.get("summary").cpu().numpy()
print(json.dumps({"steps_per_launch": int(os.environ.get("RMPC_STEPS_PER_LAUNCH", "1")), "launches": int(eng.launch_count - l0), "mpc_steps_per_s": out,
                  "qp_iters_per_step": float(s[:, rm.RS["QP_ITERS"]].sum() / (B * K)), "failures": int(s[:, rm.RS["NFAIL"]].sum())}))
