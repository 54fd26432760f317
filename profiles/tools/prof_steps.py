#!/usr/bin/env python
"""The bench workload (4096 instances, N = 100, SQP_RTI) for a profiler: 45 closed-loop steps, one launch each, so that
`ncu -k regex:rmpc_closed_loop -s 40 -c 1` captures a launch past the start-up transient.

    ncu --set full --clock-control none --import-source on -k regex:rmpc_closed_loop -s 40 -c 1 -o gpurun_out/prof python profiles/tools/prof_steps.py
"""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "profiles" / "tools"))
from run_configs import workload  # noqa: E402

import torch  # noqa: E402
import robotic_mpc_b200 as rm  # noqa: E402

B, N, K = 4096, 100, int(sys.argv[1]) if len(sys.argv) > 1 else 45
surf, z0 = workload(B, 44)
eng = rm.Engine(B, N, "SQP_RTI", dt=0.01, nlog=K)
eng.set_params(rm.pack_params(B, 0.01, surf=surf))
eng.set_state(z0)
eng.reset()
eng.closed_loop(K)
torch.cuda.synchronize()
sl = eng.get("slog").cpu().numpy()
print("steps", K, "qp iterations of step 41:", sl[:, min(40, K - 1), rm.RL["QP_ITER"]].mean())
