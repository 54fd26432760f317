"""Per-instance solve-time / iteration distribution inside steady-state launches of the bench workload."""
import json, sys
import numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "profiles/tools")
import torch
import robotic_mpc_b200 as rm
from run_configs import workload
B, N, K = 4096, 100, 140
surf, z0 = workload(B, 42 + 2)
eng = rm.Engine(B, N, "SQP_RTI", dt=0.01, nlog=K)
eng.set_params(rm.pack_params(B, 0.01, surf=surf)); eng.set_state(z0); eng.reset()
ms = []
for k in range(K):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); eng.closed_loop(1); e1.record(); e1.synchronize(); ms.append(e0.elapsed_time(e1))
sl = eng.get("slog").cpu().numpy()
for k in (0, 1, 2, 5, 20, 60, 100, 139):
    t = sl[:, k, rm.RL["TIME"]] * 1e3
    q = sl[:, k, rm.RL["QP_ITER"]].astype(int)
    print(json.dumps({"step": k, "launch_ms": ms[k], "inst_ms": {p: float(np.percentile(t, p)) for p in (1, 10, 50, 90, 99, 100)},
                      "inst_ms_mean": float(t.mean()), "qp_iter_hist": {int(a): int(b) for a, b in zip(*np.unique(q, return_counts=True))}}))
print(json.dumps({"ms_mean_20_140": float(np.mean(ms[20:])), "qp_iter_mean_20_140": float(sl[:, 20:, rm.RL["QP_ITER"]].mean())}))
