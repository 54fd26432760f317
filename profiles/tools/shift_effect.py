#!/usr/bin/env python
"""Side measurement: what the "shift" option (primal shift of the initial guess between closed-loop steps) does on the bench
workload (4096 x N = 100 x SQP_RTI x 600 steps): throughput, interior-point iterations per step, tracking RMSE.  The reference
never shifts, so the bench line itself runs with shift = 0."""
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(Path(__file__).resolve().parent))


def main():
    import torch
    import robotic_mpc_b200 as rm
    from run_configs import workload

    B, N, K = 4096, 100, 600
    surf, z0 = workload(B, 42 + 2)
    par = rm.pack_params(B, 0.01, surf=surf)
    for shift in (0, 1):
        eng = rm.Engine(B, N, "SQP_RTI", dt=0.01, nlog=0)
        eng.set_params(par)
        eng.set_option("shift", shift)
        eng.set_state(z0)
        eng.reset()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        eng.closed_loop(K)
        e1.record()
        e1.synchronize()
        ms = e0.elapsed_time(e1)
        s = eng.get("summary").cpu().numpy()
        rmse = np.sqrt(s[:, 0:5] / s[:, [rm.RS["NSAMPLES"]]])
        print(json.dumps({"shift": shift, "mpc_steps_per_s": B * K / ms * 1e3, "ms_per_step": ms / K,
                          "qp_iters_per_step": float(s[:, rm.RS["QP_ITERS"]].sum() / (B * K)), "failures": int(s[:, rm.RS["NFAIL"]].sum()),
                          "rmse_e1_to_e5_mean": [float(v) for v in rmse.mean(axis=0)]}))
        eng.close()


if __name__ == "__main__":
    main()
