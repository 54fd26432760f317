#!/usr/bin/env python
"""Where does the time of a sweep go when the GPU fills up?  Per-stage time of each interior-point phase (device timers of
the step record, RL_TIME_QP_*) as a function of the batch size B (SM load and HBM load grow together) and of the horizon N
at full batch (N = 4: the work arrays of all 4096 instances fit in the 126 MB L2, so the SM load is that of the bench
workload while HBM is idle).

    python profiles/tools/contention_probe.py            # prints one JSON line per (B, N)
"""
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "profiles" / "tools"))
from run_configs import workload  # noqa: E402


def probe(B, N, steps=40, skip=20):
    import torch
    import robotic_mpc_b200 as rm
    from robotic_mpc_b200.params import RL

    surf, z0 = workload(B, 44)
    eng = rm.Engine(B, N, "SQP_RTI", dt=0.01, nlog=steps)
    eng.set_params(rm.pack_params(B, 0.01, surf=surf))
    eng.set_state(z0)
    eng.reset()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    ev[0].record()
    for k in range(steps):
        eng.closed_loop(1)
        ev[k + 1].record()
    torch.cuda.synchronize()
    launch_ms = np.array([ev[k].elapsed_time(ev[k + 1]) for k in range(steps)])[skip:]
    sl = eng.get("slog").cpu().numpy()[:, skip:]
    it = sl[..., RL["QP_ITER"]]
    out = {"B": B, "N": N, "launch_ms_p50": float(np.median(launch_ms)), "launch_ms_mean": float(launch_ms.mean()),
           "qp_iter_mean": float(it.mean()), "qp_iter_max_mean": float(it.max(axis=0).mean()),
           "time_tot_ms_mean": float(1e3 * sl[..., RL["TIME"]].mean())}
    # per-stage microseconds of one sweep: factor (1 per iteration), forward (2), backward (1), update + residual (1 + 1 / QP)
    n_it = it.sum()
    out["us_per_stage"] = {
        "factor": float(1e6 * sl[..., RL["TIME_QP_FACT"]].sum() / (n_it * N)),
        "forward": float(1e6 * sl[..., RL["TIME_QP_FWD"]].sum() / (2 * n_it * N)),
        "backward": float(1e6 * sl[..., RL["TIME_QP_BWD"]].sum() / (n_it * N)),
        "update_residual": float(1e6 * sl[..., RL["TIME_QP_RES"]].sum() / ((n_it + it.size) * N)),
        "linearise_per_stage": float(1e6 * sl[..., RL["TIME_LIN"]].sum() / (it.size * N)),
        "globalisation_per_stage": float(1e6 * sl[..., RL["TIME_GLOB"]].sum() / (it.size * N)),
    }
    eng.close()
    return out


if __name__ == "__main__":
    for B, N in [(1, 100), (64, 100), (592, 100), (1184, 100), (2368, 100), (4096, 100), (4736, 100), (8192, 100),
                 (4096, 4), (4096, 8), (4096, 20), (4096, 50), (4096, 200)]:
        print(json.dumps(probe(B, N)), flush=True)
