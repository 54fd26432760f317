#!/usr/bin/env python
"""Device time of one MPC step under the round-2 solver options (side measurement, not the bench line): standard RTI, partial
condensing (qp_cond_N = N / 5) and the AS-RTI levels, as PREPARATION and FEEDBACK phase times of one instance (device timers of
the step record, mean over instances and steps 10..39) and as batched wall time per step, at B = 1 (N = 20) and B = 4096 (N = 100).

    python profiles/tools/options_latency.py            # one JSON line per (B, N, option)
"""
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "profiles" / "tools"))
from run_configs import workload  # noqa: E402


def run(B, N, name, opts, steps=40, skip=10):
    import torch
    import robotic_mpc_b200 as rm

    surf, z0 = workload(B, 44)
    eng = rm.Engine(B, N, "SQP_RTI", dt=0.01, nlog=1)
    eng.set_params(rm.pack_params(B, 0.01, surf=surf))
    for k, v in opts.items():
        eng.set_option(k, v)
    eng.set_state(z0)
    eng.reset()
    T = rm.RL["TIME"]
    z = torch.as_tensor(z0, device="cuda")
    prep, fb, wall, qp = [], [], [], []
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    for k in range(steps):
        ev[0].record()
        eng.solve(rti_phase=1)
        ev[1].record()
        tp = eng.get("stats")[:, T].mean().item()
        eng.set("z", z)
        ev[1].record()
        eng.solve(rti_phase=2)
        ev[2].record()
        st = eng.get("stats")
        u = eng.get("u", 0)
        # plant: exact-ZOH step of the first-order joint model (the engine's own plant does not advance in phase calls)
        w, h = 200.0, 0.01
        a22 = float(np.exp(-w * h)); a12 = (1 - a22) / w
        q, qd = z[:, :6], z[:, 6:]
        z = torch.cat([q + a12 * qd + (h - a12) * u, a22 * qd + (1 - a22) * u], 1)
        torch.cuda.synchronize()
        if k >= skip:
            prep.append(tp)
            fb.append(st[:, T].mean().item())
            qp.append(st[:, rm.RL["QP_ITER"]].mean().item())
            wall.append(ev[1].elapsed_time(ev[2]))
    eng.close()
    return {"B": B, "N": N, "option": name, "preparation_ms_per_instance": 1e3 * float(np.mean(prep)),
            "feedback_ms_per_instance": 1e3 * float(np.mean(fb)), "feedback_batched_wall_ms": float(np.mean(wall)),
            "feedback_qp_iters": float(np.mean(qp)), "solver_failures": int((st[:, 0] != 0).sum().item())}


if __name__ == "__main__":
    for B, N in ((1, 20), (4096, 100)):
        for name, opts in (("standard RTI (preparation / feedback split)", {}), ("qp_cond_N = N / 5", {"qp_cond_N": N // 5}),
                           ("AS-RTI level A", {"as_rti_level": 0}), ("AS-RTI level B, 1 iteration", {"as_rti_iter": 1, "as_rti_level": 1}),
                           ("AS-RTI level C, 1 iteration", {"as_rti_iter": 1, "as_rti_level": 2}),
                           ("AS-RTI level D, 1 iteration", {"as_rti_iter": 1, "as_rti_level": 3})):
            print(json.dumps(run(B, N, name, opts)), flush=True)
