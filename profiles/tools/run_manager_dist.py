#!/usr/bin/env python
"""SimulationManager.run_all under torchrun (one process per GPU, NCCL): the README grid (24 option tuples x S surfaces) is
sharded over the ranks group by group, the summary rows are all-gathered; rank 0 then re-runs every group un-sharded on its
own GPU and checks that the gathered rows are bit-identical (instances do not depend on batch composition).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544 \
        profiles/tools/run_manager_dist.py [--surfaces 16] [--steps 100]
"""
import argparse
import json
import os
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))


def main():
    import torch
    import torch.distributed as dist
    import robotic_mpc_b200 as rm
    from robotic_mpc_b200.simulator import run_batch

    ap = argparse.ArgumentParser()
    ap.add_argument("--surfaces", type=int, default=16)
    ap.add_argument("--steps", type=int, default=100)
    a = ap.parse_args()
    rank, ws, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if ws > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rng = np.random.default_rng(43)
    mean = {"a": -0.1, "b": 0.1, "c": -0.01, "d": 0.01, "e": 0.01, "f": 0.0}
    sets = [{k: float(v + 0.01 * rng.standard_normal()) for k, v in mean.items()} for _ in range(a.surfaces)]
    qdm = np.array([2.16, 2.16, np.pi, 3.20, 3.20, 3.20])
    base = dict(robot_name="ur10", dt=0.01, simulation_time=a.steps * 0.01, prediction_horizon=20,
                q_0=np.array([np.pi / 4, -np.pi / 3, np.pi / 4, -np.pi / 2, -np.pi / 2, 0.0]), qdot_0=np.array([1.0, 2, 1, 0, 0, 0]),
                wcv=np.full(6, 200.0), q_min=np.full(6, -2 * np.pi), q_max=np.full(6, 2 * np.pi), qdot_min=-qdm, qdot_max=qdm,
                surface_coeffs=mean, scene=False)
    mgr = rm.SimulationManager(base)
    mgr.grid_search({"prediction_horizon": [20, 30, 40], "w_qddot": [0.02, 0.05], "w_u": [0.01, 0.001],
                     "solver_options.nlp_solver_type": ["SQP", "SQP_RTI"]}, surface_coeff_sets=sets)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    mgr.run_all(return_results=False, keep_logs=False)
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    S = mgr.last_summaries
    if rank == 0:
        sims = mgr.build_simulators()
        plan = mgr.plan_groups(sims, 1)
        ref = np.zeros_like(S)
        for g in plan:
            ref[g["indices"]] = run_batch([sims[i] for i in g["indices"]], device=local, keep_logs=False)
        cols = [c for c in range(S.shape[1])]
        same = bool(np.array_equal(S[:, :10], ref[:, :10]) and np.array_equal(S[:, rm.RS["SQP_ITERS"]], ref[:, rm.RS["SQP_ITERS"]])
                    and np.array_equal(S[:, rm.RS["QP_ITERS"]], ref[:, rm.RS["QP_ITERS"]]))
        print(json.dumps({"what": "SimulationManager.run_all sharded over ranks vs un-sharded on rank 0", "world_size": ws,
                          "instances": len(sims), "closed_loop_steps": a.steps, "run_info": mgr.last_run_info, "wall_s": wall,
                          "gathered_rows_bit_identical_to_unsharded": same, "failures": int(S[:, rm.RS["NFAIL"]].sum()),
                          "max_abs_diff_sse": float(np.abs(S[:, :5] - ref[:, :5]).max())}))
    if ws > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
