#!/usr/bin/env python
"""BASELINE configs[4]: 65536 instances, horizon sweep N in {10, 50, 100, 200, 300}, SQP_RTI, 600 closed-loop steps, sharded
over the GPUs of one box (one process per GPU, torchrun); the only collectives are the gather of the [B, 20] summary rows
and the per-horizon reduction of tracking-error statistics (NCCL over NVLink), both timed separately.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 \
        profiles/tools/run_config5_dist.py [--total 65536] [--steps 600]

Each rank takes a contiguous chunk of every horizon group.  Rank 0 prints one JSON line.
"""
import argparse
import json
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(Path(__file__).resolve().parent))


def main():
    import torch
    import torch.distributed as dist

    import robotic_mpc_b200 as rm
    from robotic_mpc_b200 import distributed as rdist
    from run_configs import workload

    ap = argparse.ArgumentParser()
    ap.add_argument("--total", type=int, default=65536)
    ap.add_argument("--steps", type=int, default=600)
    ap.add_argument("--concurrent", type=int, default=1, help="1: the five horizon groups run at the same time on separate streams")
    a = ap.parse_args()
    rank, ws, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if ws > 1:
        dist.init_process_group("nccl", device_id=dev)
    horizons = [10, 50, 100, 200, 300]
    per_h = [a.total // 5 + (1 if i < a.total % 5 else 0) for i in range(5)]
    K = a.steps
    ap_conc = a.concurrent
    rows, grp, ms_h = [], [], []
    if ws > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    if ap_conc:
        # all five horizon groups of this rank in flight at once, one CUDA stream each, interleaved 25 steps at a time: no
        # group fills the GPU alone (1638 instances of 4736 resident) and the long horizons would otherwise run by themselves
        cur = torch.cuda.current_stream()
        runs = []
        for hi, (N, Bh) in enumerate(zip(horizons, per_h)):
            b, e = rdist.shard_bounds(Bh, ws)[rank]
            surf, z0 = workload(Bh, 42 + 4 + N)
            st = torch.cuda.Stream()
            st.wait_stream(cur)
            with torch.cuda.stream(st):
                eng = rm.Engine(e - b, N, "SQP_RTI", dt=0.01, nlog=0, device=local)
                eng.set_params(rm.pack_params(e - b, 0.01, surf=surf[b:e]))
                eng.set_state(z0[b:e])
                eng.reset()
            runs.append((eng, st, hi, e - b))
        for k0 in range(0, K, 25):
            for eng, st, hi, n in runs:
                with torch.cuda.stream(st):
                    eng.closed_loop(min(25, K - k0))
        for eng, st, hi, n in runs:
            with torch.cuda.stream(st):
                rows.append(eng.get("summary"))
            st.synchronize()
            cur.wait_stream(st)
            grp.append(np.full(n, hi))
            ms_h.append(None)
            eng.close()
    else:
        for hi, (N, Bh) in enumerate(zip(horizons, per_h)):
            b, e = rdist.shard_bounds(Bh, ws)[rank]
            surf, z0 = workload(Bh, 42 + 4 + N)  # the same global workload on every rank, each takes its chunk
            eng = rm.Engine(e - b, N, "SQP_RTI", dt=0.01, nlog=0, device=local)
            eng.set_params(rm.pack_params(e - b, 0.01, surf=surf[b:e]))
            eng.set_state(z0[b:e])
            eng.reset()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            eng.closed_loop(K)
            e1.record()
            e1.synchronize()
            ms_h.append(e0.elapsed_time(e1))
            rows.append(eng.get("summary"))
            grp.append(np.full(e - b, hi))
            eng.close()
    t1.record()
    t1.synchronize()
    solve_ms = torch.tensor([t0.elapsed_time(t1)], dtype=torch.float64, device=dev)
    if ws > 1:
        dist.all_reduce(solve_ms, op=dist.ReduceOp.MAX)
    # collectives: summary gather per horizon group + statistics reduction
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record()
    gathered = [rdist.all_gather_rows(r, Bh, device=dev) for r, Bh in zip(rows, per_h)]
    g1.record()
    g1.synchronize()
    local_rows = torch.cat(rows).cpu().numpy()
    rmse1 = np.sqrt(local_rows[:, 0] / local_rows[:, rm.RS["NSAMPLES"]])
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s0.record()
    stats = rdist.reduce_group_stats(rmse1, np.concatenate(grp), 5, device=dev)
    s1.record()
    s1.synchronize()
    if rank == 0:
        allrows = np.concatenate(gathered)
        print(json.dumps({
            "config": "configs[4] horizon sweep N in {10,50,100,200,300}, SQP_RTI, %d steps" % K, "instances": a.total, "n_gpus": ws, "concurrent_groups": bool(a.concurrent),
            "solve_ms_max_over_ranks": float(solve_ms.item()), "mpc_steps_per_s": a.total * K / (float(solve_ms.item()) * 1e-3),
            "per_horizon_ms_rank0": dict(zip(map(str, horizons), ms_h)),
            "summary_gather_ms": g0.elapsed_time(g1), "stats_reduce_ms": s0.elapsed_time(s1),
            "failures": int(allrows[:, rm.RS["NFAIL"]].sum()), "gathered_rows": int(allrows.shape[0]),
            "rmse_e1_per_horizon": {str(N): {"count": int(stats["count"][i]), "mean": float(stats["mean"][i]), "var": float(stats["var"][i]),
                                             "min": float(stats["min"][i]), "max": float(stats["max"][i])} for i, N in enumerate(horizons)}}))
    if ws > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
