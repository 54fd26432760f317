#!/usr/bin/env python
"""The compiled reference (oracle/_ref: unmodified acados + HPIPM + BLASFEO) timed on the host cores for the BASELINE.json
configs that are NOT the bench line -- the denominators BASELINE.md 3 lacked (missing #3):

  config2   README grid: every (N, nlp_solver_type) group of the grid, N in {20, 30, 40} x {SQP, SQP_RTI}, >= 256 instances per
            group over the grid's w_u / w_qddot values and random surfaces, K closed-loop steps from the initial state
  config4   full SQP, N = 200, terminal ellipsoid: K closed-loop steps, each a full SQP solve
  config5   horizon sweep N in {10, 50, 100, 200, 300}, SQP_RTI, >= 256 instances per horizon, K closed-loop steps

One pinned single-threaded process per host core; throughput = (instances x steps) / wall time of the slowest process (wall
includes creating each instance's solver and the plant steps; ..._solve_time_only divides by the summed acados solve times);
latency = acados' own time_tot of every solve (AC/ocp_nlp/ocp_nlp_sqp.c:645, ocp_nlp_sqp_rti.c:1171), read by the C closed
loop of oracle/ref_driver.c.  Same workload generator as profiles/tools/run_configs.py (the GPU side).  CPU only.

    python profiles/tools/cpu_side_configs.py config2|config4|config5 [--per-group 256] [--steps K]
"""
import argparse
import json
import math
import os
import sys
import time
from multiprocessing import get_context
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))


def workload(B, seed):
    rng = np.random.default_rng(seed)
    mean = np.array([-0.1, 0.1, -0.01, 0.01, 0.01, 0.0])
    surf = mean + 0.01 * rng.standard_normal((B, 6))
    q0 = np.array([math.pi / 4, -math.pi / 3, math.pi / 4, -math.pi / 2, -math.pi / 2, 0.0])
    qd0 = np.array([1.0, 2.0, 1.0, 0.0, 0.0, 0.0])
    z0 = np.concatenate([q0 + rng.uniform(-0.05, 0.05, (B, 6)), qd0 + rng.uniform(-0.1, 0.1, (B, 6))], axis=1)
    return surf, np.ascontiguousarray(z0)


def _worker(job):
    core, items, steps = job
    from oracle import pyoracle as o

    if core is not None and hasattr(os, "sched_setaffinity"):
        os.sched_setaffinity(0, {core})
    leg = o.RefLeg()
    wall, tt, it, qp, fails = 0.0, [], 0, 0, 0
    for kw, z0 in items:
        p = o.make_problem(**kw)
        t0 = time.perf_counter()
        r = leg.closed_loop(p, z0, steps)
        wall += time.perf_counter() - t0
        tt.append(r["time_tot"])
        it += int(r["sqp_iter"].sum())
        qp += int(r["qp_iter"].sum())
        fails += int((r["status"] != 0).sum())
    return wall, np.stack(tt) if tt else np.zeros((0, steps)), it, qp, fails  # time_tot [instances of this process][steps]


def run_group(label, items, steps, per_step_detail=False):
    cpus = sorted(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else list(range(os.cpu_count() or 1))
    cores = len(cpus)
    jobs = [(cpus[c], items[c::cores], steps) for c in range(cores)]
    with get_context("fork").Pool(cores) as pool:
        res = pool.map(_worker, jobs)
    tmax = max(r[0] for r in res)
    tsolve = max(float(r[1].sum()) for r in res)  # acados' own solve time only: no solver creation, no plant step
    lat = np.concatenate([r[1].ravel() for r in res]) * 1e3
    n = len(items)
    per_step = [{"step": k, "solves_per_s_solve_time_only": n / max(float(r[1][:, k].sum()) for r in res),
                 "acados_time_tot_ms_p50": float(np.percentile(np.concatenate([r[1][:, k] for r in res]) * 1e3, 50)),
                 "acados_time_tot_ms_p99": float(np.percentile(np.concatenate([r[1][:, k] for r in res]) * 1e3, 99))}
                for k in range(steps)] if per_step_detail else None
    line = {"config": label, "impl": "reference (oracle/_ref, compiled acados + HPIPM + BLASFEO)", "cores": cores, "instances": n,
            "closed_loop_steps": steps, "wall_s_slowest_process": tmax, "mpc_steps_per_s": n * steps / tmax,
            "mpc_steps_per_s_solve_time_only": n * steps / tsolve,
            "acados_time_tot_ms": {"p50": float(np.percentile(lat, 50)), "p99": float(np.percentile(lat, 99)), "max": float(lat.max())},
            "sqp_iter_per_step": sum(r[2] for r in res) / (n * steps), "qp_iter_per_step": sum(r[3] for r in res) / (n * steps),
            "nonzero_status_steps": sum(r[4] for r in res)}
    if per_step:
        line["per_step"] = per_step
    print(json.dumps(line), flush=True)
    return line


def config2(args):
    from oracle import pyoracle as o

    K, per = args.steps or 100, args.per_group or 256
    tot_steps, tot_wall = 0, 0.0
    for N in (20, 30, 40):
        for st, nm in ((o.SQP, "SQP"), (o.SQP_RTI, "SQP_RTI")):
            surf, z0 = workload(per, 42 + 1)
            combos = [(wq, wu) for wq in (0.02, 0.05) for wu in (0.01, 0.001)]  # the grid's four weight tuples, round robin
            items = [(dict(N=N, solver_type=st, surf=surf[i], w_qddot=combos[i % 4][0], w_u=combos[i % 4][1]),
                      np.concatenate([o.README_Q0, o.README_QD0])) for i in range(per)]
            ln = run_group(f"configs[1] README grid group N={N} {nm}", items, K)
            tot_steps += per * K
            tot_wall += ln["wall_s_slowest_process"]
    print(json.dumps({"config": "configs[1] README grid, all six groups one after the other", "mpc_steps_per_s": tot_steps / tot_wall,
                      "note": "every group has the same number of instances, as in the grid (4 weight tuples x 256 surfaces each)"}))


def config4(args):
    from oracle import pyoracle as o
    import robotic_mpc_b200.terminal_constraint as tc

    K, per, N = args.steps or 3, args.per_group or 256, args.horizon or 200
    surf, z0 = workload(per, 42 + 3)
    P, alpha = tc.default_terminal_set(wcv=200.0, Ts=0.01, rho=100.0)
    items = [(dict(N=N, solver_type=o.SQP, surf=surf[i], term_P=P, term_alpha=alpha), z0[i]) for i in range(per)]
    run_group(f"configs[3] full SQP N={N} + terminal ellipsoid, {K} closed-loop steps (step 0 = the failing first QP, see run_configs.py)",
              items, K, per_step_detail=True)


def config5(args):
    from oracle import pyoracle as o

    K, per = args.steps or 20, args.per_group or 256
    for N in (10, 50, 100, 200, 300):
        surf, z0 = workload(per, 42 + 4 + N)
        items = [(dict(N=N, solver_type=o.SQP_RTI, surf=surf[i]), z0[i]) for i in range(per)]
        run_group(f"configs[4] horizon sweep N={N} SQP_RTI", items, K)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("which", choices=["config2", "config4", "config5"])
    ap.add_argument("--per-group", type=int, default=0)
    ap.add_argument("--steps", type=int, default=0)
    ap.add_argument("--horizon", type=int, default=0)
    a = ap.parse_args()
    {"config2": config2, "config4": config4, "config5": config5}[a.which](a)
