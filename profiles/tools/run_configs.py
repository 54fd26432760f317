#!/usr/bin/env python
"""Side measurements of the BASELINE.json configs that are NOT the bench line (bench.py = configs[2]):

  config1   B = 1, N = 20, SQP_RTI, flat surface: single-instance step latency on the device beside one host core
  config2   README grid search, 24 option tuples x 256 random surfaces = 6144 instances, 600 closed-loop steps, through
            SimulationManager.grid_search / run_all (one batched run per (N, solver) group), summaries only
  config4   full SQP, N = 200, terminal ellipsoid, B instances: K closed-loop steps, each one full SQP solve of every
            instance; per step: solves/s and the per-solve device latency distribution (RL_TIME of each instance)
  config5   horizon sweep N in {10, 50, 100, 200, 300}, SQP_RTI, 600 steps: this GPU's 1/8 share of the 65536 instances
            (8192 instances, 1638/1639 per horizon) unless --batch says otherwise

    python profiles/tools/run_configs.py config1|config2|config4|config5 [--batch B] [--steps K]

Prints one JSON line per measurement.  Needs a GPU.
"""
import argparse
import json
import math
import sys
import time
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


def config1(args):
    """B = 1, N = 20, SQP_RTI, flat surface, README initial state, 600 steps: per-step latency of ONE instance on the device
    (a group of 16 lanes on one SM; everything else idle) beside the compiled reference on one host core."""
    import torch
    import robotic_mpc_b200 as rm
    from oracle import pyoracle as o

    K = args.steps or 600
    z0 = np.concatenate([o.README_Q0, o.README_QD0])[None]
    par = rm.pack_params(1, 0.01, surf=np.zeros((1, 6)))
    eng = rm.Engine(1, 20, "SQP_RTI", dt=0.01, nlog=K)
    eng.set_params(par)
    out = {}
    for mode in ("one_launch_per_step", "one_call"):
        eng.set_state(z0)
        eng.reset()
        torch.cuda.synchronize()
        ms = []
        t0 = time.perf_counter()
        if mode == "one_call":
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); eng.closed_loop(K); e1.record(); e1.synchronize()
            ms = [e0.elapsed_time(e1) / K]
        else:
            for _ in range(K):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); eng.closed_loop(1); e1.record(); e1.synchronize()
                ms.append(e0.elapsed_time(e1))
        wall = time.perf_counter() - t0
        sl = eng.get("slog").cpu().numpy()[0]
        out[mode] = {"ms_per_step_mean": float(np.mean(ms)), "p50": float(np.percentile(ms, 50)), "p99": float(np.percentile(ms, 99)),
                     "wall_s": wall, "device_solve_ms_mean": float(sl[:, rm.RL["TIME"]].mean() * 1e3)}
    s = eng.get("summary").cpu().numpy()[0]
    eng.close()
    leg = o.RefLeg() if o.REF_SO.exists() else o.PortLeg()
    t0 = time.perf_counter()
    r = leg.closed_loop(o.make_problem(N=20, solver_type=o.SQP_RTI, surf=np.zeros(6)), z0[0], K)
    cpu_wall = time.perf_counter() - t0
    print(json.dumps({"config": "configs[0] single run B=1 N=20 SQP_RTI flat surface, %d steps" % K, "gpu": out,
                      "gpu_failures": int(s[rm.RS["NFAIL"]]), "gpu_rmse_e1": float(np.sqrt(s[0] / s[rm.RS["NSAMPLES"]])),
                      "cpu_reference_one_core": {"kind": "reference" if o.REF_SO.exists() else "port", "ms_per_step_wall": cpu_wall / K * 1e3,
                                                 "ms_per_step_acados_timer": float(r["time_tot"].mean() * 1e3), "failures": int(r["nfail"])}}))


def config2(args):
    import torch
    import robotic_mpc_b200 as rm

    nsurf = args.batch or 256
    rng = np.random.default_rng(42 + 1)
    mean = {"a": -0.1, "b": 0.1, "c": -0.01, "d": 0.01, "e": 0.01, "f": 0.0}
    sets = [{k: float(v + 0.01 * rng.standard_normal()) for k, v in mean.items()} for _ in range(nsurf)]
    qdm = np.array([2.16, 2.16, np.pi, 3.20, 3.20, 3.20])
    base = dict(robot_name="ur10", dt=0.01, simulation_time=args.steps * 0.01 if args.steps else 6.0, prediction_horizon=20,
                q_0=np.array([np.pi / 4, -np.pi / 3, np.pi / 4, -np.pi / 2, -np.pi / 2, 0.0]), qdot_0=np.array([1.0, 2, 1, 0, 0, 0]),
                wcv=np.full(6, 200.0), q_min=np.full(6, -2 * np.pi), q_max=np.full(6, 2 * np.pi), qdot_min=-qdm, qdot_max=qdm,
                surface_coeffs=mean, scene=False)
    mgr = rm.SimulationManager(base)
    mgr.grid_search({"prediction_horizon": [20, 30, 40], "w_qddot": [0.02, 0.05], "w_u": [0.01, 0.001],
                     "solver_options.nlp_solver_type": ["SQP", "SQP_RTI"]}, surface_coeff_sets=sets)
    n = len(mgr.simulations)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    mgr.run_all(return_results=False)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    S = mgr.last_summaries
    K = int(base["simulation_time"] / 0.01)
    print(json.dumps({"config": "configs[1] README grid search", "instances": n, "closed_loop_steps": K,
                      "run_info": mgr.last_run_info,
                      "wall_s": dt, "mpc_steps_per_s_incl_host_setup": n * K / dt,
                      "failures": int(S[:, rm.RS["NFAIL"]].sum()), "sqp_iters_total": int(S[:, rm.RS["SQP_ITERS"]].sum()),
                      "qp_iters_total": int(S[:, rm.RS["QP_ITERS"]].sum())}))


def config4(args):
    import torch
    import robotic_mpc_b200 as rm
    import robotic_mpc_b200.terminal_constraint as tc

    B, N, K = args.batch or 16384, args.horizon or 200, args.steps or 5
    surf, z0 = workload(B, 42 + 3)
    P, alpha = tc.default_terminal_set(wcv=200.0, Ts=0.01, rho=args.rho)
    eng = rm.Engine(B, N, "SQP", dt=0.01, nlog=K)
    eng.set_params(rm.pack_params(B, 0.01, surf=surf))
    eng.set_terminal(P, alpha)
    eng.set_state(z0)
    eng.reset()
    torch.cuda.synchronize()
    ms = []
    for k in range(K):  # every closed-loop step is one full SQP solve of every instance (<= 100 iterations, tol 1e-6)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        eng.closed_loop(1)
        e1.record()
        e1.synchronize()
        ms.append(e0.elapsed_time(e1))
    sl = eng.get("slog").cpu().numpy()
    xN = eng.get("x", N).cpu().numpy()
    h = np.einsum("bi,ij,bj->b", xN, P, xN)
    for k in range(K):
        lat = sl[:, k, rm.RL["TIME"]] * 1e3
        print(json.dumps({
            "config": "configs[3] full SQP N=%d + terminal ellipsoid (rho=%g, alpha=%.4g), closed-loop step %d" % (N, args.rho, alpha, k),
            "note": "step 0 starts from the flat initial guess: its first QP fails in the reference too (status 4, 1 SQP / 3 QP "
                    "iterations, checked against the compiled acados) and the solve proper happens at step 1" if k == 0 else "",
            "instances": B, "batch_ms": ms[k], "solves_per_s": B / (ms[k] * 1e-3),
            "per_solve_latency_ms": {"p50": float(np.percentile(lat, 50)), "p99": float(np.percentile(lat, 99)),
                                     "max": float(lat.max()), "what": "device time of each instance's own solve (RL_TIME)"},
            "status_hist": {int(a): int(b) for a, b in zip(*np.unique(sl[:, k, 0].astype(int), return_counts=True))},
            "sqp_iter_mean": float(sl[:, k, rm.RL["SQP_ITER"]].mean()), "sqp_iter_max": int(sl[:, k, rm.RL["SQP_ITER"]].max()),
            "qp_iter_mean": float(sl[:, k, rm.RL["QP_ITER"]].mean()),
            "max_kkt_res": [float(v) for v in sl[:, k, rm.RL["RES"]:rm.RL["RES"] + 4].max(axis=0)]}))
    print(json.dumps({"config": "configs[3] summary", "terminal_row_max_h_over_alpha": float(h.max() / alpha),
                      "lq_triggers": int(eng.get("flags").cpu().numpy()[:, 2].sum()), "device_GiB": eng.device_bytes / 2 ** 30}))
    eng.close()


def config5(args):
    import torch
    import robotic_mpc_b200 as rm

    Btot, K = args.batch or 8192, args.steps or 600
    horizons = [10, 50, 100, 200, 300]
    share = [Btot // 5 + (1 if i < Btot % 5 else 0) for i in range(5)]
    tot_ms, rows = 0.0, []
    for N, B in zip(horizons, share):
        surf, z0 = workload(B, 42 + 4 + N)
        eng = rm.Engine(B, N, "SQP_RTI", dt=0.01, nlog=0)
        eng.set_params(rm.pack_params(B, 0.01, surf=surf))
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
        rows.append({"N": N, "instances": B, "ms": ms, "mpc_steps_per_s": B * K / (ms * 1e-3), "failures": int(s[:, rm.RS["NFAIL"]].sum()),
                     "rmse_e1_mean": float(rmse[:, 0].mean()), "rmse_e1_max": float(rmse[:, 0].max()),
                     "qp_iters_per_step": float(s[:, rm.RS["QP_ITERS"]].sum() / (B * K))})
        tot_ms += ms
        eng.close()
    print(json.dumps({"config": "configs[4] horizon sweep, one GPU's share", "instances": Btot, "closed_loop_steps": K,
                      "total_ms": tot_ms, "mpc_steps_per_s": Btot * K / (tot_ms * 1e-3), "per_horizon": rows}))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("which", choices=["config1", "config2", "config4", "config5"])
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--steps", type=int, default=0)
    ap.add_argument("--horizon", type=int, default=0)
    ap.add_argument("--rho", type=float, default=100.0)
    ap.add_argument("--reps", type=int, default=2)
    a = ap.parse_args()
    {"config1": config1, "config2": config2, "config4": config4, "config5": config5}[a.which](a)
