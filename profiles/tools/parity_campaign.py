#!/usr/bin/env python
"""Large-sample parity of the CUDA path against the compiled reference (oracle/_ref, the reference's own acados + HPIPM +
BLASFEO) -- a measurement kept under profiles/, not a test: every instance of a batch drawn like the bench workload
(SURVEY.md 8(d) synthetic inputs) is run closed-loop on the GPU and, instance by instance, on the host cores by the
reference; reports the worst relative deviation of the applied controls and plant states, and how often status / SQP / QP
iteration counts agree.

    python profiles/tools/parity_campaign.py [--batch 2048] [--steps 60] [--horizon 100] [--solver SQP_RTI] [--robot ur10]

Needs a GPU and oracle/_ref/librmpc_ref.so (falls back to the plain-C port and says so).
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
sys.path.insert(0, str(Path(__file__).resolve().parent))

_G = {}


def _worker(job):
    from oracle import pyoracle as o

    idx, N, st, ns, chain, term, wkw = job
    leg = o.RefLeg() if o.REF_SO.exists() else o.PortLeg()
    if chain is not None:
        leg.set_chain(*chain)
    out = []
    for i in idx:
        box = dict(q_min=_G["qbox"][0][i], q_max=_G["qbox"][1][i]) if "qbox" in _G else {}
        r = leg.closed_loop(o.make_problem(N=N, solver_type=st, surf=_G["surf"][i], **term, **wkw, **box), _G["z0"][i], ns)
        out.append((i, r["u"], r["z"], r["status"], r["sqp_iter"], r["qp_iter"]))
    return out


def main():
    import torch
    import robotic_mpc_b200 as rm
    from oracle import pyoracle as o
    from run_configs import workload

    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=2048)
    ap.add_argument("--steps", type=int, default=60)
    ap.add_argument("--horizon", type=int, default=100)
    ap.add_argument("--solver", default="SQP_RTI")
    ap.add_argument("--robot", default="ur10")
    ap.add_argument("--seed", type=int, default=44)
    ap.add_argument("--w-u", type=float, default=0.01)
    ap.add_argument("--w-qddot", type=float, default=0.02)
    ap.add_argument("--term-rho", type=float, default=0.0, help="> 0: terminal ellipsoid of the LQR with R = rho I (BASELINE configs[3])")
    ap.add_argument("--cond-N", type=int, default=0, help="qp_solver_cond_N < N: partial condensing on both sides")
    ap.add_argument("--q-box", type=float, default=0.0, help="> 0: joint-angle bounds q0 -+ U(0.06, 0.14) instead of -+2 pi (bounds active along the horizon)")
    a = ap.parse_args()
    B, N, K = a.batch, a.horizon, a.steps
    surf, z0 = workload(B, a.seed)
    chain = None
    if a.robot != "ur10":
        chain = rm.robots.chain_table(a.robot)
        z0[:, :6] += np.array([0.0, -1.8, 1.9, -1.7, -1.57, 0.0]) - np.array([np.pi / 4, -np.pi / 3, np.pi / 4, -np.pi / 2, -np.pi / 2, 0.0])
    eng = rm.Engine(B, N, a.solver, dt=0.01, nlog=K)
    if chain is not None:
        eng.set_chain(*chain)
    term = {}
    if a.term_rho > 0:
        P, alpha = rm.terminal_constraint.default_terminal_set(wcv=200.0, Ts=0.01, rho=a.term_rho)
        term = dict(term_P=P, term_alpha=alpha, term_lh=-alpha)
        eng.set_terminal(P, alpha)
    wkw = dict(w_u=a.w_u, w_qddot=a.w_qddot)
    pkw = {}
    if a.q_box > 0:
        w = np.random.default_rng(a.seed + 1).uniform(0.06, 0.14, (B, 1)) * (a.q_box / 0.1)
        pkw = dict(q_min=z0[:, :6] - w, q_max=z0[:, :6] + w)
        _G["qbox"] = (pkw["q_min"], pkw["q_max"])
    eng.set_params(rm.pack_params(B, 0.01, surf=surf, **wkw, **pkw))
    if a.cond_N:
        eng.set_option("qp_cond_N", a.cond_N)
        wkw = dict(wkw, cond_N=a.cond_N)
    eng.set_state(z0)
    eng.reset()
    torch.cuda.synchronize()
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record()
    eng.closed_loop(K)
    g1.record()
    torch.cuda.synchronize()
    gpu_s = g0.elapsed_time(g1) * 1e-3
    ulog, zlog, slog = (eng.get(k).cpu().numpy() for k in ("ulog", "zlog", "slog"))
    eng.close()
    _G["surf"], _G["z0"] = surf, z0
    cores = os.cpu_count() or 1
    st = o.SQP_RTI if a.solver == "SQP_RTI" else o.SQP
    jobs = [(list(range(c, B, cores * 4)), N, st, K, chain, term, wkw) for c in range(cores * 4)]
    from multiprocessing import get_context

    t0 = time.time()
    with get_context("fork").Pool(cores) as pool:
        res = [r for chunk in pool.map(_worker, jobs) for r in chunk]
    cpu_s = time.time() - t0
    eu = np.zeros(B)
    ez = np.zeros(B)
    same_status = same_sqp = same_qp = qp_pm1 = 0
    for i, u, z, status, sqp, qp in res:
        eu[i] = np.abs(ulog[i] - u).max() / max(np.abs(u).max(), 1e-300)
        ez[i] = np.abs(zlog[i] - z).max() / max(np.abs(z).max(), 1e-300)
        same_status += int(np.array_equal(slog[i, :, 0].astype(int), status))
        same_sqp += int(np.array_equal(slog[i, :, 1].astype(int), sqp))
        d = np.abs(slog[i, :, 2].astype(int) - qp)
        same_qp += int(d.max() == 0)
        qp_pm1 += int(d.max() <= 1)
    print(json.dumps({
        "what": "CUDA path vs compiled reference, closed loop, every instance", "oracle": "reference" if o.REF_SO.exists() else "port",
        "robot": a.robot, "terminal_rho": a.term_rho, "w_u": a.w_u, "w_qddot": a.w_qddot, "cond_N": a.cond_N, "q_box": a.q_box, "instances": B, "N": N, "solver": a.solver, "closed_loop_steps": K, "seed": a.seed,
        "u_relerr": {"max": float(eu.max()), "p99": float(np.percentile(eu, 99)), "median": float(np.median(eu))},
        "z_relerr": {"max": float(ez.max()), "p99": float(np.percentile(ez, 99)), "median": float(np.median(ez))},
        "instances_over_1e-6": int((np.maximum(eu, ez) > 1e-6).sum()),
        "status_sequences_equal": same_status, "sqp_iter_sequences_equal": same_sqp, "qp_iter_sequences_equal": same_qp,
        "qp_iter_sequences_within_1": qp_pm1, "failures_gpu": int((slog[:, :, 0] != 0).sum()),
        "cpu_cores": cores, "cpu_seconds": cpu_s, "gpu_seconds": gpu_s}))


if __name__ == "__main__":
    main()
