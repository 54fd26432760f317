#!/usr/bin/env python
"""Cycles per code section of one stage of each interior-point sweep (lane 0 of every group, SM clock), at B = 1 and at the
bench batch.  Needs the -DRMPC_SECPROF build of the library:

    nvcc ... -DRMPC_SECPROF -o robotic-mpc_b200/librmpc_secprof.so robotic-mpc_b200/csrc/rmpc_capi.cu
    RMPC_LIB=robotic-mpc_b200/librmpc_secprof.so python profiles/tools/section_profile.py
"""
import ctypes as C
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "profiles" / "tools"))
from run_configs import workload  # noqa: E402

SECTIONS = {
    "factor": (0, ["issue copies", "wait + sync", "S1: barrier terms, P+ b, Muu row, Mxx row, stores", "sync", "S2: Mxx row -> P, 6x6 Cholesky, publish",
                   "S2: Lxu row, p, stores", "sync", "S3: Schur complement row + stores"]),
    "forward_affine": (10, ["issue copies", "wait + sync", "s = l_u + Lxu' dx, Jc dx (+ sync)", "du = -Luu^-T s, dx+", "sync, stores, sync",
                            "dpi = P+ dx+ + p+", "dlam, dt, alpha, sums", "residual row (incl. sync)"]),
    "forward_stored": (20, ["issue copies", "wait + sync", "s = z_u + Lxu' dx, Jc dx (+ sync)", "du = -Luu^-T s, dx+", "sync, stores, sync",
                            "dpi = P+ dx+ + p", "dlam, dt, alpha, sums", "residual row (incl. sync)"]),
    "backward": (30, ["issue copies", "wait + sync", "tv = p+ + P+ b", "gamma (2 divisions) (+ sync)", "r_u, r_x (+ sync)",
                      "z_u = Luu^-1 r_u, r_x -= Lxu z_u (+ sync)"]),
    "residual_first": (40, ["issue copies", "wait + sync", "values, bound slot", "control row", "state row + dynamics", "terminal row, write back, sync"]),
    "update_residual": (50, ["issue copies", "wait + sync", "values, bound slot", "control row", "state row + dynamics", "terminal row, write back, sync"]),
}


def run(B, N=100, steps=30):
    import torch
    import robotic_mpc_b200 as rm
    from robotic_mpc_b200.params import RL

    lib = rm.engine.load_library()
    lib.rmpc_debug_secprof.argtypes = [C.c_void_p, C.c_int]
    surf, z0 = workload(B, 44)
    eng = rm.Engine(B, N, "SQP_RTI", dt=0.01, nlog=steps)
    eng.set_params(rm.pack_params(B, 0.01, surf=surf))
    eng.set_state(z0)
    eng.reset()
    eng.closed_loop(10)
    torch.cuda.synchronize()
    buf = np.zeros(64)
    lib.rmpc_debug_secprof(buf.ctypes.data_as(C.c_void_p), 1)
    eng.closed_loop(steps - 10)
    torch.cuda.synchronize()
    lib.rmpc_debug_secprof(buf.ctypes.data_as(C.c_void_p), 1)
    it = eng.get("slog").cpu().numpy()[:, 10:, RL["QP_ITER"]]
    n_it, n_qp = it.sum(), it.size
    # sweeps per (instance, step): factor n_it, forward_affine n_it, forward_stored >= n_it, backward >= n_it, first 1 per QP, update n_it
    calls = {"factor": n_it, "forward_affine": n_it, "forward_stored": n_it, "backward": n_it, "residual_first": n_qp, "update_residual": n_it}
    out = {"B": B, "N": N, "qp_iter_mean": float(it.mean())}
    for name, (base, labels) in SECTIONS.items():
        per_stage = buf[base:base + len(labels)] / (calls[name] * N)
        out[name] = {"cycles_per_stage_total": float(per_stage.sum()), "sections": {lab: round(float(c), 1) for lab, c in zip(labels, per_stage)}}
    eng.close()
    return out


if __name__ == "__main__":
    for B in (1, 4096):
        print(json.dumps(run(B)), flush=True)
