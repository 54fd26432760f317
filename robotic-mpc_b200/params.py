"""Packing of per-instance problem data into the engine's parameter block (host side, numpy).

Layout = the ``RP_*`` enum of ``csrc/rmpc_types.h``.  Values and defaults follow the reference:
``R/prediction_model.py:87-115`` (exact-ZOH coefficients), ``R/trajectory_optimizer.py:44-48,142-176`` (weights,
bounds), ``R/simulator.py:18-35`` (Simulator defaults), ``R/surface.py:14-17`` (surface defaults).
"""
from __future__ import annotations

import math

import numpy as np

RP = dict(SURF=0, A12=6, A22=12, B1=18, B2=24, WCV=30, WU=36, WQDD=37, PXREF=38, VYREF=39, WTASK=40, UMIN=45,
          UMAX=51, QMIN=57, QMAX=63, TEE=69, COUNT=72)
RL = dict(STATUS=0, SQP_ITER=1, QP_ITER=2, COST=3, RES=4, ALPHA=8, QP_STATUS=9, TIME=10, COUNT=12)
RK = dict(POSE=0, RPY=12, TWIST=15, ERR=21, PTZ=26, COUNT=28)
RS = dict(SSE=0, ITSE=5, NFAIL=10, SQP_ITERS=11, QP_ITERS=12, MAX_RES=13, COST_LAST=17, NSAMPLES=18, COUNT=20)

DEFAULT_SURFACE = (-0.15, 0.15, -0.01, 0.01, 0.01, 0.0)
DEFAULT_QDOT_MAX = (2.16, 2.16, math.pi, 3.20, 3.20, 3.20)


def _bc(v, B, n):
    a = np.asarray(v, dtype=np.float64)
    if a.ndim == 0:
        a = np.full((n,), float(a))
    return np.broadcast_to(a, (B, n)) if a.ndim == 1 else a.reshape(B, n)


def pack_params(B, dt, surf=DEFAULT_SURFACE, wcv=200.0, w_u=0.01, w_qddot=0.02, px_ref=0.40, vy_ref=0.05,
                w_task=50.0, q_min=-2 * math.pi, q_max=2 * math.pi, qdot_min=None, qdot_max=None,
                t_ee=(0.0, 0.0, 0.1)) -> np.ndarray:
    """Return the ``[B, 72]`` float64 parameter block; every argument broadcasts over the batch."""
    P = np.zeros((B, RP["COUNT"]))
    P[:, RP["SURF"]:RP["SURF"] + 6] = _bc(surf, B, 6)
    w = _bc(wcv, B, 6)
    # math.exp (libm) element by element, the way the reference evaluates its numpy scalars
    a22 = np.vectorize(lambda x: math.exp(-x * dt))(w)
    a12 = (1.0 - a22) / w
    P[:, RP["A12"]:RP["A12"] + 6] = a12
    P[:, RP["A22"]:RP["A22"] + 6] = a22
    P[:, RP["B1"]:RP["B1"] + 6] = dt - a12
    P[:, RP["B2"]:RP["B2"] + 6] = 1.0 - a22
    P[:, RP["WCV"]:RP["WCV"] + 6] = w
    P[:, RP["WU"]] = np.broadcast_to(np.asarray(w_u, float), (B,))
    P[:, RP["WQDD"]] = np.broadcast_to(np.asarray(w_qddot, float), (B,))
    P[:, RP["PXREF"]] = np.broadcast_to(np.asarray(px_ref, float), (B,))
    P[:, RP["VYREF"]] = np.broadcast_to(np.asarray(vy_ref, float), (B,))
    P[:, RP["WTASK"]:RP["WTASK"] + 5] = _bc(w_task, B, 5)
    qdmax = np.asarray(DEFAULT_QDOT_MAX if qdot_max is None else qdot_max, float)
    qdmin = -np.asarray(DEFAULT_QDOT_MAX) if qdot_min is None else np.asarray(qdot_min, float)
    P[:, RP["UMIN"]:RP["UMIN"] + 6] = _bc(qdmin, B, 6)
    P[:, RP["UMAX"]:RP["UMAX"] + 6] = _bc(qdmax, B, 6)
    P[:, RP["QMIN"]:RP["QMIN"] + 6] = _bc(q_min, B, 6)
    P[:, RP["QMAX"]:RP["QMAX"] + 6] = _bc(q_max, B, 6)
    P[:, RP["TEE"]:RP["TEE"] + 3] = _bc(t_ee, B, 3)
    return np.ascontiguousarray(P)
