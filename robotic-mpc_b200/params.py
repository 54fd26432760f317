"""Packing of per-instance problem data into the engine's parameter block (host side, numpy).

Layout = the ``RP_*`` enum of ``csrc/rmpc_types.h``.  Values and defaults follow the reference:
``R/prediction_model.py:87-115`` (exact-ZOH coefficients), ``R/trajectory_optimizer.py:44-48,142-176`` (weights,
bounds), ``R/simulator.py:18-35`` (Simulator defaults), ``R/surface.py:14-17`` (surface defaults).
"""
from __future__ import annotations

import math

import numpy as np

# mirrors of the enums of include/rmpc_fields.h (checked against the header by tests/test_capi_symbols.py)
RP = dict(SURF=0, A12=6, A22=12, B1=18, B2=24, WCV=30, WU=36, WQDD=42, YREF=48, PXREF=51, VYREF=52, WTASK=66, UMIN=72,
          UMAX=78, QMIN=84, QMAX=90, TEE=96, COUNT=100)
RT = dict(YREF=0, LB=18, UB=30, COUNT=42)
RL = dict(STATUS=0, SQP_ITER=1, QP_ITER=2, COST=3, RES=4, ALPHA=8, QP_STATUS=9, TIME=10, TIME_LIN=11, TIME_QP=12,
          TIME_GLOB=13, TIME_QP_FACT=14, TIME_QP_FWD=15, TIME_QP_BWD=16, TIME_QP_RES=17, TIME_XCOND=18, COUNT=20)
RK = dict(POSE=0, RPY=12, TWIST=15, ERR=21, PTZ=26, COUNT=28)
RS = dict(SSE=0, ITSE=5, NFAIL=10, SQP_ITERS=11, QP_ITERS=12, MAX_RES=13, COST_LAST=17, NSAMPLES=18, COUNT=20)
RE = dict(P=0, LH=144, UH=145, COUNT=146)
RQ = dict(RES=0, QP_STATUS=4, QP_ITER=5, ALPHA=6, COUNT=7)

DEFAULT_SURFACE = (-0.15, 0.15, -0.01, 0.01, 0.01, 0.0)
DEFAULT_QDOT_MAX = (2.16, 2.16, math.pi, 3.20, 3.20, 3.20)


def _bc(v, B, n):
    a = np.asarray(v, dtype=np.float64)
    if a.ndim == 0:
        a = np.full((n,), float(a))
    return np.broadcast_to(a, (B, n)) if a.ndim == 1 else a.reshape(B, n)


def pack_params(B, dt, surf=DEFAULT_SURFACE, wcv=200.0, w_u=0.01, w_qddot=0.02, px_ref=0.40, vy_ref=0.05,
                w_task=50.0, q_min=-2 * math.pi, q_max=2 * math.pi, qdot_min=None, qdot_max=None,
                t_ee=(0.0, 0.0, 0.1), yref=None) -> np.ndarray:
    """Return the ``[B, RP_COUNT]`` float64 parameter block; every argument broadcasts over the batch.  ``w_u`` / ``w_qddot``
    may be scalars (the reference), ``[B]`` or per joint ``[B, 6]``; ``yref`` (``[18]`` or ``[B, 18]``) overrides the
    reference's ``[0, 1, 0, px_ref, vy_ref, 0 ...]``."""
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
    def per_joint(v):
        a = np.asarray(v, dtype=np.float64)
        if a.ndim == 0:
            return np.full((B, 6), float(a))
        if a.ndim == 1 and a.shape[0] == B and B != 6:
            return np.repeat(a[:, None], 6, axis=1)
        return np.broadcast_to(a if a.ndim == 2 else (a[:, None] if a.shape[0] == B else a[None, :]), (B, 6))

    P[:, RP["WU"]:RP["WU"] + 6] = per_joint(w_u)
    P[:, RP["WQDD"]:RP["WQDD"] + 6] = per_joint(w_qddot)
    if yref is None:
        P[:, RP["YREF"] + 1] = 1.0  # g2_ref (R/trajectory_optimizer.py:108-114); every other reference is 0
        P[:, RP["PXREF"]] = np.broadcast_to(np.asarray(px_ref, float), (B,))
        P[:, RP["VYREF"]] = np.broadcast_to(np.asarray(vy_ref, float), (B,))
    else:
        P[:, RP["YREF"]:RP["YREF"] + 18] = _bc(yref, B, 18)
    P[:, RP["WTASK"]:RP["WTASK"] + 5] = _bc(w_task, B, 5)
    qdmax = np.asarray(DEFAULT_QDOT_MAX if qdot_max is None else qdot_max, float)
    qdmin = -np.asarray(DEFAULT_QDOT_MAX) if qdot_min is None else np.asarray(qdot_min, float)
    P[:, RP["UMIN"]:RP["UMIN"] + 6] = _bc(qdmin, B, 6)
    P[:, RP["UMAX"]:RP["UMAX"] + 6] = _bc(qdmax, B, 6)
    P[:, RP["QMIN"]:RP["QMIN"] + 6] = _bc(q_min, B, 6)
    P[:, RP["QMAX"]:RP["QMAX"] + 6] = _bc(q_max, B, 6)
    P[:, RP["TEE"]:RP["TEE"] + 3] = _bc(t_ee, B, 3)
    return np.ascontiguousarray(P)
