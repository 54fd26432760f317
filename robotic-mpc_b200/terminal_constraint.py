"""Terminal ellipsoid of the LQR value function (host side, offline) -- mirror of R/_terminal_constraint.py:50-116,240-289.

The set is  E = { x : x' P x <= alpha }  with P the DARE solution for (Ad, Bd, Q, R = rho I) and alpha the largest
level (times a safety factor) for which E stays inside the state box and the LQR input u = K x stays inside the input
box:  alpha = safety * min_i d_i^2 / (c_i' P^-1 c_i)  over the rows c_i' x <= d_i of [Cx; Cu K].

The reference computes (P, alpha) with scipy and never wires them into its OCP; here they feed the stage-N row
``lh_e <= x_N' P x_N <= alpha`` of the batched solver (``MPC(terminal_constraint=...)``, C-ABI fields ``P_e``,
``alpha_e``).  Same function names, argument order and return keys as the reference module; the plotting / YAML
export helpers of that file are presentation and are not mirrored.
"""
from __future__ import annotations

import numpy as np
from scipy.linalg import solve_discrete_are


def discrete_lqr(A, B, Q, R):
    """(K, P) of the discrete-time LQR, u = K x  (R/_terminal_constraint.py:50-54)."""
    A, B, Q, R = (np.asarray(M, dtype=float) for M in (A, B, Q, R))
    P = solve_discrete_are(A, B, Q, R)
    BtP = B.T @ P
    K = -np.linalg.solve(R + BtP @ B, BtP @ A)
    return K, P


def build_constraints(q_min, q_max, qdot_min, qdot_max, nx=12, nu=6):
    """Half-space form of the joint-position box on x = [q; qd] and of the input box (R/_terminal_constraint.py:57-69)."""
    if nx != 12 or nu != 6:
        raise AssertionError("Expected dimensions: nx=12, nu=6")
    half = nx // 2
    sel = np.hstack([np.eye(half), np.zeros((half, half))])
    Cx = np.vstack([sel, -sel])
    dx = np.concatenate([np.asarray(q_max, float), -np.asarray(q_min, float)])
    Cu = np.vstack([np.eye(nu), -np.eye(nu)])
    du = np.concatenate([np.asarray(qdot_max, float), -np.asarray(qdot_min, float)])
    return Cx, dx, Cu, du


def alpha_from_constraints_over_ellipsoide(P, C, d, safety=0.98):
    """Largest level of x'Px inside every half-space c_i'x <= d_i, times `safety` (R/_terminal_constraint.py:72-86).
    Returns (alpha, alpha_i, idx_active)."""
    P = np.asarray(P, float)
    C = np.atleast_2d(np.asarray(C, float))
    d = np.asarray(d, float)
    support = np.einsum("ij,ij->i", C, np.linalg.solve(P, C.T).T)  # c_i' P^-1 c_i
    alpha_i = np.empty(d.shape[0])
    for i, (s, di) in enumerate(zip(support, d)):
        if s <= 1e-15:
            alpha_i[i] = np.inf if di >= 0 else 0.0
        else:
            alpha_i[i] = di * di / s
    idx_active = int(np.argmin(alpha_i))
    return safety * float(alpha_i[idx_active]), alpha_i, idx_active


def ellipsoid_metric_J(P, alpha):
    """Volume proxy alpha^(n/2) / sqrt(det P); NaN for a non-positive determinant or level (R/_terminal_constraint.py:89-94)."""
    P = np.asarray(P, float)
    det = np.linalg.det(P)
    if det <= 0 or alpha <= 0:
        return np.nan
    return alpha ** (P.shape[0] / 2.0) / np.sqrt(det)


def terminal_ellipsoid_from_lqr(A_d, B_d, Q, rho, Cx, dx, Cu, du, safety=0.98):
    """P, K of the LQR with R = rho I and the level alpha of the largest admissible ellipsoid
    (R/_terminal_constraint.py:96-116); keys: K P alpha alpha_i idx_active J slack C d."""
    B_d = np.asarray(B_d, float)
    K, P = discrete_lqr(A_d, B_d, Q, float(rho) * np.eye(B_d.shape[1]))
    C = np.vstack([Cx, Cu @ K])
    d = np.concatenate([dx, du])
    alpha, alpha_i, idx_active = alpha_from_constraints_over_ellipsoide(P, C, d, safety=safety)
    return {"K": K, "P": P, "alpha": alpha, "alpha_i": alpha_i, "idx_active": idx_active,
            "J": ellipsoid_metric_J(P, alpha), "slack": alpha_i / max(alpha, 1e-18), "C": C, "d": d}


def tune_Q_diagonal_for_state_slacks(A_d, B_d, Q_diag_init, rho_star, Cx, dx, Cu, du, nu, safety=0.98, n_iter=12,
                                     beta=0.35, w_min=1e-3, w_max=1e6):
    """Re-weight the joint-position entries of diag(Q) so that the position half-spaces become equally tight: each pass
    scales w_j by max(slack_j, 1)^-beta, clips, and restores the geometric mean (R/_terminal_constraint.py:240-289).
    Returns (Q_diag, history)."""
    Q_diag = np.array(Q_diag_init, dtype=float)
    hist = []
    for it in range(n_iter):
        res = terminal_ellipsoid_from_lqr(A_d, B_d, np.diag(Q_diag), rho_star, Cx, dx, Cu, du, safety=safety)
        slack = res["slack"]
        slack_state = slack[:2 * nu]
        s_joint = np.maximum(slack_state[:nu], slack_state[nu:2 * nu])
        w_old = Q_diag[:nu]
        w_new = np.clip(w_old * np.maximum(s_joint, 1.0) ** (-beta), w_min, w_max)
        g_old = np.exp(np.mean(np.log(np.maximum(w_old, 1e-18))))
        g_new = np.exp(np.mean(np.log(np.maximum(w_new, 1e-18))))
        if g_new > 0:
            w_new = w_new * (g_old / g_new)
        Q_diag[:nu] = w_new
        hist.append({"iter": it, "Q_diag": Q_diag.copy(), "alpha": res["alpha"], "J": res["J"],
                     "idx_active": res["idx_active"], "slack": slack.copy(), "slack_state": slack_state.copy(),
                     "s_joint": s_joint.copy()})
    return Q_diag, hist


def zoh_matrices(wcv, Ts):
    """Ad, Bd of the exact-ZOH joint model x = [q; qd], qd' = -w qd + w u (R/prediction_model.py:87-115)."""
    w = np.broadcast_to(np.asarray(wcv, float), (6,))
    a22 = np.exp(-w * Ts)
    a12 = (1.0 - a22) / w
    Ad = np.block([[np.eye(6), np.diag(a12)], [np.zeros((6, 6)), np.diag(a22)]])
    Bd = np.vstack([np.diag(Ts - a12), np.diag(1.0 - a22)])
    return Ad, Bd


def default_terminal_set(wcv=200.0, Ts=0.01, rho=1.0, q_min=None, q_max=None, qdot_min=None, qdot_max=None,
                         Q_diag=None, safety=0.98):
    """(P, alpha) for BASELINE config 4: Q = diag(10,10,100,10,10,10, 1 x 6) (R/_terminal_constraint.py:313-318), the
    UR10 joint and velocity boxes of SURVEY.md 8(d) unless given."""
    qdm = np.array([2.16, 2.16, np.pi, 3.20, 3.20, 3.20])
    q_min = np.full(6, -2 * np.pi) if q_min is None else np.asarray(q_min, float)
    q_max = np.full(6, 2 * np.pi) if q_max is None else np.asarray(q_max, float)
    qdot_min = -qdm if qdot_min is None else np.asarray(qdot_min, float)
    qdot_max = qdm if qdot_max is None else np.asarray(qdot_max, float)
    if Q_diag is None:
        Q_diag = np.array([10.0, 10.0, 100.0, 10.0, 10.0, 10.0] + [1.0] * 6)
    Ad, Bd = zoh_matrices(wcv, Ts)
    Cx, dx, Cu, du = build_constraints(q_min, q_max, qdot_min, qdot_max)
    res = terminal_ellipsoid_from_lqr(Ad, Bd, np.diag(Q_diag), rho, Cx, dx, Cu, du, safety=safety)
    return res["P"], res["alpha"]
