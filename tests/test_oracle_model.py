"""Pins for the hand-restated model functions (oracle/ur10_model.c).  The reference obtains them from pinocchio and
CasADi, neither vendored nor installable here (PARITY UNPINNED at this boundary); they are pinned by
known answers of the UR10 chain, finite differences and the exact-ZOH matrix exponential."""
import ctypes as C

import numpy as np
import pytest
from scipy.linalg import expm


@pytest.fixture(scope="module")
def model(oracle):
    return oracle.Model(oracle.PortLeg().lib)


def test_ur10_known_answers(oracle, model):
    # values from the URDF chain, consistent with the UR10 DH home pose
    pj, aj, R, p = model.fk(np.zeros(6))
    np.testing.assert_allclose(p, [1.184300000415, 0.256141, 0.011600002126], atol=2e-9)
    np.testing.assert_allclose(R, [[-1, 0, 0], [0, 0, 1], [0, 1, 0]], atol=1e-8)
    assert abs(np.linalg.det(model.jacobian_world(np.zeros(6)))) < 1e-12
    q = oracle.README_Q0
    pj, aj, R, p = model.fk(q)
    np.testing.assert_allclose(p, [0.587237391534, 0.819084977163, 0.746316690991], atol=2e-9)
    np.testing.assert_allclose(R, [[0.707106781187, -0.683012701235, 0.183012704344],
                                   [-0.707106781187, -0.683012701235, 0.183012704344],
                                   [0, -0.25881904857, -0.96592582536]], atol=2e-9)
    J = model.jacobian_world(q)
    assert abs(np.linalg.det(J) - 0.240370519316) < 1e-9
    tw = J @ oracle.README_QD0
    np.testing.assert_allclose(tw, [-0.644816010769, -0.644816010769, 0.305999999049, -2.12132034356, 2.12132034356, 1],
                               atol=2e-9)


def test_world_jacobian_convention(oracle, model):
    # WORLD spatial Jacobian: v_P = v_O + w x p_ee must equal d p_ee / dt (finite differences)
    rng = np.random.default_rng(0)
    for _ in range(5):
        q, qd = rng.uniform(-2, 2, 6), rng.uniform(-1, 1, 6)
        tw = model.jacobian_world(q) @ qd
        p = model.fk(q)[3]
        h = 1e-6
        pd = (model.fk(q + h * qd)[3] - model.fk(q - h * qd)[3]) / (2 * h)
        np.testing.assert_allclose(tw[:3] + np.cross(tw[3:], p), pd, atol=1e-8)


def test_cost_jacobian_fd(oracle, model):
    rng = np.random.default_rng(1)
    P = oracle.make_problem().model
    for _ in range(5):
        x = np.concatenate([rng.uniform(-2, 2, 6), rng.uniform(-1, 1, 6)])
        u = rng.uniform(-1, 1, 6)
        y, Jt = model.cost_y_jac(P, x, u)  # Jt[r, c] = d y_c / d v_r, v = [u; x]
        v = np.concatenate([u, x])
        h = 1e-6
        for r in range(18):
            e = np.zeros(18); e[r] = h
            yp = model.cost_y(P, (v + e)[6:], (v + e)[:6])
            ym = model.cost_y(P, (v - e)[6:], (v - e)[:6])
            fd = (yp - ym) / (2 * h)
            np.testing.assert_allclose(Jt[r, :17], fd[:17], atol=2e-7)
        assert np.all(Jt[:, 17] == 0.0)  # weight-0 output: column returned as zero (see ur10_model.h)
        np.testing.assert_allclose(y, model.cost_y(P, x, u), atol=0)


def test_exact_zoh(oracle, model):
    P = oracle.make_problem(dt=0.01, wcv=[200, 150, 100, 250, 50, 20]).model
    BAbt = model.BAbt(P)  # (nu+nx) x nx
    w = np.array(P.wcv)
    M = np.zeros((18, 18))
    M[0:6, 6:12] = np.eye(6)
    M[6:12, 6:12] = -np.diag(w)
    M[6:12, 12:18] = np.diag(w)
    E = expm(M * 0.01)
    Ad, Bd = E[:12, :12], E[:12, 12:]
    np.testing.assert_allclose(BAbt[6:, :].T, Ad, atol=1e-13)
    np.testing.assert_allclose(BAbt[:6, :].T, Bd, atol=1e-13)
    x, u = np.arange(12) * 0.1, np.arange(6) * 0.3
    np.testing.assert_allclose(model.dyn(P, x, u), Ad @ x + Bd @ u, atol=1e-13)


def test_plant_rk4_vs_exact(oracle, model):
    wcv = np.full(6, 200.0)
    z, u = np.concatenate([oracle.README_Q0, oracle.README_QD0]), np.array([0.5, -0.2, 0.1, 0.3, 0.0, -0.4])
    zn = model.plant_rk4(wcv, 0.01, z, u)
    # RK4 on qd' = -w qd + w u is the stability polynomial R(-w dt) = 1 - 2 + 2 - 4/3 + 2/3 = 1/3 at w dt = 2
    # (the exact factor is exp(-2) = 0.135: the reference's plant and prediction model differ, on purpose or not)
    np.testing.assert_allclose(zn[6:], z[6:] / 3 + (2 / 3) * u, atol=1e-14)
    # one explicit hand evaluation of the RK4 tableau for joint 0
    f = lambda q, qd: (qd, -200.0 * qd + 200.0 * u[0])
    q, qd, h = z[0], z[6], 0.01
    k1 = f(q, qd); k2 = f(q + h / 2 * k1[0], qd + h / 2 * k1[1]); k3 = f(q + h / 2 * k2[0], qd + h / 2 * k2[1])
    k4 = f(q + h * k3[0], qd + h * k3[1])
    assert abs(zn[0] - (q + h / 6 * k1[0] + h / 3 * k2[0] + h / 3 * k3[0] + h / 6 * k4[0])) < 1e-15
    assert abs(zn[6] - (qd + h / 6 * k1[1] + h / 3 * k2[1] + h / 3 * k3[1] + h / 6 * k4[1])) < 1e-13
