"""Terminal-set construction: the host-side mirror ``robotic_mpc_b200.terminal_constraint`` against
golden values computed by the reference's own functions (tests/golden/make_terminal_golden.py evaluates
R/_terminal_constraint.py:50-116,240-289 in the build container), plus the MPC wiring of the row."""
from pathlib import Path

import numpy as np
import pytest

G = np.load(Path(__file__).resolve().parent / "golden" / "sets" / "terminal_set.npz")


@pytest.fixture(scope="module")
def tc():
    import robotic_mpc_b200.terminal_constraint as m

    return m


def test_build_constraints(tc):
    qdm = np.array([2.16, 2.16, np.pi, 3.20, 3.20, 3.20])
    Cx, dx, Cu, du = tc.build_constraints(np.full(6, -2 * np.pi), np.full(6, 2 * np.pi), -qdm, qdm)
    for a, k in ((Cx, "Cx"), (dx, "dx"), (Cu, "Cu"), (du, "du")):
        assert np.array_equal(a, G[k])
    with pytest.raises(AssertionError):
        tc.build_constraints(0, 0, 0, 0, nx=10, nu=5)


@pytest.mark.parametrize("k", [0, 1, 2])
def test_terminal_ellipsoid_from_lqr_matches_reference(tc, k):
    Ad, Bd = tc.zoh_matrices(G[f"c{k}_wcv"], float(G[f"c{k}_Ts"]))
    np.testing.assert_allclose(Ad, G[f"c{k}_Ad"], rtol=0, atol=1e-15)
    np.testing.assert_allclose(Bd, G[f"c{k}_Bd"], rtol=0, atol=1e-15)
    r = tc.terminal_ellipsoid_from_lqr(Ad, Bd, np.diag(G["Q_diag"]), float(G[f"c{k}_rho"]), G["Cx"], G["dx"], G["Cu"], G["du"])
    np.testing.assert_allclose(r["P"], G[f"c{k}_P"], rtol=1e-9)
    np.testing.assert_allclose(r["K"], G[f"c{k}_K"], rtol=1e-8, atol=1e-10)
    np.testing.assert_allclose(r["alpha"], float(G[f"c{k}_alpha"]), rtol=1e-9)
    np.testing.assert_allclose(r["alpha_i"], G[f"c{k}_alpha_i"], rtol=1e-8)
    assert r["idx_active"] == int(G[f"c{k}_idx_active"])
    np.testing.assert_allclose(r["J"], float(G[f"c{k}_J"]), rtol=1e-8)
    np.testing.assert_allclose(r["slack"], G[f"c{k}_slack"], rtol=1e-8)
    # the DARE solution really is the fixed point, and the ellipsoid really is inside every half-space
    K, P = r["K"], r["P"]
    R = float(G[f"c{k}_rho"]) * np.eye(6)
    Acl = Ad + Bd @ K
    np.testing.assert_allclose(Acl.T @ P @ Acl + np.diag(G["Q_diag"]) + K.T @ R @ K, P, rtol=1e-7)
    support = np.sqrt(r["alpha"] * np.einsum("ij,ij->i", r["C"], np.linalg.solve(P, r["C"].T).T))
    assert (support <= r["d"] + 1e-9).all()


def test_tune_Q_diagonal(tc):
    Ad, Bd = tc.zoh_matrices(200.0, 0.01)
    Q, hist = tc.tune_Q_diagonal_for_state_slacks(Ad, Bd, G["Q_diag"], 100.0, G["Cx"], G["dx"], G["Cu"], G["du"], 6, n_iter=5)
    np.testing.assert_allclose(Q, G["tune_Q"], rtol=1e-8)
    np.testing.assert_allclose([h["alpha"] for h in hist], G["tune_alpha"], rtol=1e-8)
    np.testing.assert_allclose([h["J"] for h in hist], G["tune_J"], rtol=1e-7)


def test_degenerate_inputs(tc):
    assert np.isnan(tc.ellipsoid_metric_J(-np.eye(3), 1.0))
    assert np.isnan(tc.ellipsoid_metric_J(np.eye(2), 0.0))
    a, ai, idx = tc.alpha_from_constraints_over_ellipsoide(np.eye(2), np.array([[0.0, 0.0], [1.0, 0.0]]), np.array([1.0, 2.0]))
    assert np.isinf(ai[0]) and idx == 1 and a == pytest.approx(0.98 * 4.0)


def test_mpc_terminal_wiring():
    """MPC(terminal_constraint=...) carries (P, alpha, lh) and splits batches by having a terminal row or not."""
    import robotic_mpc_b200 as rm

    z0 = np.zeros(12)
    qdm = np.array([2.16, 2.16, np.pi, 3.20, 3.20, 3.20])
    kw = dict(surface=rm.Surface(), initial_state=z0, model=None, N_horizon=20, Tf=0.2, qmin=np.full(6, -2 * np.pi),
              qmax=np.full(6, 2 * np.pi), dq_min=-qdm, dq_max=qdm, w_u=0.01, w_qddot=0.02, Ts=0.01, wcv=np.full(6, 200.0))
    a = rm.MPC(**kw)
    b = rm.MPC(**kw, terminal_constraint={"rho": 100.0})
    assert a.terminal is None and a.static_key() != b.static_key()
    np.testing.assert_allclose(b.terminal["P"], G["c1_P"], rtol=1e-9)
    assert b.terminal["alpha"] == pytest.approx(float(G["c1_alpha"]), rel=1e-9) and b.terminal["lh"] == -b.terminal["alpha"]
    c = rm.MPC(**kw, terminal_constraint={"P": np.eye(12), "alpha": 3.0, "lh": -1.0})
    assert c.terminal["lh"] == -1.0 and float(c.ocp.constraints.uh_e[0]) == 3.0
