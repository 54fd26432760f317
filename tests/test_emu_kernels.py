"""CPU-side check of the DEVICE SOURCE itself: rmpc_device.cuh compiled against the host warp emulator
(tests/simt_emu, randomised lane schedule) must reproduce the oracle.  This is a test harness for the kernels' logic
and synchronisation -- the product has no CPU path."""
import numpy as np
import pytest

from tests import emu


def _inputs(o, rm, B, seed, **kw):
    rng = np.random.default_rng(seed)
    surf = np.array([-0.1, 0.1, -0.01, 0.01, 0.01, 0.0]) + 0.01 * rng.standard_normal((B, 6))
    z0 = np.concatenate([o.README_Q0 + rng.uniform(-0.05, 0.05, (B, 6)), o.README_QD0 + rng.uniform(-0.1, 0.1, (B, 6))], 1)
    return surf, z0, rm.pack_params(B, 0.01, surf=surf, **kw)


@pytest.mark.parametrize("solver,N,ns,seed", [(1, 6, 25, 1), (1, 33, 8, 2), (0, 9, 8, 3), (1, 2, 10, 4)])
def test_emulated_kernels_match_oracle(oracle, solver, N, ns, seed):
    import robotic_mpc_b200.params as pm

    o = oracle
    B = 2
    surf, z0, par = _inputs(o, pm, B, seed, w_u=0.003, w_qddot=0.03)
    r = emu.closed_loop(par, z0, N, solver, ns, seed=seed)
    port = o.PortLeg()
    for i in range(B):
        p = o.make_problem(N=N, solver_type=solver, surf=surf[i], w_u=0.003, w_qddot=0.03)
        g = port.closed_loop(p, z0[i], ns)
        assert np.abs(g["u"] - r["u"][i]).max() < 1e-10 * max(1.0, np.abs(g["u"]).max())
        assert np.abs(g["z"] - r["z"][i]).max() < 1e-10
        assert np.array_equal(g["status"], r["slog"][i, :, 0].astype(int))
        assert np.array_equal(g["sqp_iter"], r["slog"][i, :, 1].astype(int))
        assert np.array_equal(g["qp_iter"], r["slog"][i, :, 2].astype(int))
        assert np.abs(g["cost"] - r["slog"][i, :, 3]).max() < 1e-10
        assert np.abs(g["res"] - r["slog"][i, :, 4:8]).max() < 1e-10
        assert r["flags"][i, 2] == 0  # LQ fall-back never needed


@pytest.mark.parametrize("case", ["term_rti_n20_inactive", "term_rti_n20_active", "term_sqp_n10_active", "term_sqp_n12_inactive"])
def test_emulated_kernels_terminal_row_against_reference_golden(built, case):
    """Terminal ellipsoid row: the device source, emulated, against the vectors of the compiled
    reference (acados BGH nh_e = 1 + HPIPM, incl. its LQ fall-back in the active cases)."""
    from pathlib import Path

    import robotic_mpc_b200.params as pm

    g = np.load(Path(__file__).resolve().parent / "golden" / f"{case}.npz")
    N, ns, st = int(g["N"]), min(int(g["nsteps"]), 30), int(g["solver_type"])
    par = pm.pack_params(1, 0.01, surf=g["surf"][None], w_u=float(g["w_u"]), w_qddot=float(g["w_qddot"]))
    term = np.concatenate([g["term_P"].ravel(), [float(g["term_lh"])], [float(g["term_alpha"])]])[None]
    r = emu.closed_loop(par, g["z0"][None], N, st, ns, seed=3, term=term)
    tol = 1e-6 if case.endswith("_active") else 1e-10
    assert np.abs(r["u"][0] - g["u"][:ns]).max() < tol * np.abs(g["u"]).max()
    assert np.abs(r["z"][0] - g["z"][:ns + 1]).max() < tol * np.abs(g["z"]).max()
    assert np.array_equal(r["slog"][0, :, 0].astype(int), g["status"][:ns])
    assert np.array_equal(r["slog"][0, :, 1].astype(int), g["sqp_iter"][:ns])
    assert np.array_equal(r["slog"][0, :, 2].astype(int), g["qp_iter"][:ns])  # incl. the iterations that take the LQ route
    assert np.abs(r["slog"][0, :, 3] - g["cost"][:ns]).max() < tol * np.abs(g["cost"]).max()


def test_emulated_schedule_independence(oracle):
    """Different lane interleavings (seeds) must give bit-identical results: no missing __syncwarp."""
    import robotic_mpc_b200.params as pm

    surf, z0, par = _inputs(oracle, pm, 1, 11)
    a = emu.closed_loop(par, z0, 7, 1, 6, seed=0)
    for s in (5, 9):
        b = emu.closed_loop(par, z0, 7, 1, 6, seed=s)
        assert np.array_equal(a["u"], b["u"]) and np.array_equal(a["slog"], b["slog"])


def test_emulated_tracking_summary(oracle):
    """Device-side e1..e5 sums (RS_SSE / RS_ITSE) against the oracle's post-hoc errors on the logged plant states."""
    import ctypes as C

    import robotic_mpc_b200.params as pm

    o = oracle
    surf, z0, par = _inputs(o, pm, 1, 21)
    ns = 12
    r = emu.closed_loop(par, z0, 5, 1, ns, seed=2)
    m = o.Model(o.PortLeg().lib)
    P = o.make_problem(surf=surf[0]).model
    sse, itse = np.zeros(5), np.zeros(5)
    for k in range(ns + 1):
        pose, _, tw = m.plant_logs(r["z"][0, k])
        e, _ = m.tracking_errors(P, pose, tw, 0.40, 0.05)
        sse += e ** 2
        itse += (k * 0.01) * e ** 2 * 0.01
    s = r["summary"][0]
    np.testing.assert_allclose(s[0:5], sse, rtol=1e-10)
    np.testing.assert_allclose(s[5:10], itse, rtol=1e-10)
    assert s[18] == ns + 1


@pytest.mark.parametrize("solver,N", [(1, 20), (0, 12), (1, 2)])
def test_emulated_shifted_closed_loop_matches_reference(built, solver, N):
    """Option "shift" (SURVEY.md 8(f) #1): the primal iterate moves one stage forward after every step.  The oracle is the
    compiled acados driven through its own get / set between solves (oracle/ref_driver.c: rmpc_ref_shift), or the port."""
    import oracle.pyoracle as o
    import robotic_mpc_b200 as rm

    rng = np.random.default_rng(31 + N)
    surf = np.array([-0.1, 0.1, -0.01, 0.01, 0.01, 0.0]) + 0.01 * rng.standard_normal(6)
    z0 = np.concatenate([o.README_Q0 + rng.uniform(-0.05, 0.05, 6), o.README_QD0 + rng.uniform(-0.1, 0.1, 6)])
    st = o.SQP_RTI if solver == 1 else o.SQP
    ns = 25
    leg = o.RefLeg() if o.REF_SO.exists() else o.PortLeg()
    g = leg.closed_loop(o.make_problem(N=N, solver_type=st, surf=surf, shift=True), z0, ns)
    g0 = leg.closed_loop(o.make_problem(N=N, solver_type=st, surf=surf), z0, ns)
    assert np.abs(g["u"] - g0["u"]).max() > 1e-9  # shifting changes the closed loop (converged SQP: only at tolerance level)
    p = o.PortLeg().closed_loop(o.make_problem(N=N, solver_type=st, surf=surf, shift=True), z0, ns)
    assert np.abs(p["u"] - g["u"]).max() < 1e-9 and np.array_equal(p["sqp_iter"], g["sqp_iter"])
    try:
        emu.set_shift(True)
        r = emu.closed_loop(rm.pack_params(1, 0.01, surf=surf[None]), z0[None], N, solver, ns, seed=5)
    finally:
        emu.set_shift(False)
    assert np.abs(r["u"][0] - g["u"]).max() < 1e-9 and np.abs(r["z"][0] - g["z"]).max() < 1e-9
    assert np.array_equal(r["slog"][0, :, 0].astype(int), g["status"])
    assert np.array_equal(r["slog"][0, :, 1].astype(int), g["sqp_iter"])


def _ref_as_rti_loop(o, prob, z0, ns):
    """The compiled acados driven the AS-RTI way: per step PREPARATION (advanced step), then FEEDBACK on the measured state."""
    m = o.Model(o.PortLeg().lib)
    s = o.RefSolver(prob, z0)
    z, U, Z, ST = z0.copy(), [], [z0.copy()], []
    for _ in range(ns):
        s.prepare()
        r = s.feedback(z)
        U.append(r["u"]); ST.append(r["status"])
        z = m.plant_rk4(prob.model.wcv, prob.dt, z, r["u"])
        Z.append(z.copy())
    s.close()
    return np.array(U), np.array(Z), np.array(ST)


@pytest.mark.parametrize("level,iters", [(0, 1), (1, 1), (1, 2), (3, 1), (3, 2)])
def test_emulated_as_rti_levels_match_reference(built, level, iters):
    """AS-RTI: levels A, B, D of ocp_nlp_sqp_rti_preparation_advanced_step
    (AC/ocp_nlp/ocp_nlp_sqp_rti.c:773-1100) against the compiled acados with as_rti_level / as_rti_iter set, closed loop of
    preparation + feedback per step.  Level C cannot be pinned: the reference aborts in it for this OCP's NONLINEAR_LS cost
    ("ocp_nlp_cost_nls_compute_gradient not implemented", AC/ocp_nlp/ocp_nlp_cost_nls.c) -- see the next test."""
    import oracle.pyoracle as o
    import robotic_mpc_b200 as rm

    if not o.REF_SO.exists():
        pytest.skip("oracle/_ref not built on this machine (needs /root/reference)")
    N, ns = 12, 20
    rng = np.random.default_rng(5 + level)
    surf = np.array([-0.1, 0.1, -0.01, 0.01, 0.01, 0.0]) + 0.01 * rng.standard_normal(6)
    z0 = np.concatenate([o.README_Q0 + rng.uniform(-0.05, 0.05, 6), o.README_QD0 + rng.uniform(-0.1, 0.1, 6)])
    U, Z, ST = _ref_as_rti_loop(o, o.make_problem(N=N, solver_type=o.SQP_RTI, surf=surf, as_rti_level=level, as_rti_iter=iters), z0, ns)
    plain = o.RefLeg().closed_loop(o.make_problem(N=N, solver_type=o.SQP_RTI, surf=surf), z0, ns)
    assert np.abs(plain["u"] - U).max() > 1e-5  # the advanced step changes the closed loop
    try:
        emu.set_as_rti(level, iters)
        r = emu.closed_loop(rm.pack_params(1, 0.01, surf=surf[None]), z0[None], N, 1, ns, seed=4)
    finally:
        emu.set_as_rti(None)
    assert np.abs(r["u"][0] - U).max() < 1e-10 * np.abs(U).max() and np.abs(r["z"][0] - Z).max() < 1e-10
    assert np.array_equal(r["slog"][0, :, 0].astype(int), ST)
    # nlp_mem->iter after the feedback: 1 on the first call (level D iterates there too), then 2 (A) / as_rti_iter + 1
    want = [1 + (iters if level == 3 else 0)] + [2 if level == 0 else iters + 1] * (ns - 1)
    assert list(r["slog"][0, :, 1].astype(int)) == want


def test_emulated_as_rti_level_c_runs(built):
    """Level C (exact gradient, gaps and bound residuals at the current iterate, matrices of the last linearisation).  No
    oracle: the reference exits for a NONLINEAR_LS cost.  Checked here: it runs, tracks, and with one iteration lands between
    level B (gradient through H * step) and level D (everything re-linearised) -- all three coincide to first order."""
    import oracle.pyoracle as o
    import robotic_mpc_b200 as rm

    N, ns = 12, 25
    rng = np.random.default_rng(9)
    surf = np.array([-0.1, 0.1, -0.01, 0.01, 0.01, 0.0]) + 0.01 * rng.standard_normal(6)
    z0 = np.concatenate([o.README_Q0 + rng.uniform(-0.05, 0.05, 6), o.README_QD0 + rng.uniform(-0.1, 0.1, 6)])
    out = {}
    for level in (1, 2, 3):
        try:
            emu.set_as_rti(level, 1)
            out[level] = emu.closed_loop(rm.pack_params(1, 0.01, surf=surf[None]), z0[None], N, 1, ns, seed=2)
        finally:
            emu.set_as_rti(None)
        assert out[level]["slog"][0, :, 0].max() == 0
    dB, dD = np.abs(out[2]["u"] - out[1]["u"]).max(), np.abs(out[2]["u"] - out[3]["u"]).max()
    assert 0 < dB < 5e-2 and 0 < dD < 5e-2
    assert np.abs(out[2]["summary"][0, :5] - out[3]["summary"][0, :5]).max() < 0.05 * np.abs(out[3]["summary"][0, :5]).max()


@pytest.mark.parametrize("N,N2,solver,seed", [(20, 2, 0, 1), (20, 10, 1, 2), (20, 7, 0, 3), (20, 3, 1, 4), (20, 5, 0, 5), (30, 4, 1, 6)])
def test_emulated_partial_condensing_matches_reference(built, N, N2, solver, seed):
    """qp_solver_cond_N = N2 < N: acados condenses the horizon into N2 blocks before HPIPM sees the QP.
    With tight joint bounds the REFERENCE ITSELF gives different interior-point iteration sequences (and solutions that differ
    at the QP tolerance) for different cond_N; the device must follow the one asked for.  Compared with the compiled acados
    at that cond_N: controls at 1e-10, status / SQP / QP iteration sequences equal."""
    import oracle.pyoracle as o
    import robotic_mpc_b200 as rm

    if not o.REF_SO.exists():
        pytest.skip("oracle/_ref not built on this machine (needs /root/reference)")
    rng = np.random.default_rng(100 + seed)
    ns = 6
    surf = np.array([-0.1, 0.1, -0.01, 0.01, 0.01, 0.0]) + 0.01 * rng.standard_normal(6)
    q0 = o.README_Q0 + rng.uniform(-0.05, 0.05, 6)
    z0 = np.concatenate([q0, o.README_QD0 + rng.uniform(-0.1, 0.1, 6)])
    w = rng.uniform(0.06, 0.14)  # joint-angle box around the initial posture: active along the horizon
    kw = dict(N=N, solver_type=solver, surf=surf, q_min=q0 - w, q_max=q0 + w)
    g = o.RefLeg().closed_loop(o.make_problem(cond_N=N2, **kw), z0, ns)
    par = rm.pack_params(1, 0.01, surf=surf[None], q_min=(q0 - w)[None], q_max=(q0 + w)[None])
    try:
        emu.set_cond_N(N2)
        r = emu.closed_loop(par, z0[None], N, solver, ns, seed=3)
    finally:
        emu.set_cond_N(None)
    assert np.abs(r["u"][0] - g["u"]).max() < 1e-10 * np.abs(g["u"]).max() and np.abs(r["z"][0] - g["z"]).max() < 1e-10
    assert np.array_equal(r["slog"][0, :, 0].astype(int), g["status"])
    assert np.array_equal(r["slog"][0, :, 1].astype(int), g["sqp_iter"])
    assert np.array_equal(r["slog"][0, :, 2].astype(int), g["qp_iter"])


def test_partial_condensing_changes_the_reference_iterates(built):
    """The discriminating power of the test above: over its cases the compiled acados with cond_N = N2 and with cond_N = N
    disagree in their QP iteration sequences at least once, and the un-condensed device run follows the cond_N = N one."""
    import oracle.pyoracle as o
    import robotic_mpc_b200 as rm

    if not o.REF_SO.exists():
        pytest.skip("oracle/_ref not built on this machine (needs /root/reference)")
    differ = 0
    for N, N2, solver, seed in [(20, 2, 0, 1), (20, 7, 0, 3), (20, 5, 0, 5)]:
        rng = np.random.default_rng(100 + seed)
        surf = np.array([-0.1, 0.1, -0.01, 0.01, 0.01, 0.0]) + 0.01 * rng.standard_normal(6)
        q0 = o.README_Q0 + rng.uniform(-0.05, 0.05, 6)
        z0 = np.concatenate([q0, o.README_QD0 + rng.uniform(-0.1, 0.1, 6)])
        w = rng.uniform(0.06, 0.14)
        kw = dict(N=N, solver_type=solver, surf=surf, q_min=q0 - w, q_max=q0 + w)
        a = o.RefLeg().closed_loop(o.make_problem(cond_N=N2, **kw), z0, 6)
        b = o.RefLeg().closed_loop(o.make_problem(**kw), z0, 6)
        differ += int(not np.array_equal(a["qp_iter"], b["qp_iter"]))
        r = emu.closed_loop(rm.pack_params(1, 0.01, surf=surf[None], q_min=(q0 - w)[None], q_max=(q0 + w)[None]), z0[None], N, solver, 6, seed=3)
        assert np.array_equal(r["slog"][0, :, 2].astype(int), b["qp_iter"])
    assert differ >= 1
