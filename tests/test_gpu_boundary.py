"""The named cost / constraint fields of the C ABI (lbu, ubu, lbx, ubx, W_diag, yref, surf_coeffs, wcv) with an
honoured `stage` argument, and the statistics / timing getters, through the reference's own solver-interface names
(AcadosOcpSolver.set / cost_set / constraints_set / get_stats, T/acados_ocp_solver.py:1461-1661,1743-2035).  The checker is
the compiled reference driven through ocp_nlp_cost_model_set / ocp_nlp_constraints_model_set (oracle/ref_driver.c:
rmpc_ref_model_set).  Tolerance 1e-6 relative (BASELINE.json north_star)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
RTOL = 1e-6


def relerr(a, b):
    return float(np.abs(np.asarray(a) - np.asarray(b)).max() / max(np.abs(b).max(), 1e-300))


def _inputs(o, B, seed):
    rng = np.random.default_rng(seed)
    surf = np.array([-0.1, 0.1, -0.01, 0.01, 0.01, 0.0]) + 0.01 * rng.standard_normal((B, 6))
    z0 = np.concatenate([o.README_Q0 + rng.uniform(-0.05, 0.05, (B, 6)), o.README_QD0 + rng.uniform(-0.1, 0.1, (B, 6))], 1)
    return surf, z0


def _need_ref(o):
    if not o.REF_SO.exists():
        pytest.skip("needs the compiled reference (oracle/_ref)")


class _Model:
    Ts = 0.01
    Wcv = np.full(6, 200.0)
    translation = (0.0, 0.0, 0.1)


def _mpcs(rm, surf, z0, N, solver):
    from robotic_mpc_b200.surface import Surface

    qd = np.array([2.16, 2.16, np.pi, 3.20, 3.20, 3.20])
    out = []
    for i in range(len(z0)):
        s = Surface(coefficients=dict(zip("abcdef", surf[i])))
        m = rm.MPC(s, z0[i], _Model(), N, N * 0.01, -2 * np.pi * np.ones(6), 2 * np.pi * np.ones(6), -qd, qd, 0.01, 0.02, px_ref=0.40,
                   vy_ref=0.05)
        m.ocp.solver_options.nlp_solver_type = solver
        out.append(m)
    return out


@pytest.mark.parametrize("solver", ["SQP_RTI", "SQP"])
def test_stage_yref_and_bounds_mid_loop_match_reference(oracle, solver):
    """An externally driven loop that, between two solves, changes `yref` at single stages (a time-varying reference), then
    tightens `ubu` at one stage and `lbx` / `ubx` on q at another -- through BatchedSolver.cost_set / constraints_set / set --
    against the compiled reference doing the same through ocp_nlp_cost_model_set / ocp_nlp_constraints_model_set."""
    import robotic_mpc_b200 as rm

    o = oracle
    _need_ref(o)
    B, N, ns = 6, 20, 8
    surf, z0 = _inputs(o, B, 31)
    bs = rm.BatchedSolver(_mpcs(rm, surf, z0, N, solver))
    st = o.SQP_RTI if solver == "SQP_RTI" else o.SQP
    refs = [o.RefSolver(o.make_problem(N=N, solver_type=st, surf=surf[i]), z0[i]) for i in range(B)]
    model = o.Model(refs[0].lib)
    z = z0.copy()
    wcv = np.full(6, 200.0)
    yref = np.zeros(18)
    yref[1], yref[3], yref[4] = 1.0, 0.40, 0.05
    worst = 0.0
    for k in range(ns):
        if k == 2:  # a different px / vy reference on the second half of the horizon, stage by stage; one u reference
            for s in range(N // 2, N):
                y = yref.copy()
                y[3], y[4] = 0.40 + 0.002 * (s - N // 2), 0.03
                if s == N - 3:
                    y[5 + 1] = 0.3
                bs.cost_set(s, "yref", y)
                for r in refs:
                    r.model_set(s, "yref", y)
        if k == 4:  # per-stage bounds: u of stage 1 capped, q box of stage 5 tightened around the current iterate
            ub = np.array([2.16, 1.2, np.pi, 3.20, 3.20, 3.20])
            bs.constraints_set(1, "ubu", ub)
            x5 = bs.get(5, "x")[:, :6]
            bs.set(5, "lbx", x5 - 0.01)
            bs.constraints_set(5, "ubx", x5 + 0.01)
            for i, r in enumerate(refs):
                r.model_set(1, "ubu", ub)
                r.model_set(5, "lbx", x5[i] - 0.01)
                r.model_set(5, "ubx", x5[i] + 0.01)
        bs.set(0, "lbx", z)
        bs.set(0, "ubx", z)
        status = bs.solve()
        u = bs.get(0, "u")
        for i, r in enumerate(refs):
            g = r.step(z[i])
            assert status[i] == g["status"], (k, i, status[i], g["status"])
            assert relerr(u[i], g["u"]) < RTOL, (k, i, relerr(u[i], g["u"]))
            assert bs.get_stats("sqp_iter")[i] == g["sqp_iter"] and bs.get_stats("qp_iter")[i] == g["qp_iter"], (k, i)
            assert abs(bs.get_cost()[i] - g["cost"]) <= RTOL * abs(g["cost"])
            worst = max(worst, relerr(u[i], g["u"]))
            z[i] = model.plant_rk4(wcv, 0.01, z[i], g["u"])
    xr, ur, _ = refs[0].iterate()
    assert relerr(bs.get(N, "x")[0], xr[N]) < RTOL and relerr(bs.get(N // 2, "u")[0], ur[N // 2]) < RTOL
    # the setters are visible through the getters; stage -1 (every stage) resets the per-stage overrides of that field
    assert np.allclose(bs.engine.get("yref", stage=N - 3).cpu().numpy()[:, 6], 0.3)
    bs.engine.set("yref", np.broadcast_to(yref, (B, 18)), stage=-1)
    assert np.array_equal(bs.engine.get("yref", stage=N - 3).cpu().numpy(), np.broadcast_to(yref, (B, 18)))
    assert np.allclose(bs.engine.get("ubu", stage=1).cpu().numpy()[:, 1], 1.2)
    for r in refs:
        r.close()
    print(f"per-stage setters, {solver}: worst relative deviation of u0 {worst:.2e}")


def test_per_instance_named_fields_equal_the_parameter_block(oracle):
    """W_diag / surf_coeffs / wcv / lbu / ubu / lbx / ubx / yref with stage = -1 write the same parameter block pack_params
    builds (bit for bit, incl. the exact-ZOH coefficients libm's exp gives), and a closed loop run after them is identical."""
    import robotic_mpc_b200 as rm

    o = oracle
    B, N, ns = 8, 12, 6
    surf, z0 = _inputs(o, B, 5)
    rng = np.random.default_rng(9)
    wcv = 200.0 + 20.0 * rng.standard_normal((B, 6))
    w_u, w_q = 0.01 * (1 + rng.uniform(0, 1, (B, 6))), 0.02 * (1 + rng.uniform(0, 1, (B, 6)))
    qd = np.array([2.0, 2.0, 3.0, 3.1, 3.1, 3.1])
    par = rm.pack_params(B, 0.01, surf=surf, wcv=wcv, w_u=w_u, w_qddot=w_q, px_ref=0.41, vy_ref=0.04, w_task=[50, 40, 30, 20, 60],
                         q_min=-3.0, q_max=3.0, qdot_min=-qd, qdot_max=qd)
    a = rm.Engine(B, N, "SQP_RTI", dt=0.01, nlog=ns)
    a.set_params(par)
    b = rm.Engine(B, N, "SQP_RTI", dt=0.01, nlog=ns)
    b.set_params(rm.pack_params(B, 0.01))
    W = np.zeros((B, 18))
    W[:, :5], W[:, 5:11], W[:, 11:17] = [50, 40, 30, 20, 60], 2 * w_u, w_q
    y = np.zeros((B, 18))
    y[:, 1], y[:, 3], y[:, 4] = 1.0, 0.41, 0.04
    b.set("W_diag", W, stage=-1)
    b.set("surf_coeffs", surf, stage=-1)
    b.set("wcv", wcv, stage=-1)
    b.set("yref", y, stage=-1)
    b.set("lbu", np.broadcast_to(-qd, (B, 6)), stage=-1)
    b.set("ubu", np.broadcast_to(qd, (B, 6)), stage=-1)
    b.set("lbx", np.full((B, 6), -3.0), stage=-1)
    b.set("ubx", np.full((B, 6), 3.0), stage=-1)
    assert np.array_equal(b.get("par").cpu().numpy(), par)
    for e in (a, b):
        e.set_state(z0)
        e.reset()
        e.closed_loop(ns)
    assert np.array_equal(a.get("ulog").cpu().numpy(), b.get("ulog").cpu().numpy())
    with pytest.raises(rm.RmpcError):
        b.set("W_diag", W, stage=3)  # per instance, not per stage
    with pytest.raises(rm.RmpcError):
        b.set("yref", y, stage=N)  # stage N has no cost
    a.close()
    b.close()


def test_statistics_table_and_timers_match_acados_layout(oracle):
    """get_stats('statistics') has acados' layout and values (row 0 iteration index; SQP: res_stat, res_eq, res_ineq,
    res_comp, qp_stat, qp_iter, alpha per iteration; SQP_RTI: qp_stat, qp_iter), and the device timers split time_tot the way
    acados splits its own: time_lin + time_qp + time_glob <= time_tot, time_qp = time_qp_solver_call + time_qp_xcond."""
    import robotic_mpc_b200 as rm

    o = oracle
    _need_ref(o)
    B, N = 4, 20
    surf, z0 = _inputs(o, B, 12)
    for solver, st in (("SQP", o.SQP), ("SQP_RTI", o.SQP_RTI)):
        bs = rm.BatchedSolver(_mpcs(rm, surf, z0, N, solver))
        bs.set(0, "lbx", z0)
        bs.solve()
        tabs = bs.get_stats("statistics")
        for i in range(B):
            r = o.RefSolver(o.make_problem(N=N, solver_type=st, surf=surf[i]), z0[i])
            r.step(z0[i])
            ref = r.statistics()
            assert tabs[i].shape == ref.shape, (solver, tabs[i].shape, ref.shape)
            assert np.array_equal(tabs[i][0], ref[0])
            if solver == "SQP":
                assert np.abs(tabs[i][1:5] - ref[1:5]).max() < 1e-8          # KKT residuals per iteration
                assert np.array_equal(tabs[i][5:7], ref[5:7])                # qp_stat, qp_iter per iteration
                assert np.abs(tabs[i][7] - ref[7]).max() < 1e-12             # accepted step lengths
            else:
                assert np.array_equal(tabs[i], ref)
            r.close()
        assert bs.get_stats("stat_n") == (7 if solver == "SQP" else 2)
        view = bs.view(1)
        assert np.array_equal(view.get_stats("statistics"), tabs[1])
        t = {k: bs.get_stats(k) for k in ("time_tot", "time_lin", "time_qp", "time_glob", "time_qp_solver_call", "time_qp_xcond")}
        ph = bs.get_stats("time_qp_phases")
        assert np.all(t["time_tot"] > 0) and np.all(t["time_lin"] > 0) and np.all(t["time_qp"] > 0) and np.all(t["time_glob"] > 0)
        assert np.all(t["time_lin"] + t["time_qp"] + t["time_glob"] <= t["time_tot"] * (1 + 1e-9) + 2e-6)
        assert np.allclose(t["time_qp_solver_call"] + t["time_qp_xcond"], t["time_qp"])
        assert np.all(ph > 0) and np.all(ph.sum(axis=1) <= t["time_qp"] * (1 + 1e-9) + 2e-6)
        assert isinstance(view.get_stats("time_lin"), float) and bs.get_stats("time_sim").shape == (B,)


@pytest.mark.parametrize("graph", [1, 0])
def test_step_host_equals_set_solve_get(oracle, graph):
    """rmpc_step_host (host x0 in, host u0 + step record out, CUDA-graph replay or plain enqueue) gives bit for bit what the
    set('z') / solve() / get('u', 0) sequence gives, in an externally driven loop whose plant runs on the host; changing a
    per-stage reference in between (new launch arguments) re-captures the graph."""
    import robotic_mpc_b200 as rm

    o = oracle
    B, N, ns = 5, 16, 6
    surf, z0 = _inputs(o, B, 77)
    par = rm.pack_params(B, 0.01, surf=surf)
    a = rm.Engine(B, N, "SQP_RTI", dt=0.01, nlog=1)
    b = rm.Engine(B, N, "SQP_RTI", dt=0.01, nlog=1)
    b.set_option("graph", graph)
    for e in (a, b):
        e.set_params(par)
        e.set_state(z0)
        e.reset()
    model = o.Model(o.PortLeg().lib)
    z = z0.copy()
    u = np.empty((B, 6))
    st = np.empty((B, rm.RL["COUNT"]))
    for k in range(ns):
        if k == 3:
            y = np.zeros((B, 18))
            y[:, 1], y[:, 3], y[:, 4] = 1.0, 0.42, 0.03
            for e in (a, b):
                e.set("yref", y, stage=N - 2)
        a.set_state(z)
        a.solve()
        ua = a.get("u", 0).cpu().numpy()
        sa = a.get("stats").cpu().numpy()
        n0 = b.launch_count
        assert b.step_host(z, u, st) is u
        assert b.launch_count - n0 == 1  # one solve kernel per step, captured or not
        assert np.array_equal(u, ua), k
        tcols = slice(rm.RL["TIME"], rm.RL["TIME_XCOND"] + 1)
        sa[:, tcols] = st[:, tcols] = 0.0  # device times differ from run to run
        assert np.array_equal(st, sa), k
        for i in range(B):
            z[i] = model.plant_rk4(np.full(6, 200.0), 0.01, z[i], u[i])
    with pytest.raises(ValueError):
        b.step_host(z[:2])
    c = rm.Engine(B, N, "SQP_RTI", dt=0.01, nlog=0)
    with pytest.raises(rm.RmpcError):
        c.step_host(z, u, st)  # the step record needs nlog > 0
    for e in (a, b, c):
        e.close()


def test_step_host_graph_follows_a_chain_change(oracle):
    """The captured step of rmpc_step_host bakes in the kinematic chain (a by-value kernel argument): after rmpc_set_chain
    the graph must be re-captured, so graph = 1 and graph = 0 keep giving the same controls."""
    import robotic_mpc_b200 as rm

    o = oracle
    B, N = 4, 12
    surf, z0 = _inputs(o, B, 91)
    par = rm.pack_params(B, 0.01, surf=surf)
    res = {}
    for graph in (0, 1):
        e = rm.Engine(B, N, "SQP_RTI", dt=0.01, nlog=1)
        e.set_option("graph", graph)
        e.set_params(par)
        e.set_state(z0)
        e.reset()
        u10 = e.step_host(z0).copy()
        e.set_chain("ur5")
        e.set_state(z0)
        e.reset()
        u5 = e.step_host(z0).copy()
        e.set_chain(None)  # back to the built-in UR10
        e.set_state(z0)
        e.reset()
        res[graph] = (u10, u5, e.step_host(z0).copy())
        e.close()
    assert not np.array_equal(res[0][0], res[0][1])  # the two arms really give different controls
    for a, b in zip(res[0], res[1]):
        assert np.array_equal(a, b, equal_nan=True)
    assert np.array_equal(res[1][0], res[1][2], equal_nan=True)


def test_run_host_discards_per_stage_overrides(oracle):
    """rmpc_run_host uploads a new parameter block: like rmpc_set "par" it drops every per-stage override made before."""
    import robotic_mpc_b200 as rm

    o = oracle
    B, N, ns = 4, 12, 5
    surf, z0 = _inputs(o, B, 92)
    par = rm.pack_params(B, 0.01, surf=surf)
    fresh = rm.Engine(B, N, "SQP_RTI", dt=0.01)
    want, u_want = fresh.run_host(par, z0, ns)
    e = rm.Engine(B, N, "SQP_RTI", dt=0.01)
    e.set_params(par)
    e.set("ubu", np.full((B, 6), 0.05), stage=1)
    e.set("yref", np.ones((B, 18)), stage=3)
    got, u_got = e.run_host(par, z0, ns)
    assert np.array_equal(e.get("ubu", stage=1).cpu().numpy(), par[:, rm.RP["UMAX"]:rm.RP["UMAX"] + 6])
    assert np.array_equal(got, want) and np.array_equal(u_got, u_want)
    for x in (fresh, e):
        x.close()


def test_step_host_refuses_as_rti(oracle):
    """One combined solve per call is what acados refuses while an AS-RTI level is set; so do rmpc_solve and rmpc_step_host."""
    import robotic_mpc_b200 as rm

    o = oracle
    B, N = 3, 10
    surf, z0 = _inputs(o, B, 93)
    e = rm.Engine(B, N, "SQP_RTI", dt=0.01)
    e.set_params(rm.pack_params(B, 0.01, surf=surf))
    e.set_state(z0)
    e.reset()
    e.set_option("as_rti_level", 1)
    with pytest.raises(rm.RmpcError) as ei:
        e.step_host(z0)
    assert ei.value.code == -1 and "AS-RTI" in str(ei.value)
    e.set_option("as_rti_level", 4)
    assert np.isfinite(e.step_host(z0)).all()
    e.close()


def test_every_stage_setter_keeps_the_other_fields_overrides(oracle):
    """stage = -1 for one field drops that field's per-stage overrides only: lbu and lbx share the lower-bound group of the
    per-stage table.  Engine.set without a stage means every stage for these fields and never touches the plant state."""
    import robotic_mpc_b200 as rm

    o = oracle
    B, N = 3, 10
    surf, z0 = _inputs(o, B, 94)
    par = rm.pack_params(B, 0.01, surf=surf)
    e = rm.Engine(B, N, "SQP_RTI", dt=0.01)
    e.set_params(par)
    e.set_state(z0)
    lbx3, ubx4, lbu2 = np.full((B, 6), -0.3), np.full((B, 6), 0.4), np.full((B, 6), -0.7)
    e.set("lbx", lbx3, stage=3)
    e.set("ubx", ubx4, stage=4)
    e.set("lbu", lbu2, stage=2)
    e.set("lbu", np.full((B, 6), -1.1))  # every stage
    e.set("ubu", np.full((B, 6), 1.3), stage=-1)
    g = lambda f, s: e.get(f, stage=s).cpu().numpy()
    assert np.array_equal(g("lbx", 3), lbx3) and np.array_equal(g("ubx", 4), ubx4)
    assert np.array_equal(g("lbu", 2), np.full((B, 6), -1.1)) and np.array_equal(g("lbu", 7), np.full((B, 6), -1.1))
    assert np.array_equal(g("ubu", 5), np.full((B, 6), 1.3))
    e.set("lbx", np.full((B, 6), -2.0))  # every stage: the override of stage 3 goes, the control bounds stay
    assert np.array_equal(g("lbx", 3), np.full((B, 6), -2.0)) and np.array_equal(g("lbu", 3), np.full((B, 6), -1.1))
    assert np.array_equal(e.get("z").cpu().numpy(), z0)
    assert np.array_equal(g("yref", 3), par[:, rm.RP["YREF"]:rm.RP["YREF"] + 18])
    e.close()
