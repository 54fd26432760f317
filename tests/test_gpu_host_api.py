"""The reference-facing Python surface (Simulator / SimulationManager / MPC.solver) on the GPU, checked against the
reference's golden vectors and the oracle.  Reads like a session with the reference: build a config dict, queue runs,
run_all(), read summaries.  Tolerance 1e-6 relative (BASELINE.json north_star)."""
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = Path(__file__).resolve().parent / "golden"
RTOL = 1e-6


def relerr(a, b):
    return float(np.abs(np.asarray(a) - np.asarray(b)).max() / max(np.abs(b).max(), 1e-300))


def _base(o, **kw):
    b = dict(robot_name="ur10", dt=0.01, simulation_time=0.2, prediction_horizon=20, q_0=o.README_Q0.copy(),
             qdot_0=o.README_QD0.copy(), wcv=np.full(6, 200.0), q_min=-2 * np.pi * np.ones(6), q_max=2 * np.pi * np.ones(6),
             qdot_min=-np.array([2.16, 2.16, np.pi, 3.2, 3.2, 3.2]), qdot_max=np.array([2.16, 2.16, np.pi, 3.2, 3.2, 3.2]),
             surface_limits=((-2, 2), (-2, 2)), surface_origin=np.zeros(3), surface_orientation_rpy=np.zeros(3),
             w_qddot=0.02, scene=False)
    b.update(kw)
    return b


def test_config1_single_run_against_reference_golden(oracle):
    """BASELINE configs[0]: one closed-loop run, dt=0.01, 6 s, N=20, SQP_RTI, flat surface -- through Simulator.run()."""
    import robotic_mpc_b200 as rm

    o = oracle
    g = np.load(GOLD / "config1_rti_n20_flat.npz")
    sim = rm.Simulator(**_base(o, simulation_time=6.0, surface_coeffs=dict(a=0, b=0, c=0, d=0, e=0, f=0),
                               solver_options={"nlp_solver_type": "SQP_RTI", "qp_solver": "PARTIAL_CONDENSING_HPIPM"}))
    assert sim.Nsim == 600
    sim.run()
    assert relerr(sim.simulation_model.z.T, g["z"]) < RTOL
    assert relerr(sim.simulation_model.u[:, 1:].T, g["u"]) < RTOL
    assert np.array_equal(sim.solver_status, g["status"]) and np.array_equal(sim.sqp_iter, g["sqp_iter"])
    assert relerr(sim.cost_history, g["cost"]) < RTOL
    assert np.array_equal(sim.residuals, np.zeros((600, 4)))  # RTI logs no residuals
    # plant logs and tracking errors against the oracle's restatement of R/simulation_model.py / R/simulator.py:265-344
    m = o.Model(o.PortLeg().lib)
    P = o.make_problem(surf=np.zeros(6)).model
    err = sim.errors
    for k in (0, 1, 57, 300, 600):
        pose, rpy, tw = m.plant_logs(g["z"][k])
        e, ptz = m.tracking_errors(P, pose, tw, 0.40, 0.05)
        assert np.abs(sim.simulation_model.ee_pose(k) - pose).max() < 1e-7
        assert np.abs(sim.simulation_model.ee_orientation_euler(k) - rpy).max() < 1e-7
        assert np.abs(sim.simulation_model.ee_velocity(k) - tw).max() < 1e-6
        assert np.abs(np.array([err[f"e{j + 1}"][k] for j in range(5)]) - e).max() < 1e-7
        assert abs(err["p_task_z"][k] - ptz) < 1e-7
    # the reference's result dicts
    data, ana, summ = sim.get_data(), sim.get_analysis(), sim.get_summary()
    assert set(summ) == set(rm.simulator.SUMMARY_KEYS)
    assert data["q"].shape == (6, 601) and data["u"].shape == (6, 601) and data["ee_pose"].shape == (12, 601) and data["N"] == 20
    assert {"e1", "weighted_rmse", "rmse", "itse", "kkt_residuals", "avg_mpc_time", "p_ee_y"} <= set(ana)
    assert summ["num_failures"] == 0 and summ["total_sqp_iterations"] == 600 and summ["avg_solver_time"] > 0
    # device-accumulated sums agree with the post-hoc arithmetic on the logs
    dev = rm.summary_from_row(sim._summary_row, sim.mpc.task_weights())
    for k in ("rmse_e1", "rmse_e3", "rmse_e5", "itse_e2", "itse_e4", "weighted_rmse"):
        assert np.isclose(dev[k], summ[k], rtol=1e-9)


def test_run_all_grid_matches_oracle_per_instance(oracle):
    """A small README-style grid (horizon x weights x solver type x surfaces): grouped into one launch per static shape,
    every instance equal to the oracle's sequential run of the same configuration."""
    import robotic_mpc_b200 as rm

    o = oracle
    rng = np.random.default_rng(11)
    coeff_sets = [dict(zip("abcdef", np.array([-0.1, 0.1, -0.01, 0.01, 0.01, 0.0]) + 0.01 * rng.standard_normal(6))) for _ in range(3)]
    m = rm.SimulationManager(_base(o, simulation_time=0.15))
    m.grid_search({"prediction_horizon": [8, 12], "w_u": [0.01, 0.001],
                   "solver_options": [{"nlp_solver_type": "SQP"}, {"nlp_solver_type": "SQP_RTI"}]},
                  surface_coeff_sets=coeff_sets,
                  name_template=lambda p, k: f"s{k}_H{p['prediction_horizon']}_wu{p['w_u']}_{p['solver_options']['nlp_solver_type']}")
    res = m.run_all()
    assert len(res) == 24 and m.last_run_info["launches"] == 4  # (H=8|12) x (SQP|SQP_RTI)
    port = o.checker_leg()  # the compiled reference where it was built
    z0 = np.concatenate([o.README_Q0, o.README_QD0])
    for r, spec in zip(res, m.simulations):
        c = spec["config"]
        assert r["name"] == spec["name"] and r["simulator"].name == spec["name"]
        st = o.SQP_RTI if c["solver_options"]["nlp_solver_type"] == "SQP_RTI" else o.SQP
        surf = np.array([c["surface_coeffs"][k] for k in "abcdef"])
        gld = port.closed_loop(o.make_problem(N=c["prediction_horizon"], solver_type=st, surf=surf, w_u=c["w_u"]), z0, 15)
        assert relerr(r["data"]["u"][:, 1:].T, gld["u"]) < RTOL
        assert relerr(np.vstack([r["data"]["q"], r["data"]["qdot"]]).T, gld["z"]) < RTOL
        assert np.array_equal(r["analysis"]["sqp_iterations"], gld["sqp_iter"])
        assert np.abs(r["analysis"]["residuals"] - gld["res"]).max() < 1e-8
        assert r["summary"]["num_failures"] == int((gld["status"] != 0).sum())
    # summaries-only mode keeps nothing but the device-accumulated rows, same numbers
    res2 = m.run_all(keep_logs=False)
    for a, b in zip(res, res2):
        assert b["data"] is None and np.isclose(a["summary"]["weighted_rmse"], b["summary"]["weighted_rmse"], rtol=1e-9)
        assert a["summary"]["total_sqp_iterations"] == b["summary"]["total_sqp_iterations"]
    assert m.run_all(return_results=False) is None


def test_trajectory_optimizer_solver_interface(oracle):
    """mpc.solver.set(0,'lbx'/'ubx',x); solve(); get(0,'u'); get_stats / get_residuals / get_cost (R/simulator.py:210-221)."""
    import robotic_mpc_b200 as rm

    o = oracle
    z0 = np.concatenate([o.README_Q0, o.README_QD0])
    surf = rm.Surface(coefficients={"a": -0.12})
    model = rm.SixDofRobot(Ts=0.01, Wcv=np.full(6, 200.0))
    x = z0 + 0.01
    mpc = rm.MPC(surface=surf, initial_state=x, model=model, N_horizon=15, Tf=0.15, qmin=-2 * np.pi * np.ones(6),
                 qmax=2 * np.pi * np.ones(6), dq_min=-3 * np.ones(6), dq_max=3 * np.ones(6), w_u=0.01, w_qddot=0.02,
                 px_ref=0.4, vy_ref=0.05)
    mpc.ocp.solver_options.nlp_solver_type = "SQP"
    mpc.finalize_solver()  # initial guess x_k = initial_state, u_k = 0 (acados_solver.in.c:2744-2777)
    mpc.solver.set(0, "lbx", x)
    mpc.solver.set(0, "ubx", x)
    status = mpc.solver.solve()
    u = mpc.solver.get(0, "u")
    p = o.make_problem(N=15, solver_type=o.SQP, surf=surf.coeff_vector(), qdot_min=-3 * np.ones(6), qdot_max=3 * np.ones(6))
    gld = o.checker_leg().closed_loop(p, x, 1, full_traj=True)
    assert status == gld["status"][0] and mpc.solver.get_stats("sqp_iter") == gld["sqp_iter"][0]
    assert relerr(u, gld["u"][0]) < RTOL
    assert relerr(mpc.solver.get(7, "x"), gld["xtraj"][0][7]) < RTOL
    assert np.abs(mpc.solver.get_residuals() - gld["res"][0]).max() < 1e-8
    assert np.isclose(mpc.solver.get_cost(), gld["cost"][0], rtol=RTOL)
    assert mpc.solver.get_stats("time_tot") > 0 and mpc.solver.get_stats("qp_iter") >= 1
    # batched form: three instances, one call each
    mpcs = []
    for a in (-0.12, -0.10, -0.08):
        mm = rm.MPC(surface=rm.Surface(coefficients={"a": a}), initial_state=x, model=model, N_horizon=15, Tf=0.15,
                    qmin=-2 * np.pi * np.ones(6), qmax=2 * np.pi * np.ones(6), dq_min=-3 * np.ones(6), dq_max=3 * np.ones(6),
                    w_u=0.01, w_qddot=0.02, px_ref=0.4, vy_ref=0.05)
        mpcs.append(mm)
    bs = rm.BatchedSolver(mpcs)
    st = bs.solve()
    assert st.shape == (3,) and relerr(bs.get(0, "u")[0], gld["u"][0]) < RTOL
    assert bs.get_residuals().shape == (3, 4) and bs.get_cost().shape == (3,)
    with pytest.raises(ValueError):
        mpcs[1].ocp.solver_options.nlp_solver_type = "SQP_RTI"
        rm.BatchedSolver(mpcs)


def test_run_all_over_arms_and_options(oracle, tmp_path):
    """SimulationManager over robot_name x shift_initial_guess x terminal set: one launch per combination (they are part of
    the batching key), every instance equal to the oracle's sequential run on the same chain / option, poses reported for the
    arm's own end-effector frame; export writes one table row per simulation."""
    import robotic_mpc_b200 as rm
    import robotic_mpc_b200.robots as rb

    o = oracle
    rng = np.random.default_rng(12)
    coeff_sets = [dict(zip("abcdef", np.array([-0.1, 0.1, -0.01, 0.01, 0.01, 0.0]) + 0.01 * rng.standard_normal(6))) for _ in range(2)]
    base = _base(o, simulation_time=0.12)
    base["q_0"] = np.array([0.0, -1.8, 1.9, -1.7, -1.57, 0.0])
    base["qdot_0"] = np.array([0.2, 0.1, -0.1, 0.0, 0.0, 0.0])
    base["prediction_horizon"] = 10
    base["solver_options"] = {"nlp_solver_type": "SQP_RTI"}
    m = rm.SimulationManager(base)
    m.grid_search({"robot_name": ["ur5", "ur10"], "shift_initial_guess": [False, True],
                   "terminal_constraint": [None, {"rho": 100.0}]}, surface_coeff_sets=coeff_sets)
    res = m.run_all()
    assert len(res) == 16 and m.last_run_info["launches"] == 8
    leg = o.RefLeg() if o.REF_SO.exists() else o.PortLeg()
    z0 = np.concatenate([base["q_0"], base["qdot_0"]])
    try:
        for r, spec in zip(res, m.simulations):
            c = spec["config"]
            chain = rb.chain_table(c["robot_name"])
            leg.set_chain(*chain)
            tkw = {}
            if c["terminal_constraint"] is not None:
                P, alpha = rm.terminal_constraint.default_terminal_set(wcv=200.0, Ts=0.01, rho=100.0)
                tkw = dict(term_P=P, term_alpha=alpha)
            surf = np.array([c["surface_coeffs"][k] for k in "abcdef"])
            g = leg.closed_loop(o.make_problem(N=10, solver_type=o.SQP_RTI, surf=surf, shift=c["shift_initial_guess"], **tkw), z0, 12)
            assert relerr(r["data"]["u"][:, 1:].T, g["u"]) < RTOL
            assert relerr(np.vstack([r["data"]["q"], r["data"]["qdot"]]).T, g["z"]) < RTOL
            assert np.array_equal(r["analysis"]["solver_status"], g["status"])
            p, R = rb.forward_kinematics(chain[0], chain[1], r["data"]["q"][:, -1])
            assert np.abs(r["data"]["ee_pose"][:3, -1] - p).max() < 1e-12
    finally:
        leg.set_chain(None)
    files = m.export(tmp_path / "out", res)
    import pandas as pd

    t = pd.read_parquet(files[0])
    assert len(t) == 16 and set(t["robot_name"]) == {"ur5", "ur10"} and len(files) == 2 + 16
