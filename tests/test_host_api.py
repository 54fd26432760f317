"""Host-side mirror of the reference's Python surface (R/simulator.py, R/trajectory_optimizer.py): queueing, naming,
grouping into batches, option handling, summary arithmetic and the multi-rank sharding / gather (gloo, world_size 2).
No GPU needed: nothing here computes an MPC step (and the product refuses to, without a GPU)."""
import os
import socket

import numpy as np
import pytest

import robotic_mpc_b200 as rm
from robotic_mpc_b200 import distributed as rdist

BASE = dict(robot_name="ur10", dt=0.01, simulation_time=0.1, prediction_horizon=20,
            q_0=np.array([np.pi / 4, -np.pi / 3, np.pi / 4, -np.pi / 2, -np.pi / 2, 0.0]), qdot_0=np.array([1.0, 2, 1, 0, 0, 0]),
            wcv=np.full(6, 200.0), q_min=-2 * np.pi * np.ones(6), q_max=2 * np.pi * np.ones(6),
            qdot_min=-np.array([2.16, 2.16, np.pi, 3.2, 3.2, 3.2]), qdot_max=np.array([2.16, 2.16, np.pi, 3.2, 3.2, 3.2]),
            surface_limits=((-2, 2), (-2, 2)), surface_origin=np.zeros(3), surface_orientation_rpy=np.zeros(3),
            w_qddot=0.02, scene=False)


def test_grid_search_names_and_configs_follow_the_reference():
    m = rm.SimulationManager(BASE)
    m.grid_search({"prediction_horizon": [20, 30], "w_u": [0.01, 0.001]})
    assert [s["name"] for s in m.simulations] == ["prediction_horizon=20_w_u=0.01", "prediction_horizon=20_w_u=0.001",
                                                  "prediction_horizon=30_w_u=0.01", "prediction_horizon=30_w_u=0.001"]
    assert m.simulations[3]["config"]["prediction_horizon"] == 30 and m.simulations[3]["config"]["w_u"] == 0.001
    m.clear()
    m.grid_search({"surface_coeffs.a": [0.1]}, surface_coeff_sets=[{"a": -0.2, "b": 0.2}, {"a": -0.1, "b": 0.1}])
    assert [s["name"] for s in m.simulations] == ["coeffs0_a=0.1", "coeffs1_a=0.1"]
    assert m.simulations[1]["config"]["surface_coeffs"] == {"a": -0.1, "b": 0.1}  # the coefficient set wins (R/simulator.py:621-622)
    m.clear()
    m.grid_search({"w_u": [1, 2]}, name_template=lambda p, k: f"{k}:{p['w_u']}")
    m.grid_search({"w_u": [3]}, name_template=lambda p: f"one-arg {p['w_u']}")  # TypeError fallback (R/simulator.py:627-631)
    assert [s["name"] for s in m.simulations] == ["0:1", "0:2", "one-arg 3"]
    assert "w_u" not in BASE  # base config untouched


def test_sweep_and_readme_aliases():
    m = rm.SimulationManager(BASE)
    m.sweep("surface_coeffs.a", [-0.2, 0.0])
    m.sweep("prediction_horizon", [50], name_template="H={}")
    m.sweep_parameter("prediction_horizon", [10, 50], name_template="H={}")
    m.add_manual(name="H=200", params={"prediction_horizon": 200})
    names = [s["name"] for s in m.simulations]
    assert names == ["surface_coeffs.a=-0.2", "surface_coeffs.a=0.0", "H=50", "H=10", "H=50", "H=200"]
    assert m.simulations[0]["config"]["surface_coeffs"] == {"a": -0.2}
    assert m.simulations[-1]["config"]["prediction_horizon"] == 200
    m.clear()
    assert m.simulations == []


def test_simulator_defaults_attributes_and_guards():
    s = rm.Simulator(**BASE)
    assert s.Nsim == 10 and s.name == "ur10" and s.px_ref == 0.40 and s.vy_ref == 0.05 and s.w_u == 0.01
    assert s.mpc.ocp.solver_options.nlp_solver_type == "SQP"  # the reference's default (R/trajectory_optimizer.py:60)
    assert s.mpc.w_origin_task == 50.0 and s.mpc.N_horizon == 20 and abs(s.mpc.Tf - 0.2) < 1e-15
    assert s.mpc_time.shape == (10,) and s.residuals.shape == (10, 4) and s.sqp_iter.dtype.kind == "i"
    assert s.simulation_model.z.shape == (12, 11) and np.array_equal(s.simulation_model.u[:, 0], BASE["qdot_0"])
    assert s.scene is None
    for attr in ("errors", "metrics", "solver_stats", "timings"):
        with pytest.raises(RuntimeError, match="Must call run"):
            getattr(s, attr)
    with pytest.raises(RuntimeError):
        s.get_summary()
    row = s.mpc.param_row()
    assert row[rm.RP["SURF"]] == -0.15 and row[rm.RP["VYREF"]] == 0.05 and row[rm.RP["WQDD"]] == 0.02
    assert np.allclose(row[rm.RP["A22"]:rm.RP["A22"] + 6], np.exp(-2.0))


def test_solver_options_handling(capsys):
    s = rm.Simulator(**{**BASE, "solver_options": [{"nlp_solver_type": "SQP_RTI", "qp_solver": "PARTIAL_CONDENSING_HPIPM",
                                                     "bogus_option": 3}]})
    assert s.mpc.ocp.solver_options.nlp_solver_type == "SQP_RTI"
    assert "Unknown solver option 'bogus_option'" in capsys.readouterr().out  # R/simulator.py:129-135
    with pytest.raises(NotImplementedError):
        rm.Simulator(**{**BASE, "solver_options": {"qp_solver": "FULL_CONDENSING_QPOASES"}})
    with pytest.raises(ValueError):
        rm.Simulator(**{**BASE, "robot_name": "ur3"})  # no built-in chain table
    s5 = rm.Simulator(**{**BASE, "robot_name": "ur5"})
    assert s5.static_key() != s.static_key()  # another chain = another batched launch
    # round 2: partial condensing and AS-RTI are solver options like in acados (T/acados_ocp_options.py:72-84,1129-1208)
    sc = rm.Simulator(**{**BASE, "solver_options": {"nlp_solver_type": "SQP_RTI", "qp_solver_cond_N": 5}})
    assert sc.mpc.ocp.solver_options.qp_solver_cond_N == 5 and sc.static_key() != s.static_key()
    sc.mpc.ocp.solver_options.validate()
    sa = rm.Simulator(**{**BASE, "solver_options": {"nlp_solver_type": "SQP_RTI", "as_rti_level": 1, "as_rti_iter": 2}})
    sa.mpc.ocp.solver_options.validate()
    assert sa.static_key() != s.static_key()
    for bad in ({"qp_solver_cond_N": 0}, {"qp_solver_cond_N": 10 ** 6}, {"nlp_solver_type": "SQP_RTI", "as_rti_level": 7},
                {"nlp_solver_type": "SQP", "as_rti_level": 1}, {"nlp_solver_type": "SQP_RTI", "as_rti_iter": -1}):
        with pytest.raises(ValueError):
            rm.Simulator(**{**BASE, "solver_options": bad})  # validated at construction (R/simulator.py:129-135 sets them)
    ssh = rm.Simulator(**{**BASE, "shift_initial_guess": True})
    assert ssh.mpc.shift_initial_guess and ssh.static_key() != rm.Simulator(**BASE).static_key()


def test_plan_groups_by_static_shape_and_rank():
    m = rm.SimulationManager(BASE)
    m.grid_search({"prediction_horizon": [20, 30, 40], "w_qddot": [0.02, 0.05], "w_u": [0.01, 0.001],
                   "solver_options": [{"nlp_solver_type": "SQP"}, {"nlp_solver_type": "SQP_RTI"}]},
                  surface_coeff_sets=[{"a": -0.1 + 0.01 * k} for k in range(5)])
    sims = m.build_simulators()
    assert len(sims) == 24 * 5
    plan = m.plan_groups(sims, world_size=1)
    assert len(plan) == 6 and all(len(g["indices"]) == 20 for g in plan)  # README grid -> 6 launches by (N, type)
    plan = m.plan_groups(sims, world_size=3, max_batch=4)
    for g in plan:
        assert [len(c) for c in g["rank_chunks"]] == [7, 7, 6]
        assert sum(g["rank_chunks"], []) == g["indices"]
        assert all(len(b) <= 4 for r in g["launches"] for b in r)
        assert sum((sum(r, []) for r in g["launches"]), []) == g["indices"]


def test_summary_from_row_matches_reference_formulas():
    rng = np.random.default_rng(3)
    e = rng.standard_normal((61, 5)) * 0.1
    dt = 0.01
    row = np.zeros(rm.RS["COUNT"])
    row[0:5] = (e ** 2).sum(0)
    row[5:10] = ((np.arange(61) * dt)[:, None] * e ** 2).sum(0) * dt
    row[rm.RS["NSAMPLES"]] = 61
    row[rm.RS["SQP_ITERS"]] = 120
    row[rm.RS["NFAIL"]] = 2
    row[rm.RS["MAX_RES"]:rm.RS["MAX_RES"] + 4] = [1e-7, 3e-7, 0, 2e-7]
    s = rm.summary_from_row(row, [50.0] * 5)
    assert set(s) == set(rm.simulator.SUMMARY_KEYS)
    for k in range(5):
        assert np.isclose(s[f"rmse_e{k + 1}"], np.sqrt(np.mean(e[:, k] ** 2)))  # R/simulator.py:368-374
        assert np.isclose(s[f"itse_e{k + 1}"], np.sum(np.arange(61) * dt * e[:, k] ** 2) * dt)  # :361-366
    assert np.isclose(s["weighted_rmse"], np.sqrt(np.mean(50 * (e ** 2).sum(1))))  # :377-379
    assert s["avg_sqp_iterations"] == 2.0 and s["num_failures"] == 2 and s["max_kkt_residual"] == 3e-7


def test_shard_bounds():
    assert rdist.shard_bounds(10, 4) == [(0, 3), (3, 6), (6, 8), (8, 10)]
    assert rdist.shard_bounds(2, 4) == [(0, 1), (1, 2), (2, 2), (2, 2)]
    assert rdist.shard_bounds(4096, 8)[-1] == (3584, 4096)
    assert rdist.world() == (0, 1)
    st = rdist.reduce_group_stats([1.0, 3.0, 10.0], [0, 0, 1], 2)  # single process: local reduction
    assert list(st["count"]) == [2, 1] and np.allclose(st["mean"], [2.0, 10.0]) and np.allclose(st["var"], [1.0, 0.0])


def _gloo_worker(rank, ws, port, q):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    try:
        assert rdist.world() == (rank, ws)
        n = 7  # uneven: rank 0 holds 4 rows, rank 1 holds 3
        b, e = rdist.shard_bounds(n, ws)[rank]
        full = np.arange(n * 20, dtype=float).reshape(n, 20)
        out = rdist.all_gather_rows(full[b:e], n)
        ok = np.array_equal(out, full)
        # planning is identical on every rank and the chunks tile the group
        m = rm.SimulationManager(BASE)
        m.sweep("surface_coeffs.a", [0.01 * k for k in range(5)])
        plan = m.plan_groups(m.build_simulators(), ws)
        ok = ok and plan[0]["rank_chunks"] == [[0, 1, 2], [3, 4]]
        # per-group statistics over all ranks (tracking-error stats per horizon of the horizon sweep)
        vals, grp = np.arange(n, dtype=float) ** 2, np.array([0, 1, 0, 1, 0, 1, 3])
        st = rdist.reduce_group_stats(vals[b:e], grp[b:e], 4)
        ok = ok and list(st["count"]) == [3, 3, 0, 1]
        ok = ok and np.allclose(st["mean"][[0, 1, 3]], [vals[grp == 0].mean(), vals[grp == 1].mean(), 36.0])
        ok = ok and np.allclose(st["var"][[0, 1]], [vals[grp == 0].var(), vals[grp == 1].var()])
        ok = ok and st["min"][1] == 1.0 and st["max"][0] == 16.0 and np.isnan(st["mean"][2])
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def test_two_rank_gather_on_gloo():
    import torch.multiprocessing as mp

    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == [(0, True), (1, True)]


def test_results_export_table_and_arrays(tmp_path):
    """summary table (parquet / csv) and per-simulation arrays (.npz) of a run_all() result list, built
    here from synthetic device logs (no GPU): keys are those of R/simulator.py:450-547."""
    import pandas as pd

    m = rm.SimulationManager(BASE)
    m.grid_search({"prediction_horizon": [20, 30], "solver_options.nlp_solver_type": ["SQP", "SQP_RTI"]},
                  surface_coeff_sets=[{"a": -0.1, "b": 0.1, "c": 0.0, "d": 0.0, "e": 0.0, "f": 0.0}])
    sims = m.build_simulators()
    rng = np.random.default_rng(0)
    results = []
    for i, s in enumerate(sims):
        ns = s.Nsim
        row = np.abs(rng.standard_normal(rm.RS["COUNT"]))
        row[rm.RS["NSAMPLES"]] = ns + 1
        if i % 2 == 0:  # logs on this rank
            slog = np.abs(rng.standard_normal((ns, rm.RL["COUNT"])))
            slog[:, 0] = 0
            s._load_logs(rng.standard_normal((ns + 1, 12)), rng.standard_normal((ns, 6)), slog,
                         rng.standard_normal((ns + 1, rm.RK["COUNT"])), row)
            results.append({"name": s.name, "simulator": s, "data": s.get_data(), "analysis": s.get_analysis(), "summary": s.get_summary()})
        else:           # solved by another rank: summary only
            results.append({"name": s.name, "simulator": None, "data": None, "analysis": None,
                            "summary": rm.summary_from_row(row, s.mpc.task_weights(), ns)})
    table = m.summary_table(results)
    assert len(table) == 4 and {"name", "prediction_horizon", "solver_options.nlp_solver_type", "surface_coeffs.a", "rmse_e1",
                                "weighted_rmse", "num_failures", "total_computation_time"} <= set(table.columns)
    assert list(table["prediction_horizon"]) == [20, 20, 30, 30]
    files = m.export(tmp_path / "out", results)
    assert [f.name for f in files[:2]] == ["summaries.parquet", "summaries.csv"] and len(files) == 4
    back = pd.read_parquet(files[0])
    assert np.allclose(back["rmse_e3"], table["rmse_e3"]) and list(back["name"]) == [r["name"] for r in results]
    z = np.load(files[2])
    d = results[0]["data"]
    assert np.array_equal(z["data.q"], d["q"]) and np.array_equal(z["data.ee_pose"], d["ee_pose"]) and z["data.N"] == 20
    assert np.array_equal(z["analysis.e1"], results[0]["analysis"]["e1"])
    assert np.isclose(z["analysis.rmse.e2"], results[0]["analysis"]["rmse"]["e2"])
    assert np.isclose(z["summary.weighted_rmse"], results[0]["summary"]["weighted_rmse"])
