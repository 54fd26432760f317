"""UR5 through the reference-facing Simulator API on the GPU: 64 random surfaces x {ur5, ur10}, 200 closed-loop steps."""
import json, sys, time
import numpy as np
sys.path.insert(0, ".")
import robotic_mpc_b200 as rm

qdm = np.array([2.16, 2.16, np.pi, 3.20, 3.20, 3.20])
base = dict(robot_name="ur5", dt=0.01, simulation_time=2.0, prediction_horizon=20, q_0=np.array([0.0, -1.8, 1.9, -1.7, -1.57, 0.0]),
            qdot_0=np.array([0.2, 0.1, -0.1, 0, 0, 0]), wcv=np.full(6, 200.0), q_min=np.full(6, -2 * np.pi), q_max=np.full(6, 2 * np.pi),
            qdot_min=-qdm, qdot_max=qdm, surface_coeffs={"a": -0.1, "b": 0.1, "c": -0.01, "d": 0.01, "e": 0.01, "f": 0.0}, scene=False,
            solver_options={"nlp_solver_type": "SQP_RTI"})
rng = np.random.default_rng(3)
sets = [{k: float(v + 0.01 * rng.standard_normal()) for k, v in base["surface_coeffs"].items()} for _ in range(64)]
mgr = rm.SimulationManager(base)
mgr.grid_search({"robot_name": ["ur5", "ur10"]}, surface_coeff_sets=sets)
t0 = time.time()
res = mgr.run_all()
S = mgr.last_summaries
out = {"run_info": mgr.last_run_info, "wall_s": time.time() - t0}
for name in ("ur5", "ur10"):
    rows = [r for r in res if r["simulator"].robot_name == name]
    out[name] = {"n": len(rows), "failures": int(sum(r["summary"]["num_failures"] for r in rows)),
                 "rmse_e1_mean": float(np.mean([r["summary"]["rmse_e1"] for r in rows])),
                 "final_e1_abs_mean": float(np.mean([abs(r["analysis"]["e1"][-1]) for r in rows])),
                 "final_e4_abs_mean": float(np.mean([abs(r["analysis"]["e4"][-1]) for r in rows]))}
print(json.dumps(out))
