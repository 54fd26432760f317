"""Closed-loop simulator and experiment manager -- host-side mirror of ``R/simulator.py`` (``Simulator``,
``SimulationManager``) over the batched engine.

Same constructor keywords, attributes, method names and result-dict keys as the reference
(``R/simulator.py:18-35,59-65,265-547,562-676``); what changes is WHERE the work happens:

* ``SimulationManager.run_all`` groups the queued configurations by static shape (horizon, dt, number of steps, solver
  options), and runs each group as ONE batch on the GPU: a single ``rmpc_closed_loop`` launch carries every instance of
  the group through all ``Nsim`` closed-loop steps (solve, plant RK4, logging) with no host round trip.  The reference
  runs them one after another, rebuilding a solver each time and sleeping to real time (``R/simulator.py:228-238,654-674``).
* Under ``torchrun`` each rank takes a contiguous chunk of every group; the summary rows are all-gathered at the end.
* ``Simulator.run()`` on its own is a batch of one.

No CPU path: without the CUDA library / a GPU, ``run()`` raises ``RmpcError``.  No sleeps and no MeshCat scene
(``scene`` is accepted and ignored: there is no graphics stack).
"""
from __future__ import annotations

import time
from itertools import product

import numpy as np

from . import distributed as rdist
from .engine import RMPC_ENOMEM, RmpcError
from .params import RK, RL, RS
from .prediction_model import SixDofRobot
from .simulation_model import Robot
from .surface import Surface
from .trajectory_optimizer import MPC, bind_chain, bind_terminal

SUMMARY_KEYS = ("rmse_e1", "rmse_e2", "rmse_e3", "rmse_e4", "rmse_e5", "itse_e1", "itse_e2", "itse_e3", "itse_e4",
                "itse_e5", "weighted_rmse", "total_sqp_iterations", "avg_sqp_iterations", "num_failures",
                "max_kkt_residual", "total_solver_time", "avg_mpc_time", "avg_solver_time", "avg_integration_time",
                "total_computation_time")


def _solver_options_dict(opts):
    """The reference passes a dict, or -- in R/main.py:29-32 -- a list holding one dict."""
    if opts is None:
        return {}
    if isinstance(opts, (list, tuple)):
        if len(opts) != 1 or not isinstance(opts[0], dict):
            raise TypeError("solver_options must be a dict (or a list holding one dict)")
        return dict(opts[0])
    return dict(opts)


class Simulator:
    def __init__(self, robot_name, dt, simulation_time, prediction_horizon, q_0, qdot_0, wcv, q_min, q_max, qdot_min, qdot_max,
                 surface_limits=None, surface_origin=None, surface_orientation_rpy=None, w_qddot=0.02, surface_coeffs=None,
                 solver_options=None, w_u=0.01, px_ref=0.40, vy_ref=0.05, scene=True, terminal_constraint=None,
                 shift_initial_guess=False):
        self.name = robot_name
        self.robot_name = robot_name
        self.dt = dt
        self.simulation_time = simulation_time
        self.Nsim = int(simulation_time / dt)
        self.prediction_horizon = prediction_horizon
        self.q_0, self.qdot_0 = np.asarray(q_0, float), np.asarray(qdot_0, float)
        self.wcv = np.asarray(wcv, float)
        self.q_min, self.q_max = np.asarray(q_min, float), np.asarray(q_max, float)
        self.qdot_min, self.qdot_max = np.asarray(qdot_min, float), np.asarray(qdot_max, float)
        self.w_u, self.w_qddot = w_u, w_qddot
        self.initial_state = np.hstack((self.q_0, self.qdot_0))
        self.px_ref, self.vy_ref = px_ref, vy_ref

        self.mpc_time = np.zeros(self.Nsim)
        self.integration_time = np.zeros(self.Nsim)
        self.sqp_iter = np.zeros(self.Nsim, dtype=int)
        self.solver_status = np.zeros(self.Nsim, dtype=int)
        self.residuals = np.zeros((self.Nsim, 4))
        self.solver_time = np.zeros(self.Nsim)
        self.cost_history = np.zeros(self.Nsim)
        self.qp_iter = np.zeros(self.Nsim, dtype=int)  # extra: interior-point iterations per step

        self.surface = Surface(position=surface_origin, orientation_rpy=surface_orientation_rpy, limits=surface_limits,
                               coefficients=surface_coeffs)
        self.simulation_model = Robot(z0=self.initial_state, u0=self.qdot_0, dt=self.dt, Nsim=self.Nsim, wcv=self.wcv,
                                      integration_method="RK4")
        self.prediction_model = SixDofRobot(Ts=self.dt, Wcv=self.wcv, robot_name=robot_name)
        self.translation = self.prediction_model.translation
        self.mpc = MPC(surface=self.surface, initial_state=self.initial_state, model=self.prediction_model,
                       N_horizon=self.prediction_horizon, Tf=self.dt * self.prediction_horizon, qmin=self.q_min,
                       qmax=self.q_max, dq_min=self.qdot_min, dq_max=self.qdot_max, w_u=self.w_u, w_qddot=self.w_qddot,
                       px_ref=self.px_ref, vy_ref=self.vy_ref, terminal_constraint=terminal_constraint)
        self.mpc.shift_initial_guess = bool(shift_initial_guess)  # extension: primal shift between steps (off = reference)
        so = _solver_options_dict(solver_options)
        if so:
            self._apply_solver_options(so)
        self.mpc.ocp.solver_options.validate()
        self.scene = None  # no graphics stack on a GPU box; the MeshCat path of the reference is out of scope
        self._data_computed = False
        self._summary_row = None
        self._cache = {}

    def _apply_solver_options(self, options):
        for key, value in options.items():
            if hasattr(self.mpc.ocp.solver_options, key):
                setattr(self.mpc.ocp.solver_options, key, value)
            else:
                print(f"Warning: Unknown solver option '{key}'")

    # ------------------------------------------------------------------ batching hooks
    def static_key(self):
        return (self.Nsim,) + self.mpc.static_key()

    def _invalidate_cache(self):
        self._cache = {}

    def _load_logs(self, zlog, ulog, slog, plog, summary_row):
        """Per-instance slices of the device logs of a finished closed loop."""
        self.simulation_model._load(zlog, ulog, plog, RK)
        self.solver_status = slog[:, RL["STATUS"]].astype(int)
        self.sqp_iter = slog[:, RL["SQP_ITER"]].astype(int)
        self.qp_iter = slog[:, RL["QP_ITER"]].astype(int)
        self.cost_history = slog[:, RL["COST"]].copy()
        self.residuals = slog[:, RL["RES"]:RL["RES"] + 4].copy()
        self.solver_time = slog[:, RL["TIME"]].copy()  # device time of this instance's solve (GPU global timer)
        self.mpc_time = self.solver_time.copy()        # set x0 / solve / get u0 are one fused device step
        self.integration_time = np.zeros(self.Nsim)    # the RK4 plant step is fused into the same kernel
        self._errors_log = plog[:, RK["ERR"]:RK["ERR"] + 5].copy()
        self._ptz_log = plog[:, RK["PTZ"]].copy()
        self._summary_row = np.asarray(summary_row, float)
        self._invalidate_cache()
        self._data_computed = True

    def _load_summary_only(self, summary_row):
        self._summary_row = np.asarray(summary_row, float)
        self._invalidate_cache()
        self._data_computed = "summary"

    def run(self, device=0):
        """Run the closed loop of this one instance on the GPU (a batch of one)."""
        self._invalidate_cache()
        run_batch([self], device=device)

    # ------------------------------------------------------------------ results (R/simulator.py:265-448)
    def _require(self, what, full=True):
        if not self._data_computed or (full and self._data_computed == "summary"):
            raise RuntimeError(f"Must call run() before accessing {what}" if not self._data_computed else
                               f"{what} needs the per-step logs; this run kept summaries only")

    @property
    def errors(self):
        self._require("errors")
        if "errors" not in self._cache:
            e = self._errors_log
            self._cache["errors"] = {"e1": e[:, 0], "e2": e[:, 1], "e3": e[:, 2], "e4": e[:, 3], "e5": e[:, 4],
                                     "p_task_z": self._ptz_log, "p_ee_y": self.simulation_model._ee_pose_log[1, :]}
        return self._cache["errors"]

    @property
    def metrics(self):
        self._require("metrics")
        if "metrics" not in self._cache:
            errors = self.errors
            itse, rmse = {}, {}
            for key in ("e1", "e2", "e3", "e4", "e5"):
                sig = errors[key]
                tvec = np.arange(len(sig)) * self.dt
                itse[key] = np.sum(tvec * sig ** 2) * self.dt
                rmse[key] = float(np.sqrt(np.mean(sig ** 2)))
            w = self.mpc.task_weights()
            wsum = sum(w[m] * errors[f"e{m + 1}"] ** 2 for m in range(5))
            self._cache["metrics"] = {"weighted_rmse": float(np.sqrt(np.mean(wsum))), "rmse": rmse, "itse": itse}
        return self._cache["metrics"]

    @property
    def solver_stats(self):
        self._require("solver_stats")
        kkt = np.max(self.residuals, axis=1)
        return {"sqp_iterations": self.sqp_iter, "solver_status": self.solver_status, "residuals": self.residuals,
                "kkt_residuals": kkt, "res_stat": self.residuals[:, 0], "res_eq": self.residuals[:, 1],
                "res_ineq": self.residuals[:, 2], "res_comp": self.residuals[:, 3], "cost_history": self.cost_history,
                "total_sqp_iterations": int(np.sum(self.sqp_iter)), "avg_sqp_iterations": float(np.mean(self.sqp_iter)),
                "num_failures": int(np.sum(self.solver_status != 0)), "total_solver_time": float(np.sum(self.solver_time)),
                "max_kkt_residual": float(np.max(kkt))}

    @property
    def timings(self):
        self._require("timings")
        total = self.mpc_time + self.integration_time
        return {"mpc_time": self.mpc_time, "integration_time": self.integration_time, "solver_time": self.solver_time,
                "total_computation_time": total, "avg_mpc_time": float(np.mean(self.mpc_time)),
                "avg_solver_time": float(np.mean(self.solver_time)),
                "avg_integration_time": float(np.mean(self.integration_time)), "avg_total_time": float(np.mean(total)),
                "computational_time_sim": float(np.sum(total))}

    def get_data(self):
        self._require("data")
        q, qdot, u = self.simulation_model.z[:6, :], self.simulation_model.z[6:, :], self.simulation_model.u
        wcv = np.asarray(self.wcv).reshape(-1, 1)
        qddot = -wcv * qdot + wcv * u
        qddot_fd = np.empty_like(qdot)
        qddot_fd[:, :-1] = (qdot[:, 1:] - qdot[:, :-1]) / self.dt
        qddot_fd[:, -1] = qddot_fd[:, -2]
        return {"time": np.arange(q.shape[1]) * self.dt, "q": q, "qdot": qdot, "qddot": qddot, "qddot_fd": qddot_fd, "u": u,
                "ee_pose": self.simulation_model._ee_pose_log, "px_ref": self.mpc.px_ref, "vy_ref": self.mpc.vy_ref,
                "N": self.prediction_horizon}

    def get_analysis(self):
        self._require("analysis")
        return {**self.errors, **self.metrics, **self.solver_stats, **self.timings}

    def get_summary(self):
        """The 20 scalars of R/simulator.py:509-547.  With per-step logs they are computed exactly as the reference does;
        a summaries-only run derives them from the sums the kernel accumulated (RS_* row)."""
        self._require("summary", full=False)
        if self._data_computed == "summary":
            return summary_from_row(self._summary_row, self.mpc.task_weights())
        m, s, t = self.metrics, self.solver_stats, self.timings
        out = {f"rmse_e{k}": m["rmse"][f"e{k}"] for k in range(1, 6)}
        out.update({f"itse_e{k}": m["itse"][f"e{k}"] for k in range(1, 6)})
        out.update({"weighted_rmse": m["weighted_rmse"], "total_sqp_iterations": s["total_sqp_iterations"],
                    "avg_sqp_iterations": s["avg_sqp_iterations"], "num_failures": s["num_failures"],
                    "max_kkt_residual": s["max_kkt_residual"], "total_solver_time": s["total_solver_time"],
                    "avg_mpc_time": t["avg_mpc_time"], "avg_solver_time": t["avg_solver_time"],
                    "avg_integration_time": t["avg_integration_time"], "total_computation_time": t["computational_time_sim"]})
        return out


def _flatten_arrays(prefix, d, out):
    """Numeric leaves of a (nested) result dict as a flat {name: ndarray} for numpy.savez."""
    for k, v in d.items():
        key = f"{prefix}{k}"
        if isinstance(v, dict):
            _flatten_arrays(key + ".", v, out)
        elif v is not None:
            a = np.asarray(v)
            if a.dtype.kind in "fiub":
                out[key] = a


def save_result_npz(path, result):
    """One simulation's ``get_data()`` / ``get_analysis()`` / ``get_summary()`` arrays (the inputs of the reference's plot
    pipeline, R/main.py:84-342) as one ``.npz``: keys ``data.<k>``, ``analysis.<k>`` (nested dicts dotted), ``summary.<k>``."""
    out = {}
    for part in ("data", "analysis", "summary"):
        if result.get(part) is not None:
            _flatten_arrays(part + ".", result[part], out)
    np.savez_compressed(path, name=np.array(result.get("name", "")), **out)
    return sorted(out)


def summary_from_row(row, w_task, nsteps=None):
    """get_summary() keys from a device summary row (RS_*): RMSE = sqrt(SSE / samples), ITSE as accumulated."""
    row = np.asarray(row, float)
    ns = max(row[RS["NSAMPLES"]], 1.0)
    steps = (ns - 1.0) if nsteps is None else float(nsteps)
    sse, itse = row[RS["SSE"]:RS["SSE"] + 5], row[RS["ITSE"]:RS["ITSE"] + 5]
    out = {f"rmse_e{k + 1}": float(np.sqrt(sse[k] / ns)) for k in range(5)}
    out.update({f"itse_e{k + 1}": float(itse[k]) for k in range(5)})
    out["weighted_rmse"] = float(np.sqrt(np.dot(np.asarray(w_task, float), sse) / ns))
    out["total_sqp_iterations"] = int(row[RS["SQP_ITERS"]])
    out["avg_sqp_iterations"] = float(row[RS["SQP_ITERS"]] / max(steps, 1.0))
    out["num_failures"] = int(row[RS["NFAIL"]])
    out["max_kkt_residual"] = float(np.max(row[RS["MAX_RES"]:RS["MAX_RES"] + 4]))
    for k in ("total_solver_time", "avg_mpc_time", "avg_solver_time", "avg_integration_time", "total_computation_time"):
        out[k] = float("nan")  # per-step device times live in the step log, which a summaries-only run does not keep
    return out


class _BatchRun:
    """One batched closed loop in flight: engine + CUDA stream + the simulators it serves.  ``start`` uploads the problem
    data, ``step(n)`` enqueues n closed-loop steps (asynchronous), ``finish`` downloads the results into the Simulator
    objects and frees the engine.  Several of them can be in flight on different streams of one GPU."""

    def __init__(self, sims, device=0, keep_logs=True, stream=None):
        from .engine import Engine
        from .trajectory_optimizer import bind_chain, bind_terminal
        import torch

        key = sims[0].static_key()
        if any(s.static_key() != key for s in sims):
            raise ValueError("run_batch: all simulators must share Nsim, horizon, dt, solver options, arm and terminal-set flag")
        self.sims, self.keep_logs, self.torch = sims, keep_logs, torch
        s0 = sims[0]
        o = s0.mpc.ocp.solver_options
        self.Nsim, self.done = s0.Nsim, 0
        self.stream = stream
        self.eng = Engine(len(sims), s0.prediction_horizon, o.nlp_solver_type, dt=s0.mpc.Tf / s0.mpc.N_horizon,
                          nlog=self.Nsim if keep_logs else 0, device=device, max_iter=o.nlp_solver_max_iter,
                          qp_iter_max=o.qp_solver_iter_max,
                          tol=[o.nlp_solver_tol_stat, o.nlp_solver_tol_eq, o.nlp_solver_tol_ineq, o.nlp_solver_tol_comp],
                          qp_tol=o.qp_tol, plant_dt=s0.dt)
        try:
            with self._on_stream():
                bind_chain(self.eng, [s.mpc for s in sims])
                bind_terminal(self.eng, [s.mpc for s in sims])
                self.eng.set_params(np.stack([s.mpc.param_row() for s in sims]))
                self.eng.set_state(np.stack([s.initial_state for s in sims]))
                self.eng.reset()
        except Exception:
            self.eng.close()
            raise

    def _on_stream(self):
        import contextlib

        return self.torch.cuda.stream(self.stream) if self.stream is not None else contextlib.nullcontext()

    def step(self, n=None):
        n = self.Nsim - self.done if n is None else min(int(n), self.Nsim - self.done)
        if n > 0:
            with self._on_stream():
                self.eng.closed_loop(n)
            self.done += n
        return self.done < self.Nsim

    def finish(self):
        try:
            self.step()
            with self._on_stream():
                summ = self.eng.get("summary").cpu().numpy()
                if self.keep_logs:
                    zlog, ulog = self.eng.get("zlog").cpu().numpy(), self.eng.get("ulog").cpu().numpy()
                    slog, plog = self.eng.get("slog").cpu().numpy(), self.eng.get("plog").cpu().numpy()
            for i, s in enumerate(self.sims):
                if self.keep_logs:
                    s._load_logs(zlog[i], ulog[i], slog[i], plog[i], summ[i])
                else:
                    s._load_summary_only(summ[i])
        finally:
            self.eng.close()
        return summ


def run_batch(sims, device=0, keep_logs=True):
    """Run the closed loops of ``sims`` (all sharing one static key) as one batch on one GPU and load the results into
    the Simulator objects.  Returns the ``[B, RS_COUNT]`` summary array."""
    if not sims:
        return np.zeros((0, RS["COUNT"]))
    return _BatchRun(sims, device=device, keep_logs=keep_logs).finish()


def run_batches_concurrently(batches, device=0, keep_logs=True, chunk=25):
    """Run several batches (lists of simulators, one static key each) of one GPU at the same time, each on its own CUDA stream:
    a grid search usually yields a handful of groups that are each too small to fill the GPU (1024 instances occupy a fifth of
    a B200), so their kernels are interleaved ``chunk`` closed-loop steps at a time.  Returns the summary arrays in order."""
    import torch

    batches = [b for b in batches]
    if len(batches) <= 1:
        return [run_batch(b, device=device, keep_logs=keep_logs) for b in batches]
    runs = []
    try:
        with torch.cuda.device(device):
            cur = torch.cuda.current_stream()
            try:
                for b in batches:
                    st = torch.cuda.Stream()
                    st.wait_stream(cur)
                    runs.append(_BatchRun(b, device=device, keep_logs=keep_logs, stream=st))
            except (RmpcError, torch.cuda.OutOfMemoryError) as err:
                if isinstance(err, RmpcError) and err.code != RMPC_ENOMEM:
                    raise
                for r in runs:  # not enough device memory for all groups at once: one after the other instead
                    r.eng.close()
                runs = []
                torch.cuda.empty_cache()  # the engines' cudaMalloc competes with torch's cached free blocks
                return [run_batch(b, device=device, keep_logs=keep_logs) for b in batches]
            active = list(runs)
            while active:
                active = [r for r in active if r.step(chunk)]
            out = [r.finish() for r in runs]
            runs = []
            return out
    finally:
        for r in runs:  # only on an error path: free what is still allocated
            r.eng.close()


class SimulationManager:
    """Same queueing API as R/simulator.py:549-716 (plus the two README-only names, R/readme.md:64-69)."""

    def __init__(self, base_config):
        self.base_config = base_config.copy()
        self.simulations = []
        self.last_summaries = None  # [n_sims, RS_COUNT] of the last run_all, gathered over all ranks
        self.last_run_info = None

    def sweep(self, param_path, values, name_template=None):
        for val in values:
            config = self.base_config.copy()
            if "." in param_path:
                head, leaf = param_path.split(".", 1)
                if head == "surface_coeffs":
                    config["surface_coeffs"] = dict(config.get("surface_coeffs") or {})
                    config["surface_coeffs"][leaf] = val
            else:
                config[param_path] = val
            name = name_template.format(val) if name_template else f"{param_path}={val}"
            self.simulations.append({"name": name, "config": config})

    def sweep_parameter(self, param, values, name_template=None):
        """README name for ``sweep`` (R/readme.md:64)."""
        self.sweep(param, values, name_template=name_template)

    def add_manual(self, name, params):
        """README ``add_manual(name=..., params={...})`` (R/readme.md:68-69): one simulation = base config + overrides."""
        config = self.base_config.copy()
        config.update(params)
        self.simulations.append({"name": name, "config": config})

    def grid_search(self, param_grid, surface_coeff_sets=None, name_template=None):
        paths = list(param_grid.keys())
        value_lists = [param_grid[p] for p in paths]
        coeff_sets = [None] if surface_coeff_sets is None else surface_coeff_sets
        for coeff_idx, coeff_set in enumerate(coeff_sets):
            for values in product(*value_lists):
                config = self.base_config.copy()
                for path, val in zip(paths, values):
                    if "." in path:
                        head, leaf = path.split(".", 1)
                        config[head] = dict(config.get(head) or {})
                        config[head][leaf] = val
                    else:
                        config[path] = val
                if coeff_set is not None:
                    config["surface_coeffs"] = coeff_set.copy()
                chosen = dict(zip(paths, values))
                if name_template:
                    try:
                        name = name_template(chosen, coeff_idx)
                    except TypeError:
                        name = name_template(chosen)
                else:
                    pstr = "_".join(f"{p.split('.')[-1]}={v}" for p, v in zip(paths, values))
                    name = f"coeffs{coeff_idx}_{pstr}" if coeff_set is not None else pstr
                self.simulations.append({"name": name, "config": config})

    def clear(self):
        self.simulations = []

    # ------------------------------------------------------------------ results surface
    def summary_table(self, results):
        """One row per simulation: ``name``, the scalar entries of its queued config (nested ``solver_options`` /
        ``surface_coeffs`` dotted) and the 20 ``get_summary()`` scalars (R/simulator.py:509-547) as a pandas DataFrame --
        the table the comparison plots of R/main.py are drawn from."""
        import pandas as pd

        rows = []
        for spec, res in zip(self.simulations, results):
            row = {"name": res["name"]}
            for k, v in spec["config"].items():
                if isinstance(v, (list, tuple)) and len(v) == 1 and isinstance(v[0], dict):
                    v = v[0]
                if isinstance(v, dict):
                    row.update({f"{k}.{kk}": vv for kk, vv in v.items() if np.isscalar(vv)})
                elif np.isscalar(v):
                    row[k] = v
            row.update(res["summary"])
            rows.append(row)
        return pd.DataFrame(rows)

    def export(self, directory, results, arrays=True):
        """Write ``summaries.parquet`` (+ ``summaries.csv``) and, with ``arrays``, one ``<index>_<name>.npz`` per simulation
        whose logs are on this rank (``save_result_npz``).  Returns the list of files written."""
        from pathlib import Path

        d = Path(directory)
        d.mkdir(parents=True, exist_ok=True)
        table = self.summary_table(results)
        files = [d / "summaries.parquet", d / "summaries.csv"]
        table.to_parquet(files[0], index=False)
        table.to_csv(files[1], index=False)
        if arrays:
            for i, res in enumerate(results):
                if res.get("data") is None:
                    continue
                safe = "".join(c if (c.isalnum() or c in "-_.=") else "_" for c in str(res["name"]))[:80]
                f = d / f"{i:05d}_{safe}.npz"
                save_result_npz(f, res)
                files.append(f)
        return files

    # ------------------------------------------------------------------ planning (pure host logic)
    def build_simulators(self):
        sims = []
        for spec in self.simulations:
            sim = Simulator(**spec["config"].copy())
            sim.name = spec["name"]
            sims.append(sim)
        return sims

    @staticmethod
    def plan_groups(sims, world_size=1, max_batch=None):
        """Group simulator indices by static key (order of first appearance), deal each group to the ranks in contiguous
        balanced chunks, and split chunks larger than ``max_batch``.  Returns
        ``[{'key', 'indices', 'rank_chunks': [[idx...] per rank], 'launches': [[idx...], ...] per rank}]``."""
        groups = {}
        for i, s in enumerate(sims):
            groups.setdefault(s.static_key(), []).append(i)
        plan = []
        for key, idx in groups.items():
            chunks = [idx[b:e] for b, e in rdist.shard_bounds(len(idx), world_size)]
            launches = []
            for ch in chunks:
                if max_batch and len(ch) > max_batch:
                    launches.append([ch[k:k + max_batch] for k in range(0, len(ch), max_batch)])
                else:
                    launches.append([ch] if ch else [])
            plan.append({"key": key, "indices": idx, "rank_chunks": chunks, "launches": launches})
        return plan

    def run_all(self, return_results=True, device=None, max_batch=None, keep_logs=True, concurrent=True):
        """Run every queued simulation; one batched GPU run per (static shape, rank), all of a rank's runs in flight at the same
        time on separate CUDA streams unless ``concurrent=False`` or ``max_batch`` splits them (memory-bound sweeps).

        ``return_results=True`` returns the reference's list of ``{'name','simulator','data','analysis','summary'}``
        (for instances solved by another rank only ``name`` and ``summary`` are filled).  ``keep_logs=False`` keeps
        only the device-accumulated summaries (large sweeps)."""
        rank, ws = rdist.world()
        if device is None:
            import torch

            device = torch.cuda.current_device() if torch.cuda.is_available() else 0
        t0 = time.time()
        sims = self.build_simulators()
        plan = self.plan_groups(sims, ws, max_batch)
        summaries = np.zeros((len(sims), RS["COUNT"]))
        launches = 0
        concurrent = concurrent and not max_batch
        if concurrent:
            # all launches of this rank in flight at once, one stream each (groups of a grid search rarely fill the GPU alone)
            flat = [[sims[i] for i in batch_idx] for g in plan for batch_idx in g["launches"][rank]]
            rows = run_batches_concurrently(flat, device=device, keep_logs=keep_logs)
            launches = len(flat)
            it = iter(rows)
        for g in plan:
            local_rows = []
            for batch_idx in g["launches"][rank]:
                if concurrent:
                    local_rows.append(next(it))
                else:
                    local_rows.append(run_batch([sims[i] for i in batch_idx], device=device, keep_logs=keep_logs))
                    launches += 1
            local = np.concatenate(local_rows, 0) if local_rows else np.zeros((0, RS["COUNT"]))
            allrows = rdist.all_gather_rows(local, len(g["indices"])) if ws > 1 else local
            summaries[g["indices"]] = allrows
        self.last_summaries = summaries
        self.last_run_info = {"groups": len(plan), "launches": launches, "world_size": ws, "seconds": time.time() - t0}
        if not return_results:
            return None
        results = []
        for i, sim in enumerate(sims):
            if sim._data_computed is True:
                results.append({"name": sim.name, "simulator": sim, "data": sim.get_data(), "analysis": sim.get_analysis(),
                                "summary": sim.get_summary()})
            else:
                results.append({"name": sim.name, "simulator": sim if sim._data_computed else None, "data": None,
                                "analysis": None, "summary": summary_from_row(summaries[i], sim.mpc.task_weights(), sim.Nsim)})
        return results
