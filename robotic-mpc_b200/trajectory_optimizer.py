"""OCP formulation and solver handles -- host-side mirror of ``R/trajectory_optimizer.py`` (class ``MPC``) and of the
``AcadosOcpSolver`` methods the reference calls on it (``T/acados_ocp_solver.py:468-476,939-989,1461-1723,1743-1819``).

``MPC`` keeps the reference's constructor signature and attributes (``w_*``, ``px_ref``, ``vy_ref``,
``ocp.solver_options.<attr>``, ``finalize_solver()``, ``.solver``).  Nothing is code-generated: the formulation is fixed
(UR10 surface tracking) and lives in the CUDA kernels; the objects below only carry its
per-instance DATA to the engine's parameter block.

``BatchedSolver`` is the solver interface for B instances at once (same method names, ``[B, ...]`` arrays);
``SolverView`` is one instance's window onto it so that ``sim.mpc.solver.get(0, 'u')`` keeps working.
There is no CPU path: creating a solver without the CUDA library / a GPU raises ``RmpcError``.
"""
from __future__ import annotations

import uuid

import numpy as np

from .params import RL, RQ, pack_params

_SUPPORTED_FIXED = {  # option -> the only value(s) the engine implements (the reference never uses another one)
    "qp_solver": ("PARTIAL_CONDENSING_HPIPM",),
    "hessian_approx": ("GAUSS_NEWTON",),
    "integrator_type": ("DISCRETE",),
    "globalization": ("MERIT_BACKTRACKING",),
    "qp_solver_warm_start": (2,),
    "nlp_solver_warm_start_first_qp": (True,),
}


class SolverOptions:
    """The subset of ``AcadosOcpOptions`` the hot path reads, with the values ``MPC.__init__`` / acados resolve
    (``R/trajectory_optimizer.py:60-70``, ``T/acados_ocp_options.py:45-165``)."""

    def __init__(self, N_horizon=None, tf=None):
        self.nlp_solver_type = "SQP"
        self.hessian_approx = "GAUSS_NEWTON"
        self.print_level = 0
        self.qp_solver = "PARTIAL_CONDENSING_HPIPM"
        self.qp_tol = 1e-8
        self.integrator_type = "DISCRETE"
        self.nlp_solver_warm_start_first_qp = True
        self.qp_solver_warm_start = 2
        self.nlp_solver_max_iter = 100
        self.globalization = "MERIT_BACKTRACKING"
        self.N_horizon = N_horizon
        self.tf = tf
        self.qp_solver_iter_max = 50
        self.qp_solver_cond_N = None
        self.nlp_solver_tol_stat = 1e-6
        self.nlp_solver_tol_eq = 1e-6
        self.nlp_solver_tol_ineq = 1e-6
        self.nlp_solver_tol_comp = 1e-6
        self.as_rti_level = 4  # STANDARD_RTI (T/acados_ocp_options.py:137,1189-1208); 0..3 = AS-RTI level A..D
        self.as_rti_iter = 1

    def validate(self):
        if self.nlp_solver_type not in ("SQP", "SQP_RTI"):
            raise NotImplementedError(f"nlp_solver_type {self.nlp_solver_type!r}: the engine implements SQP and SQP_RTI")
        for key, allowed in _SUPPORTED_FIXED.items():
            if getattr(self, key) not in allowed:
                raise NotImplementedError(f"solver option {key}={getattr(self, key)!r} is not implemented "
                                          f"(supported: {allowed}); see DESIGN.md 'out of scope'")
        if self.as_rti_level not in (0, 1, 2, 3, 4):
            raise ValueError('Invalid as_rti_level value must be in [0, 1, 2, 3, 4].')
        if not (isinstance(self.as_rti_iter, int) and self.as_rti_iter >= 0):
            raise ValueError('Invalid as_rti_iter value. as_rti_iter must be a nonnegative int.')
        if self.as_rti_level != 4 and self.nlp_solver_type != "SQP_RTI":
            raise ValueError("as_rti_level != 4 needs nlp_solver_type 'SQP_RTI'")
        if self.qp_solver_cond_N is not None and not (isinstance(self.qp_solver_cond_N, int) and 1 <= self.qp_solver_cond_N <= self.N_horizon):
            raise ValueError('Invalid qp_solver_cond_N value. qp_solver_cond_N must be a positive int not larger than N_horizon.')

    def static_key(self):
        """What must be equal for two instances to share one batched launch."""
        return (self.nlp_solver_type, int(self.nlp_solver_max_iter), int(self.qp_solver_iter_max), float(self.qp_tol),
                float(self.nlp_solver_tol_stat), float(self.nlp_solver_tol_eq), float(self.nlp_solver_tol_ineq),
                float(self.nlp_solver_tol_comp), int(self.as_rti_level), int(self.as_rti_iter),
                int(self.qp_solver_cond_N or self.N_horizon or 0))


class _Namespace:
    pass


class MPC:
    """Same constructor as ``R/trajectory_optimizer.py:13-14``.  ``surface`` needs ``coeff_vector()`` (our ``Surface``) or
    a ``coeffs`` dict; ``model`` needs ``Ts`` and ``Wcv`` (our ``SixDofRobot`` stand-in) or may be ``None`` when ``Ts`` /
    ``wcv`` are passed explicitly."""

    def __init__(self, surface, initial_state, model, N_horizon, Tf, qmin, qmax, dq_min, dq_max, w_u, w_qddot,
                 px_ref=0.40, vy_ref=0.01, Ts=None, wcv=None, terminal_constraint=None):
        self.surface = surface
        self.model = model
        self.surface_pos_world = surface.get_position() if hasattr(surface, "get_position") else None
        self.surface_rpy_world = surface.get_orientation_rpy() if hasattr(surface, "get_orientation_rpy") else None
        self.N_horizon = int(N_horizon)
        self.Tf = float(Tf)
        self.px_ref, self.vy_ref = float(px_ref), float(vy_ref)
        self._instance_id = uuid.uuid4().hex[:8]
        # task weights (R/trajectory_optimizer.py:44-48)
        self.w_origin_task = 50.0
        self.w_normal_alignment_task = 50.0
        self.w_x_alignment_task = 50.0
        self.w_fixed_x_task = 50.0
        self.w_fixed_vy_task = 50.0
        self.w_u = float(w_u)
        self.w_qddot = float(w_qddot)
        self.nx, self.nu = 12, 6
        self.ny = self.nx + self.nu
        self.Ts = float(Ts if Ts is not None else getattr(model, "Ts", self.Tf / self.N_horizon))
        self.wcv = np.asarray(wcv if wcv is not None else getattr(model, "Wcv"), dtype=float).reshape(6)
        self.translation = np.asarray(getattr(model, "translation", (0.0, 0.0, 0.1)), dtype=float).reshape(3)
        self.initial_state = np.asarray(initial_state, dtype=float).reshape(12)
        self.qmin, self.qmax = np.asarray(qmin, float).reshape(6), np.asarray(qmax, float).reshape(6)
        self.dq_min, self.dq_max = np.asarray(dq_min, float).reshape(6), np.asarray(dq_max, float).reshape(6)
        self.ocp = _Namespace()
        self.ocp.solver_options = SolverOptions(N_horizon=self.N_horizon, tf=self.Tf)
        self.ocp.constraints = _Namespace()
        self.ocp.constraints.x0 = self.initial_state
        self.ocp.constraints.lbu, self.ocp.constraints.ubu = self.dq_min, self.dq_max
        self.ocp.constraints.lbx, self.ocp.constraints.ubx = self.qmin, self.qmax
        self.solver = None
        self.terminal = None
        # shift the primal initial guess by one stage after every closed-loop step; the reference's loop
        # does not (R/simulator.py:199-241), so the default is off
        self.shift_initial_guess = False
        if terminal_constraint is not None:
            self.set_terminal_constraint(terminal_constraint)

    def set_terminal_constraint(self, spec):
        """Terminal ellipsoid ``lh_e <= x_N' P x_N <= alpha`` at stage N.  The reference computes (P, alpha)
        in ``R/_terminal_constraint.py`` but never wires them into its OCP; this is the acados wiring one would add
        (``ocp.model.con_h_expr_e = x.T @ P @ x``, ``ocp.constraints.lh_e / uh_e``).  ``spec``: a dict with ``P`` (12x12) and
        ``alpha`` (and optionally ``lh``, default ``-alpha``), or a dict with ``rho`` (and optionally ``Q_diag``,
        ``safety``) from which ``terminal_ellipsoid_from_lqr`` builds them for this instance's (Ad, Bd) and boxes."""
        from . import terminal_constraint as tc

        spec = dict(spec)
        if "P" in spec:
            P, alpha = np.asarray(spec["P"], float).reshape(12, 12), float(spec["alpha"])
        else:
            P, alpha = tc.default_terminal_set(wcv=self.wcv, Ts=self.Ts, rho=float(spec["rho"]), q_min=self.qmin, q_max=self.qmax,
                                               qdot_min=self.dq_min, qdot_max=self.dq_max, Q_diag=spec.get("Q_diag"),
                                               safety=float(spec.get("safety", 0.98)))
        lh = float(spec.get("lh", -alpha))
        self.terminal = {"P": 0.5 * (P + P.T), "alpha": alpha, "lh": lh}
        self.ocp.constraints.lh_e, self.ocp.constraints.uh_e = np.array([lh]), np.array([alpha])

    def task_weights(self):
        return np.array([self.w_origin_task, self.w_normal_alignment_task, self.w_x_alignment_task, self.w_fixed_x_task,
                         self.w_fixed_vy_task])

    def param_row(self) -> np.ndarray:
        """This instance's row of the engine parameter block (``RP_*``)."""
        surf = self.surface.coeff_vector() if hasattr(self.surface, "coeff_vector") else \
            np.array([self.surface.coeffs[k] for k in "abcdef"], float)
        return pack_params(1, self.Ts, surf=surf, wcv=self.wcv, w_u=self.w_u, w_qddot=self.w_qddot, px_ref=self.px_ref,
                           vy_ref=self.vy_ref, w_task=self.task_weights(), q_min=self.qmin, q_max=self.qmax,
                           qdot_min=self.dq_min, qdot_max=self.dq_max, t_ee=self.translation)[0]

    def static_key(self):
        o = self.ocp.solver_options
        chain = self.model.chain_key() if hasattr(self.model, "chain_key") else None
        return (self.N_horizon, round(self.Tf / self.N_horizon, 15), self.terminal is not None, chain,
                bool(self.shift_initial_guess)) + o.static_key()

    def finalize_solver(self, device=0):
        """Creates the solver handle (a batch of one).  ``SimulationManager.run_all`` binds many ``MPC`` objects to one
        shared batch instead and never calls this."""
        self.ocp.solver_options.validate()
        bs = BatchedSolver([self], device=device)
        self.solver = bs.view(0)
        return self


def bind_chain(engine, mpcs):
    """Hand the batch-wide settings that are part of the static key to the engine: kinematic chain, shift option, AS-RTI, partial condensing."""
    chain = getattr(mpcs[0].model, "chain", None)
    if chain is not None:
        engine.set_chain(*chain)
    engine.set_option("shift", 1 if mpcs[0].shift_initial_guess else 0)
    o = mpcs[0].ocp.solver_options
    if o.qp_solver_cond_N is not None:
        engine.set_option("qp_cond_N", int(o.qp_solver_cond_N))
    if o.nlp_solver_type == "SQP_RTI" and o.as_rti_level != 4:
        engine.set_option("as_rti_iter", o.as_rti_iter)
        engine.set_option("as_rti_level", o.as_rti_level)


def bind_terminal(engine, mpcs):
    """Hand the terminal sets of a batch to the engine (all instances of a batch have one, or none has)."""
    if mpcs[0].terminal is None:
        return
    engine.set_terminal(np.stack([m.terminal["P"] for m in mpcs]), np.array([m.terminal["alpha"] for m in mpcs]),
                        np.array([m.terminal["lh"] for m in mpcs]))


class BatchedSolver:
    """``AcadosOcpSolver`` for B instances that share (N, dt, solver options): one engine handle on one GPU.

    set / get take and return ``[B, n]`` arrays (numpy or torch); ``solve()`` runs ONE SQP / SQP_RTI solve of every
    instance (``rmpc_solve``) and returns the ``[B]`` status vector."""

    def __init__(self, mpcs, device=0, nlog=1):
        from .engine import Engine

        if not mpcs:
            raise ValueError("BatchedSolver needs at least one MPC")
        key = mpcs[0].static_key()
        for m in mpcs:
            m.ocp.solver_options.validate()
            if m.static_key() != key:
                raise ValueError("all instances of a batch must share N, dt and the solver options")
        self.mpcs = list(mpcs)
        self.B = len(mpcs)
        o = mpcs[0].ocp.solver_options
        self.N = mpcs[0].N_horizon
        self.dt = mpcs[0].Tf / mpcs[0].N_horizon
        self.engine = Engine(self.B, self.N, o.nlp_solver_type, dt=self.dt, nlog=max(int(nlog), 1), device=device,
                             max_iter=o.nlp_solver_max_iter, qp_iter_max=o.qp_solver_iter_max,
                             tol=[o.nlp_solver_tol_stat, o.nlp_solver_tol_eq, o.nlp_solver_tol_ineq, o.nlp_solver_tol_comp],
                             qp_tol=o.qp_tol, plant_dt=mpcs[0].Ts)
        self.engine.set_params(np.stack([m.param_row() for m in mpcs]))
        bind_chain(self.engine, mpcs)
        bind_terminal(self.engine, mpcs)
        self.engine.set_state(np.stack([m.initial_state for m in mpcs]))
        self.engine.reset()
        self._stats = None
        self.rti_phase = 0
        self._last_phase = 0
        if o.nlp_solver_type == "SQP":
            self.engine.set_option("statistics", 1)  # get_stats('statistics') of the last solve, like acados keeps it

    def options_set(self, field, value):
        """``AcadosOcpSolver.options_set`` (T/acados_ocp_solver.py:2220-2340) for the runtime options the hot path has:
        ``rti_phase`` (0 preparation + feedback, 1 preparation, 2 feedback; SQP_RTI only, like acados), ``as_rti_level``
        (0..3 = A..D, 4 = standard RTI) and ``as_rti_iter``."""
        if field in ("as_rti_level", "as_rti_iter"):  # T/acados_ocp_solver.py:2294-2300: int options forwarded to the solver
            if self.mpcs[0].ocp.solver_options.nlp_solver_type != "SQP_RTI":
                raise ValueError(f"AcadosOcpSolver.options_set(): {field!r} needs an SQP_RTI solver")
            self.engine.set_option(field, int(value))
            return
        if field != "rti_phase":
            raise NotImplementedError(f"options_set({field!r}): 'rti_phase', 'as_rti_level' and 'as_rti_iter' can be changed after creation")
        if value not in (0, 1, 2):
            raise ValueError("AcadosOcpSolver.options_set(): argument 'rti_phase' can take only values 0, 1, 2")
        if value and self.mpcs[0].ocp.solver_options.nlp_solver_type != "SQP_RTI":
            raise ValueError("AcadosOcpSolver.options_set(): argument 'rti_phase' can take only value 0 for SQP-type solvers")
        self.rti_phase = int(value)

    def view(self, i):
        return SolverView(self, i)

    # ---- AcadosOcpSolver.set / cost_set / constraints_set (T/acados_ocp_solver.py:1743-1819,1863-2035)
    def set(self, stage, field, value):
        """Iterate fields 'x', 'u', 'pi'; everything else is routed like acados does: cost fields to ``cost_set``,
        constraint fields to ``constraints_set``."""
        if field in ("x", "u", "pi"):
            self.engine.set(field, value, stage=stage)
            self._stats = None
        elif field in ("yref", "W"):
            self.cost_set(stage, field, value)
        elif field in ("lbx", "ubx", "lbu", "ubu"):
            self.constraints_set(stage, field, value)
        elif field == "p":
            raise NotImplementedError("set(stage, 'p', .): the formulation has no runtime parameter vector; the surface is "
                                      "set per instance with set_surface(coeffs)")
        else:
            raise NotImplementedError(f"set(..., {field!r})")

    def cost_set(self, stage, field, value):
        """``AcadosOcpSolver.cost_set`` -> ``ocp_nlp_cost_model_set`` (acados_c/ocp_nlp_interface.c:516-560).
        'yref': ``[B, 18]`` (or ``[18]``) reference of ``stage`` (0..N-1).  'W': ``[18, 18]`` / ``[B, 18, 18]`` diagonal weight
        matrix (or its ``[B, 18]`` diagonal); the engine keeps W per INSTANCE, so a call for any stage sets every stage --
        like the reference, whose W is the same matrix on all stages (R/trajectory_optimizer.py:139-156)."""
        if field == "yref":
            self.engine.set("yref", np.broadcast_to(np.asarray(value, float), (self.B, 18)), stage=stage)
        elif field == "W":
            w = np.asarray(value, float)
            if w.shape[-2:] == (18, 18) and w.ndim >= 2:
                if np.abs(w - np.einsum("...ii->...i", w)[..., None] * np.eye(18)).max() > 0:
                    raise NotImplementedError("cost_set('W'): only diagonal weight matrices (the reference's W is diagonal)")
                w = np.einsum("...ii->...i", w)
            w = np.broadcast_to(w, (self.B, 18))
            for blk in (w[:, 5:11], w[:, 11:17]):
                if np.any(blk < 0):
                    raise ValueError("cost_set('W'): weights must be non-negative")
            self.engine.set("W_diag", w, stage=-1)
        else:
            raise NotImplementedError(f"cost_set(..., {field!r}): fields 'yref' and 'W'")
        self._stats = None

    def constraints_set(self, stage, field, value):
        """``AcadosOcpSolver.constraints_set`` -> ``ocp_nlp_constraints_model_set`` (acados_c/ocp_nlp_interface.c:637-760).
        'lbx' / 'ubx' at stage 0: the initial-state constraint ``[B, 12]`` (the reference sets both to the plant state,
        R/simulator.py:210-211); at stages 1..N-1: bounds on q ``[B, 6]``; 'lbu' / 'ubu' at stages 0..N-1: ``[B, 6]``."""
        if field not in ("lbx", "ubx", "lbu", "ubu"):
            raise NotImplementedError(f"constraints_set(..., {field!r}): fields 'lbx', 'ubx', 'lbu', 'ubu'")
        n = 12 if (field in ("lbx", "ubx") and stage == 0) else 6
        self.engine.set(field, np.broadcast_to(np.asarray(value, float), (self.B, n)), stage=stage)
        self._stats = None

    def set_surface(self, coeffs):
        """Surface coefficients a..f per instance (``[B, 6]``): what the reference bakes into the generated cost function."""
        self.engine.set("surf_coeffs", np.broadcast_to(np.asarray(coeffs, float), (self.B, 6)), stage=-1)
        self._stats = None

    def solve(self):
        self.engine.solve(rti_phase=self.rti_phase)
        self._last_phase = self.rti_phase
        self._stats = self.engine.get("stats").cpu().numpy()
        return self._stats[:, RL["STATUS"]].astype(int)

    # ---- AcadosOcpSolver.get (T/acados_ocp_solver.py:939-989)
    def get(self, stage, field):
        if field not in ("x", "u", "pi", "lam"):
            raise NotImplementedError(f"get(..., {field!r})")
        return self.engine.get(field, stage=stage).cpu().numpy()

    def _need_stats(self):
        if self._stats is None:
            raise RuntimeError("call solve() first")
        return self._stats

    _TIME_FIELDS = {"time_tot": "TIME", "time_lin": "TIME_LIN", "time_qp": "TIME_QP", "time_glob": "TIME_GLOB",
                    "time_qp_xcond": "TIME_XCOND"}
    _ZERO_TIME_FIELDS = ("time_sim", "time_sim_ad", "time_sim_la", "time_reg", "time_qpscaling", "time_solution_sensitivities")

    def get_stats(self, field):
        """``AcadosOcpSolver.get_stats`` (T/acados_ocp_solver.py:1461-1661) for the batch: ``[B]`` arrays; 'statistics' a list of
        B tables in acados' layout (row 0 = iteration index, then for SQP: res_stat, res_eq, res_ineq, res_comp, qp_stat,
        qp_iter, alpha; for SQP_RTI: qp_stat, qp_iter).  Times are DEVICE times of each instance's own solve (GPU global
        timer), split like acados splits its own (AC/ocp_nlp/ocp_nlp_common.c:4610-4681); 'time_qp_phases' adds the
        interior-point breakdown ``[B, 4]`` = factorisation, forward sweeps, backward substitutions, update + residuals."""
        s = self._need_stats()
        if field in ("sqp_iter", "nlp_iter"):
            return s[:, RL["SQP_ITER"]].astype(int)
        if field == "qp_iter":
            return s[:, RL["QP_ITER"]].astype(int)
        if field == "qp_stat":
            return s[:, RL["QP_STATUS"]].astype(int)
        if field == "alpha":
            return s[:, RL["ALPHA"]].copy()
        if field == "residuals":
            return self.get_residuals()
        if field in self._TIME_FIELDS:
            return s[:, RL[self._TIME_FIELDS[field]]].copy()
        if field == "time_qp_solver_call":
            return s[:, RL["TIME_QP"]] - s[:, RL["TIME_XCOND"]]
        if field in self._ZERO_TIME_FIELDS:
            return np.zeros(self.B)  # no integrator, regularisation, scaling or sensitivity module on this path
        if field == "time_preparation":
            return s[:, RL["TIME_LIN"]].copy() if self._last_phase != 2 else np.zeros(self.B)
        if field == "time_feedback":
            return s[:, RL["TIME"]] - (s[:, RL["TIME_LIN"]] if self._last_phase == 0 else 0.0)
        if field == "time_qp_phases":
            return s[:, RL["TIME_QP_FACT"]:RL["TIME_QP_RES"] + 1].copy()
        if field == "record":
            return s.copy()  # the engine's own RL_* record of the last solve
        if field in ("statistics", "stat_m", "stat_n"):
            rti = self.mpcs[0].ocp.solver_options.nlp_solver_type == "SQP_RTI"
            if field == "stat_n":
                return 2 if rti else RQ["COUNT"]
            if field == "stat_m":
                return 2 if rti else self.mpcs[0].ocp.solver_options.nlp_solver_max_iter + 1
            it = s[:, RL["SQP_ITER"]].astype(int)
            if rti:
                return [np.array([[0.0, 1.0], [0.0, s[i, RL["QP_STATUS"]]], [0.0, s[i, RL["QP_ITER"]]]])[:, :it[i] + 1]
                        for i in range(self.B)]
            tab = self.engine.get("sqp_stat").cpu().numpy()
            return [np.vstack([np.arange(it[i] + 1, dtype=float)[None, :], tab[i, :it[i] + 1, :].T]) for i in range(self.B)]
        raise NotImplementedError(f"get_stats({field!r})")

    def get_residuals(self):
        return self._need_stats()[:, RL["RES"]:RL["RES"] + 4].copy()

    def get_cost(self):
        return self._need_stats()[:, RL["COST"]].copy()

    def get_status(self):
        return self._need_stats()[:, RL["STATUS"]].astype(int)


class SolverView:
    """One instance of a ``BatchedSolver`` behind the scalar ``AcadosOcpSolver`` method names."""

    def __init__(self, batch, index):
        self.batch, self.i = batch, int(index)

    def _row_update(self, getter_field, value, stage, setter):
        cur = self.batch.engine.get(getter_field, stage=stage).cpu().numpy()
        cur[self.i] = np.asarray(value, dtype=float).reshape(-1)
        setter(cur)

    def set(self, stage, field, value):
        if field in ("x", "u", "pi"):
            self._row_update(field, value, stage, lambda a: self.batch.engine.set(field, a, stage=stage))
        elif field == "yref":
            self.cost_set(stage, field, value)
        elif field in ("lbx", "ubx", "lbu", "ubu"):
            self.constraints_set(stage, field, value)
        else:
            raise NotImplementedError(f"set(..., {field!r})")
        self.batch._stats = None

    def cost_set(self, stage, field, value):
        if field != "yref":
            raise NotImplementedError("cost_set on one instance of a batch: 'yref' (W is set for the whole batch)")
        self._row_update("yref", value, stage, lambda a: self.batch.engine.set("yref", a, stage=stage))
        self.batch._stats = None

    def constraints_set(self, stage, field, value):
        if field in ("lbx", "ubx") and stage == 0:
            self._row_update("z", value, 0, self.batch.engine.set_state)
        elif field in ("lbx", "ubx", "lbu", "ubu"):
            self._row_update(field, value, stage, lambda a: self.batch.engine.set(field, a, stage=stage))
        else:
            raise NotImplementedError(f"constraints_set(..., {field!r})")
        self.batch._stats = None

    def options_set(self, field, value):
        self.batch.options_set(field, value)

    def solve(self):
        return int(self.batch.solve()[self.i])

    def get(self, stage, field):
        return self.batch.get(stage, field)[self.i]

    def get_stats(self, field):
        if field in ("stat_m", "stat_n"):
            return self.batch.get_stats(field)
        v = self.batch.get_stats(field)[self.i]
        return v if isinstance(v, np.ndarray) and v.ndim > 0 else (v.item() if hasattr(v, "item") else v)

    def get_residuals(self):
        return self.batch.get_residuals()[self.i]

    def get_cost(self):
        return float(self.batch.get_cost()[self.i])

    def get_status(self):
        return int(self.batch.get_status()[self.i])
