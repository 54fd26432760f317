"""ctypes bindings for the two oracle legs (TEST INFRASTRUCTURE ONLY).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import this module, and only as the checker / reported CPU baseline.  The product package
(``robotic-mpc_b200``) never imports anything under ``oracle/``.

Legs
----
``ref``   oracle/_ref/librmpc_ref.so : the UNMODIFIED acados + HPIPM + BLASFEO of the reference,
          compiled by oracle/ref.mk, driven by oracle/ref_driver.c (needs /root/reference to BUILD,
          not to run -- the built .so travels to the GPU box).
``port``  oracle/_build/librmpc_oracle.so : plain-C restatement of the same algorithm
          (oracle/rmpc_oracle.c), buildable anywhere with gcc.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
REF_SO = HERE / "_ref" / "librmpc_ref.so"
PORT_SO = HERE / "_build" / "librmpc_oracle.so"

c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int)


class Ur10Params(C.Structure):
    _fields_ = [("surf", C.c_double * 6), ("wcv", C.c_double * 6), ("Ts", C.c_double),
                ("t_ee", C.c_double * 3)]


class RmpcProblem(C.Structure):
    _fields_ = [("N", C.c_int), ("solver_type", C.c_int), ("max_iter", C.c_int),
                ("qp_iter_max", C.c_int), ("tol", C.c_double * 4), ("qp_tol", C.c_double),
                ("dt", C.c_double), ("w_task", C.c_double * 5), ("w_u", C.c_double),
                ("w_qddot", C.c_double), ("px_ref", C.c_double), ("vy_ref", C.c_double),
                ("q_min", C.c_double * 6), ("q_max", C.c_double * 6), ("u_min", C.c_double * 6),
                ("u_max", C.c_double * 6), ("model", Ur10Params),
                ("term", C.c_int), ("term_P", C.c_double * 144), ("term_lh", C.c_double), ("term_uh", C.c_double),
                ("shift", C.c_int), ("cond_N", C.c_int), ("as_rti_level_p1", C.c_int), ("as_rti_iter", C.c_int)]


class RmpcTrace(C.Structure):
    _fields_ = [("z", c_dp), ("u", c_dp), ("status", c_ip), ("sqp_iter", c_ip), ("qp_iter", c_ip),
                ("res", c_dp), ("cost", c_dp), ("time_tot", c_dp), ("xtraj", c_dp), ("utraj", c_dp),
                ("pitraj", c_dp)]


SQP, SQP_RTI = 0, 1
README_Q0 = np.array([np.pi / 4, -np.pi / 3, np.pi / 4, -np.pi / 2, -np.pi / 2, 0.0])
README_QD0 = np.array([1.0, 2.0, 1.0, 0.0, 0.0, 0.0])


def make_problem(N=20, solver_type=SQP_RTI, dt=0.01, surf=None, wcv=200.0, w_u=0.01, w_qddot=0.02,
                 px_ref=0.40, vy_ref=0.05, q_min=None, q_max=None, qdot_min=None, qdot_max=None,
                 max_iter=100, qp_iter_max=50, tol=1e-6, qp_tol=1e-8, term_P=None, term_alpha=None,
                 term_lh=None, shift=False, cond_N=None, as_rti_level=None, as_rti_iter=1) -> RmpcProblem:
    """Problem with the reference defaults unless overridden."""
    p = RmpcProblem()
    p.N, p.solver_type, p.max_iter, p.qp_iter_max = int(N), int(solver_type), int(max_iter), int(qp_iter_max)
    p.tol[:] = [tol] * 4
    p.qp_tol = qp_tol
    p.dt = dt
    p.w_task[:] = [50.0] * 5
    p.w_u, p.w_qddot, p.px_ref, p.vy_ref = w_u, w_qddot, px_ref, vy_ref
    qdm = np.array([2.16, 2.16, np.pi, 3.20, 3.20, 3.20])
    p.q_min[:] = list(np.full(6, -2 * np.pi) if q_min is None else np.asarray(q_min, float))
    p.q_max[:] = list(np.full(6, 2 * np.pi) if q_max is None else np.asarray(q_max, float))
    p.u_min[:] = list(-qdm if qdot_min is None else np.asarray(qdot_min, float))
    p.u_max[:] = list(qdm if qdot_max is None else np.asarray(qdot_max, float))
    s = [-0.15, 0.15, -0.01, 0.01, 0.01, 0.0] if surf is None else list(np.asarray(surf, float))
    p.model.surf[:] = s
    p.model.wcv[:] = list(np.broadcast_to(np.asarray(wcv, float), (6,)))
    p.model.Ts = dt
    p.model.t_ee[:] = [0.0, 0.0, 0.1]
    p.shift = int(bool(shift))
    # reference leg only: partial condensing and AS-RTI level (acados enum: 0 = A .. 3 = D, 4 = standard RTI)
    p.cond_N = 0 if cond_N is None else int(cond_N)
    p.as_rti_level_p1 = 0 if as_rti_level is None else int(as_rti_level) + 1
    p.as_rti_iter = int(as_rti_iter)
    if term_P is not None:
        # terminal ellipsoid x_N' P x_N <= alpha (R/_terminal_constraint.py:96-116); lower side -alpha unless given
        P = np.asarray(term_P, float).reshape(12, 12)
        p.term = 1
        p.term_P[:] = list(np.ascontiguousarray(0.5 * (P + P.T)).ravel())
        p.term_uh = float(term_alpha)
        p.term_lh = float(-term_alpha if term_lh is None else term_lh)
    return p


def build(which=("port", "ref"), quiet=True) -> dict:
    """Compile the oracle legs. The ref leg needs /root/reference; absent => uses the prebuilt .so."""
    out = {}
    kw = dict(stdout=subprocess.DEVNULL) if quiet else {}
    if "port" in which:
        subprocess.run(["make", "-s", "-C", str(HERE), "port"], check=True, **kw)
        out["port"] = PORT_SO.exists()
    if "ref" in which:
        ref_root = Path(os.environ.get("RMPC_REFERENCE", "/root/reference"))
        if (ref_root / "acados" / "acados").is_dir():
            subprocess.run(["make", "-s", "-C", str(HERE), "-f", "ref.mk", f"REF={ref_root}",
                            f"-j{os.cpu_count() or 4}", "ref-driver"], check=True, **kw)
        out["ref"] = REF_SO.exists()
    return out


def _np(a, dtype=np.float64):
    return np.ascontiguousarray(a, dtype=dtype)


def _dp(a):
    return a.ctypes.data_as(c_dp)


def _ip(a):
    return a.ctypes.data_as(c_ip)


class _Leg:
    prefix = ""
    so = None

    def __init__(self):
        if not self.so.exists():
            raise FileNotFoundError(f"{self.so} not built (python -c 'import oracle.pyoracle as o; o.build()')")
        self.lib = C.CDLL(str(self.so))
        f = getattr(self.lib, f"{self.prefix}_closed_loop")
        f.restype = C.c_int
        f.argtypes = [C.POINTER(RmpcProblem), c_dp, C.c_int, C.POINTER(RmpcTrace)]
        self._closed_loop = f

    def set_chain(self, xyzR=None, axis=None):
        """Kinematic chain used by this leg's model functions (process-global inside the .so); None restores the UR10.
        Same table as rmpc_set_chain (include/rmpc.h)."""
        if xyzR is None:
            self.lib.ur10_model_set_chain(None, None)
            return
        t, a = _np(np.asarray(xyzR).reshape(8, 12)), _np(axis, np.int32)
        self.lib.ur10_model_set_chain(_dp(t), _ip(a))

    def closed_loop(self, prob: RmpcProblem, z0, nsteps: int, full_traj=False) -> dict:
        N = prob.N
        z0 = _np(z0)
        o = dict(z=np.zeros((nsteps + 1, 12)), u=np.zeros((nsteps, 6)),
                 status=np.zeros(nsteps, np.int32), sqp_iter=np.zeros(nsteps, np.int32),
                 qp_iter=np.zeros(nsteps, np.int32), res=np.zeros((nsteps, 4)), cost=np.zeros(nsteps),
                 time_tot=np.zeros(nsteps))
        if full_traj:
            o.update(xtraj=np.zeros((nsteps, N + 1, 12)), utraj=np.zeros((nsteps, N, 6)),
                     pitraj=np.zeros((nsteps, N, 12)))
        tr = RmpcTrace()
        for k, v in o.items():
            setattr(tr, k, _ip(v) if v.dtype == np.int32 else _dp(v))
        nfail = self._closed_loop(C.byref(prob), _dp(z0), int(nsteps), C.byref(tr))
        if nfail < 0:
            raise RuntimeError(f"{self.prefix}_closed_loop failed to create the solver")
        o["nfail"] = nfail
        return o


class RefLeg(_Leg):
    """The reference's own compiled acados/HPIPM/BLASFEO."""
    prefix = "rmpc_ref"
    so = REF_SO


class PortLeg(_Leg):
    """Plain-C restatement."""
    prefix = "rmpc_oracle"
    so = PORT_SO


def checker_leg() -> _Leg:
    """The leg parity tests compare the device with: the reference's own compiled code wherever oracle/_ref was built (it
    travels to the GPU box with the repo), the plain-C port only where it was not."""
    return RefLeg() if REF_SO.exists() else PortLeg()


# ------------------------------------------------------------------ model functions (libur10 inside either .so)
class Model:
    def __init__(self, lib: C.CDLL):
        self.lib = lib

    def fk(self, q):
        q = _np(q)
        pj, aj, R, pe = np.zeros((6, 3)), np.zeros((6, 3)), np.zeros(9), np.zeros(3)
        self.lib.ur10_fk(_dp(q), _dp(pj), _dp(aj), _dp(R), _dp(pe))
        return pj, aj, R.reshape(3, 3), pe

    def jacobian_world(self, q):
        J = np.zeros(36)
        self.lib.ur10_jacobian_world(_dp(_np(q)), _dp(J))
        return J.reshape(6, 6)

    def cost_y(self, P: Ur10Params, x, u):
        y = np.zeros(18)
        self.lib.ur10_cost_y(C.byref(P), _dp(_np(x)), _dp(_np(u)), _dp(y))
        return y

    def cost_y_jac(self, P: Ur10Params, x, u):
        y, Jt = np.zeros(18), np.zeros(324)
        self.lib.ur10_cost_y_jac(C.byref(P), _dp(_np(x)), _dp(_np(u)), _dp(y), _dp(Jt))
        return y, Jt.reshape(18, 18).T.copy()  # [row v, col y]

    def dyn(self, P: Ur10Params, x, u):
        xn = np.zeros(12)
        self.lib.ur10_dyn(C.byref(P), _dp(_np(x)), _dp(_np(u)), _dp(xn))
        return xn

    def BAbt(self, P: Ur10Params):
        M = np.zeros(18 * 12)
        self.lib.ur10_dyn_jac_BAbt(C.byref(P), _dp(M))
        return M.reshape(12, 18).T.copy()  # (nu+nx) x nx

    def plant_rk4(self, wcv, dt, z, u):
        zn = np.zeros(12)
        self.lib.ur10_plant_rk4(_dp(_np(wcv)), C.c_double(dt), _dp(_np(z)), _dp(_np(u)), _dp(zn))
        return zn

    def plant_logs(self, z):
        pose, rpy, tw = np.zeros(12), np.zeros(3), np.zeros(6)
        self.lib.ur10_plant_logs(_dp(_np(z)), _dp(pose), _dp(rpy), _dp(tw))
        return pose, rpy, tw

    def tracking_errors(self, P: Ur10Params, pose, twist, px_ref, vy_ref):
        e, ptz = np.zeros(5), C.c_double(0)
        self.lib.ur10_tracking_errors(C.byref(P), _dp(_np(pose)), _dp(_np(twist)), C.c_double(px_ref),
                                      C.c_double(vy_ref), _dp(e), C.byref(ptz))
        return e, ptz.value


class RefSolver:
    """Step-wise handle on the compiled reference (oracle/_ref): create -> [model_set ...] -> step ... -> close, the way
    R/simulator.py:203-226 drives one AcadosOcpSolver.  TEST INFRASTRUCTURE."""

    def __init__(self, prob: RmpcProblem, x0):
        self.lib = RefLeg().lib
        self.lib.rmpc_ref_create.restype = C.c_void_p
        self.lib.rmpc_ref_destroy.argtypes = [C.c_void_p]
        self.lib.rmpc_ref_step.argtypes = [C.c_void_p, c_dp, c_dp, c_ip, c_ip, c_dp, c_dp, c_dp]
        self.lib.rmpc_ref_model_set.argtypes = [C.c_void_p, C.c_int, C.c_char_p, c_dp]
        self.lib.rmpc_ref_statistics.argtypes = [C.c_void_p, c_dp, C.c_int, c_ip]
        self.lib.rmpc_ref_times.argtypes = [C.c_void_p, c_dp]
        self.lib.rmpc_ref_get_iterate.argtypes = [C.c_void_p, c_dp, c_dp, c_dp]
        self.N = prob.N
        self.prob = prob
        x0 = _np(x0)
        self.h = C.c_void_p(self.lib.rmpc_ref_create(C.byref(prob), _dp(x0)))
        if not self.h:
            raise RuntimeError("rmpc_ref_create failed")

    def model_set(self, stage, field, value):
        v = _np(value)
        if field == "W":
            v = _np(np.asarray(value, float).reshape(18, 18).T)  # column-major
        rc = self.lib.rmpc_ref_model_set(self.h, int(stage), field.encode(), _dp(v))
        if rc != 0:
            raise RuntimeError(f"rmpc_ref_model_set({stage}, {field}) -> {rc}")

    def step(self, x):
        x, u = _np(x), np.zeros(6)
        sqp, qp = C.c_int(0), C.c_int(0)
        res, cost, tt = np.zeros(4), C.c_double(0), C.c_double(0)
        st = self.lib.rmpc_ref_step(self.h, _dp(x), _dp(u), C.byref(sqp), C.byref(qp), _dp(res), C.byref(cost), C.byref(tt))
        return dict(u=u, status=int(st), sqp_iter=sqp.value, qp_iter=qp.value, res=res, cost=cost.value, time_tot=tt.value)

    def prepare(self):
        """options_set('rti_phase', 1); solve(): the preparation phase (advanced-step preparation when the problem carries an
        as_rti_level), stage-0 bounds untouched."""
        self.lib.rmpc_ref_prepare.argtypes = [C.c_void_p]
        return int(self.lib.rmpc_ref_prepare(self.h))

    def feedback(self, x):
        """options_set('rti_phase', 2); set(0, 'lbx' / 'ubx', x); solve(); get(0, 'u')."""
        self.lib.rmpc_ref_set_rti_phase.argtypes = [C.c_void_p, C.c_int]
        self.lib.rmpc_ref_set_rti_phase(self.h, 2)
        r = self.step(x)
        self.lib.rmpc_ref_set_rti_phase(self.h, 0)
        return r

    def statistics(self):
        """AcadosOcpSolver.get_stats('statistics') of the last solve: [(stat_n + 1), rows]."""
        buf, sn = np.zeros(16 * 256), C.c_int(0)
        rows = self.lib.rmpc_ref_statistics(self.h, _dp(buf), buf.size, C.byref(sn))
        if rows < 0:
            raise RuntimeError("statistics buffer too small")
        return buf[:(sn.value + 1) * rows].reshape(sn.value + 1, rows).copy()

    def times(self):
        t = np.zeros(6)
        self.lib.rmpc_ref_times(self.h, _dp(t))
        return dict(zip(("time_tot", "time_lin", "time_qp", "time_qp_solver_call", "time_qp_xcond", "time_glob"), t))

    def iterate(self):
        x, u, pi = np.zeros((self.N + 1, 12)), np.zeros((self.N, 6)), np.zeros((self.N, 12))
        self.lib.rmpc_ref_get_iterate(self.h, _dp(x), _dp(u), _dp(pi))
        return x, u, pi

    def close(self):
        if self.h:
            self.lib.rmpc_ref_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
