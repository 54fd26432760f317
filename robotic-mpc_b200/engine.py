"""ctypes binding of ``librmpc.so`` (C ABI in ``include/rmpc.h``) -- the batched MPC engine on one B200.

PyTorch is used for device memory and streams only: tensors are handed to the library as raw device pointers.
There is NO CPU path: constructing an :class:`Engine` without the compiled library or without a CUDA device raises.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

from .params import RK, RL, RP, RQ, RS

HERE = Path(__file__).resolve().parent
LIB_PATH = Path(os.environ.get("RMPC_LIB", HERE / "librmpc.so"))  # RMPC_LIB: A/B builds of the same library

SQP, SQP_RTI = 0, 1
_SOLVER_TYPES = {"SQP": SQP, "SQP_RTI": SQP_RTI, SQP: SQP, SQP_RTI: SQP_RTI}


class RmpcDims(C.Structure):
    _fields_ = [("B", C.c_int), ("N", C.c_int), ("solver_type", C.c_int), ("max_iter", C.c_int),
                ("qp_iter_max", C.c_int), ("tol", C.c_double * 4), ("qp_tol", C.c_double), ("dt", C.c_double),
                ("plant_dt", C.c_double)]


_lib = None


def load_library() -> C.CDLL:
    """Load librmpc.so; raises if it has not been built (``python -c 'import __graft_entry__ as g; g.build()'``)."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise RuntimeError(f"{LIB_PATH} is missing: build the CUDA extension first (__graft_entry__.build()); "
                           "robotic-mpc_b200 has no CPU fallback")
    lib = C.CDLL(str(LIB_PATH))
    vp, cp, dp, ip = C.c_void_p, C.c_char_p, C.c_void_p, C.c_int
    lib.rmpc_create.argtypes = [C.POINTER(RmpcDims), ip, ip, C.POINTER(vp)]
    lib.rmpc_destroy.argtypes = [vp]
    lib.rmpc_last_error.argtypes = [vp]
    lib.rmpc_last_error.restype = cp
    lib.rmpc_device_bytes.argtypes = [vp]
    lib.rmpc_device_bytes.restype = C.c_longlong
    lib.rmpc_set.argtypes = [vp, cp, ip, dp, vp]
    lib.rmpc_reset.argtypes = [vp, vp]
    lib.rmpc_solve.argtypes = [vp, vp]
    lib.rmpc_solve_rti_phase.argtypes = [vp, ip, vp]
    lib.rmpc_set_chain.argtypes = [vp, dp, dp]
    lib.rmpc_set_option.argtypes = [vp, cp, C.c_double]
    lib.rmpc_state_size.argtypes = [vp]
    lib.rmpc_state_size.restype = C.c_longlong
    lib.rmpc_store_state.argtypes = [vp, dp, vp]
    lib.rmpc_load_state.argtypes = [vp, dp, vp]
    lib.rmpc_closed_loop.argtypes = [vp, ip, vp]
    lib.rmpc_get.argtypes = [vp, cp, ip, dp, vp]
    lib.rmpc_steps_done.argtypes = [vp]
    lib.rmpc_run_host.argtypes = [vp, dp, dp, ip, dp, dp, vp]
    lib.rmpc_step_host.argtypes = [vp, dp, dp, dp, vp]
    lib.rmpc_launch_count.argtypes = [vp]
    lib.rmpc_launch_count.restype = C.c_longlong
    lib.rmpc_solve_static.argtypes = [C.POINTER(RmpcDims), ip, dp, dp, dp, dp, vp]
    _lib = lib
    return lib


EXPORTED_SYMBOLS = ["rmpc_create", "rmpc_destroy", "rmpc_last_error", "rmpc_device_bytes", "rmpc_set", "rmpc_reset",
                    "rmpc_solve", "rmpc_solve_rti_phase", "rmpc_closed_loop", "rmpc_get", "rmpc_steps_done", "rmpc_run_host",
                    "rmpc_launch_count", "rmpc_set_chain", "rmpc_set_option", "rmpc_state_size",
                    "rmpc_store_state", "rmpc_load_state", "rmpc_step_host", "rmpc_solve_static"]


class RmpcError(RuntimeError):
    """Error of the C ABI; ``code`` is the RMPC_E* return value (include/rmpc.h), None for host-side errors."""

    def __init__(self, msg, code=None):
        super().__init__(msg)
        self.code = code


RMPC_EINVAL, RMPC_ENODEV, RMPC_ENOMEM, RMPC_ECUDA = -1, -2, -3, -4


def solve_static(par, x0, N, solver_type="SQP_RTI", dt=0.01, u0=None, stats=None, max_iter=100, qp_iter_max=50, tol=1e-6,
                 qp_tol=1e-8):
    """``rmpc_solve_static``: one cold-started solve of every instance with NO device allocation inside the library -- for
    hosts that must not call cudaMalloc where the solver runs (a hard-deadline control thread; a process whose device memory
    belongs to another pool).  ``par`` [B, RP_COUNT] and ``x0`` [B, 12] are CUDA float64 tensors the caller already holds;
    ``u0`` [B, 6] and ``stats`` [B, RL_COUNT] are filled in place when given (pass them to keep this wrapper allocation-free
    too: without them torch allocates the outputs from its own pool).  Runs on torch's current stream of ``par``'s device and
    returns ``(u0, stats)`` without synchronising.  B is limited by the library's fixed scratch array (98 instances at
    N = 100, 454 at N = 20; ``RmpcError`` with code RMPC_EINVAL beyond); calls share that array and must not overlap on the
    GPU.  The first call on a device does the one-time set-up: make it during start-up."""
    import torch

    lib = load_library()
    for t, n in ((par, RP["COUNT"]), (x0, 12)):
        if not (torch.is_tensor(t) and t.is_cuda and t.dtype == torch.float64 and t.is_contiguous() and t.dim() == 2
                and t.shape[1] == n and t.shape[0] == par.shape[0]):
            raise ValueError(f"solve_static needs contiguous CUDA float64 tensors par [B, {RP['COUNT']}] and x0 [B, 12]")
    B, dev = int(par.shape[0]), par.device
    if u0 is None:
        u0 = torch.empty(B, 6, dtype=torch.float64, device=dev)
    if stats is None:
        stats = torch.empty(B, RL["COUNT"], dtype=torch.float64, device=dev)
    for t, n in ((u0, 6), (stats, RL["COUNT"])):
        if not (t.is_cuda and t.device == dev and t.dtype == torch.float64 and t.is_contiguous() and tuple(t.shape) == (B, n)):
            raise ValueError(f"output buffers must be contiguous CUDA float64 tensors [B, {n}] on {dev}")
    d = RmpcDims()
    d.B, d.N, d.solver_type, d.max_iter, d.qp_iter_max = B, int(N), _SOLVER_TYPES[solver_type], int(max_iter), int(qp_iter_max)
    d.tol[:] = [float(tol)] * 4
    d.qp_tol, d.dt, d.plant_dt = float(qp_tol), float(dt), float(dt)
    rc = lib.rmpc_solve_static(C.byref(d), dev.index or 0, C.c_void_p(par.data_ptr()), C.c_void_p(x0.data_ptr()),
                               C.c_void_p(u0.data_ptr()), C.c_void_p(stats.data_ptr()),
                               C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
    if rc != 0:
        raise RmpcError(f"rmpc_solve_static failed ({rc}): {lib.rmpc_last_error(None).decode()}", rc)
    return u0, stats


class Engine:
    """A batch of B independent UR10 surface-tracking MPC instances resident on one GPU.

    Mirrors, for a batch, what one ``AcadosOcpSolver`` + plant ``Robot`` pair does in ``R/simulator.py:199-241``.
    """

    def __init__(self, B, N, solver_type="SQP_RTI", dt=0.01, nlog=0, device=0, max_iter=100, qp_iter_max=50, tol=1e-6,
                 qp_tol=1e-8, plant_dt=None):
        import torch

        if not torch.cuda.is_available():
            raise RmpcError("robotic-mpc_b200 needs a CUDA device (no CPU fallback)")
        self.torch = torch
        self.lib = load_library()
        self.B, self.N, self.nlog = int(B), int(N), int(nlog)
        self.device = int(device)
        self.dt = float(dt)
        self.solver_type = _SOLVER_TYPES[solver_type]
        d = RmpcDims()
        d.B, d.N, d.solver_type, d.max_iter, d.qp_iter_max = self.B, self.N, self.solver_type, int(max_iter), int(qp_iter_max)
        tols = [tol] * 4 if np.isscalar(tol) else list(tol)
        d.tol[:] = [float(t) for t in tols]
        d.qp_tol, d.dt, d.plant_dt = float(qp_tol), float(dt), float(dt if plant_dt is None else plant_dt)
        self.dims = d
        h = C.c_void_p()
        rc = self.lib.rmpc_create(C.byref(d), self.device, self.nlog, C.byref(h))
        if rc != 0:
            raise RmpcError(f"rmpc_create failed ({rc}): {self.lib.rmpc_last_error(None).decode()}", rc)
        self.h = h
        self.tdev = torch.device("cuda", self.device)

    # ------------------------------------------------------------------ helpers
    def _check(self, rc, what):
        if rc != 0:
            raise RmpcError(f"{what} failed ({rc}): {self.lib.rmpc_last_error(self.h).decode()}", rc)

    def _stream(self):
        return C.c_void_p(self.torch.cuda.current_stream(self.tdev).cuda_stream)

    def _dev(self, a, shape):
        t = self.torch.as_tensor(np.asarray(a) if not self.torch.is_tensor(a) else a, dtype=self.torch.float64)
        t = t.to(self.tdev).contiguous()
        if tuple(t.shape) != tuple(shape):
            raise ValueError(f"expected shape {tuple(shape)}, got {tuple(t.shape)}")
        return t

    def close(self):
        if getattr(self, "h", None) is not None and self.h:
            self.lib.rmpc_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ setters
    def set(self, field, value, stage=None):
        """``rmpc_set``: copy a ``[B, n]`` array into the engine.  Per-instance blocks ("par", "z", the terminal set), the
        iterate ("x", "u", "pi" of ``stage``) and the named cost / constraint fields: "yref" [B, 18], "lbu" / "ubu" [B, 6],
        "lbx" / "ubx" [B, 6] (on q; stage 0: [B, 12] = the initial state) with ``stage = -1`` for every stage or a stage
        index for that stage only; "W_diag" [B, 18], "surf_coeffs" [B, 6], "wcv" [B, 6] per instance (``stage = -1``).
        ``stage=None`` means -1 (every stage) for the named cost / constraint fields and 0 for the iterate fields, so that
        ``set('lbu', v)`` bounds every stage and ``set('lbx', v)`` never lands on the initial-state slot by accident."""
        if stage is None:
            stage = -1 if field in ("yref", "lbu", "ubu", "lbx", "ubx", "W_diag", "surf_coeffs", "wcv") else 0
        n = {"par": RP["COUNT"], "z": 12, "x": 12, "u": 6, "pi": 12, "P_e": 144, "alpha_e": 1, "lh_e": 1, "term": 146,
             "yref": 18, "lbu": 6, "ubu": 6, "lbx": 6, "ubx": 6, "W_diag": 18, "surf_coeffs": 6, "wcv": 6}[field]
        if field in ("lbx", "ubx") and stage == 0:
            n = 12
        t = self._dev(value, (self.B, n))
        self._check(self.lib.rmpc_set(self.h, field.encode(), int(stage), C.c_void_p(t.data_ptr()), self._stream()),
                    f"rmpc_set({field})")
        self.torch.cuda.current_stream(self.tdev).synchronize()  # t may be a temporary

    def set_params(self, par):
        self.set("par", par)

    def set_state(self, z):
        self.set("z", z)

    def set_terminal(self, P, alpha, lh=None):
        """Terminal ellipsoid ``lh <= x_N' P x_N <= alpha`` at stage N (R/_terminal_constraint.py:96-116 gives P, alpha).
        P: [12,12] or [B,12,12]; alpha, lh: scalars or [B]; lh defaults to -alpha (never active, keeps the row two-sided)."""
        P = np.broadcast_to(np.asarray(P, float), (self.B, 12, 12))
        P = 0.5 * (P + np.swapaxes(P, 1, 2))
        alpha = np.broadcast_to(np.asarray(alpha, float), (self.B,))
        lh = -alpha if lh is None else np.broadcast_to(np.asarray(lh, float), (self.B,))
        blk = np.concatenate([P.reshape(self.B, 144), lh[:, None], alpha[:, None]], axis=1)
        self.set("term", blk)

    def store_state(self):
        """Checkpoint of the whole batch (solver memory, plant state, summaries, step counter) as one device tensor; see
        ``load_state``.  ``tensor.cpu().numpy()`` / ``numpy.save`` make it a file (AcadosOcpSolver.store_iterate's role)."""
        t = self.torch.empty(int(self.lib.rmpc_state_size(self.h)), dtype=self.torch.float64, device=self.tdev)
        self._check(self.lib.rmpc_store_state(self.h, C.c_void_p(t.data_ptr()), self._stream()), "rmpc_store_state")
        return t

    def load_state(self, blob):
        """Resume from a ``store_state`` blob of an engine with the same B and N: the continued closed loop is bit-identical
        to an uninterrupted one."""
        t = self.torch.as_tensor(blob, dtype=self.torch.float64).to(self.tdev).contiguous()
        if t.numel() != int(self.lib.rmpc_state_size(self.h)):
            raise ValueError(f"state blob has {t.numel()} doubles, this engine needs {int(self.lib.rmpc_state_size(self.h))}")
        self._check(self.lib.rmpc_load_state(self.h, C.c_void_p(t.data_ptr()), self._stream()), "rmpc_load_state")
        self.torch.cuda.current_stream(self.tdev).synchronize()  # t may be a temporary

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False

    def set_option(self, name, value):
        """Runtime option of the handle: 'shift' (0 | 1) = shift the primal iterate by one stage after every closed-loop step;
        'statistics' (0 | 1) = keep the per-iteration table of the last full-SQP solve (``get('sqp_stat')``); 'graph' (0 | 1) =
        ``step_host`` replays a captured CUDA graph (default) or enqueues its copies and kernel one by one; 'as_rti_level'
        (0..3 = A..D, 4 = standard RTI) and 'as_rti_iter' = acados' advanced-step RTI options (SQP_RTI handles); 'qp_cond_N' =
        qp_solver_cond_N, the horizon of the partially condensed QP (N = none)."""
        self._check(self.lib.rmpc_set_option(self.h, str(name).encode(), float(value)), f"rmpc_set_option({name})")

    def set_chain(self, xyzR=None, axis=None):
        """Kinematic chain of the arm for every instance: a ``robots.chain_table`` / ``robots.chain_from_urdf`` pair, or a
        robot name ('ur10', 'ur5'); None restores the built-in UR10 (R/loader.py:24-36, R/prediction_model.py:22-31)."""
        if isinstance(xyzR, str):
            from .robots import chain_table
            xyzR, axis = chain_table(xyzR)
        if xyzR is None or axis is None:
            self._check(self.lib.rmpc_set_chain(self.h, None, None), "rmpc_set_chain")
            return
        t = np.ascontiguousarray(np.asarray(xyzR, dtype=np.float64).reshape(8, 12))
        a = np.ascontiguousarray(np.asarray(axis, dtype=np.int32).reshape(6))
        self._check(self.lib.rmpc_set_chain(self.h, t.ctypes.data_as(C.c_void_p), a.ctypes.data_as(C.c_void_p)), "rmpc_set_chain")

    def reset(self):
        self._check(self.lib.rmpc_reset(self.h, self._stream()), "rmpc_reset")

    # ------------------------------------------------------------------ compute
    def solve(self, rti_phase=0):
        """One solve of every instance; rti_phase 1 / 2 = preparation / feedback only (SQP_RTI, acados 'rti_phase')."""
        if rti_phase:
            self._check(self.lib.rmpc_solve_rti_phase(self.h, int(rti_phase), self._stream()), "rmpc_solve_rti_phase")
        else:
            self._check(self.lib.rmpc_solve(self.h, self._stream()), "rmpc_solve")

    def closed_loop(self, nsteps):
        self._check(self.lib.rmpc_closed_loop(self.h, int(nsteps), self._stream()), "rmpc_closed_loop")

    def run_host(self, par, z0, nsteps, want_u=True):
        """Host buffers in, host buffers out (the end-to-end path timed by bench.py)."""
        par = np.ascontiguousarray(par, dtype=np.float64)
        z0 = np.ascontiguousarray(z0, dtype=np.float64)
        assert par.shape == (self.B, RP["COUNT"]) and z0.shape == (self.B, 12)
        summ = np.zeros((self.B, RS["COUNT"]))
        u = np.zeros((self.B, 6))
        self._check(self.lib.rmpc_run_host(self.h, par.ctypes.data_as(C.c_void_p), z0.ctypes.data_as(C.c_void_p),
                                           int(nsteps), summ.ctypes.data_as(C.c_void_p),
                                           u.ctypes.data_as(C.c_void_p) if want_u else None, self._stream()),
                    "rmpc_run_host")
        return summ, u

    def step_host(self, x0, u0=None, stats=None, stream=True):
        """One step of an externally driven loop with host buffers: x0 [B, 12] in, u0 [B, 6] (and, with ``stats`` [B, RL_COUNT]
        and nlog > 0, the step record) out -- ``solver.set(0,'lbx',x); solver.set(0,'ubx',x); solver.solve();
        solver.get(0,'u')`` of R/simulator.py:209-217 for the batch in one call.  ``u0`` / ``stats`` are filled in place when
        given (float64, C-contiguous); returns u0.  The step runs on torch's current stream, like every other Engine call, so it
        is ordered after a pending ``set`` / ``reset`` (when that stream is the legacy default stream, which a CUDA graph cannot
        be captured on, the library uses the handle's own blocking stream, which the default stream orders implicitly);
        ``stream=False`` forces the handle's own stream whatever torch's current stream is."""
        x0 = np.ascontiguousarray(x0, dtype=np.float64)
        if x0.shape != (self.B, 12):
            raise ValueError(f"expected shape {(self.B, 12)}, got {x0.shape}")
        if u0 is None:
            u0 = np.empty((self.B, 6))
        for a, shp in ((u0, (self.B, 6)), (stats, (self.B, RL["COUNT"]))):
            if a is not None and not (a.dtype == np.float64 and a.flags.c_contiguous and a.shape == shp):
                raise ValueError(f"output buffer must be a C-contiguous float64 array of shape {shp}")
        self._check(self.lib.rmpc_step_host(self.h, x0.ctypes.data_as(C.c_void_p), u0.ctypes.data_as(C.c_void_p),
                                            stats.ctypes.data_as(C.c_void_p) if stats is not None else None,
                                            self._stream() if stream else None), "rmpc_step_host")
        return u0

    # ------------------------------------------------------------------ getters
    def get(self, field, stage=0):
        B, nl = self.B, self.nlog
        shape = {"x": (B, 12), "u": (B, 6), "pi": (B, 12), "lam": (B, 24), "lamx0": (B, 24), "z": (B, 12),
                 "stats": (B, RL["COUNT"]), "summary": (B, RS["COUNT"]), "flags": (B, 4),
                 "zlog": (B, nl + 1, 12), "ulog": (B, nl, 6), "slog": (B, nl, RL["COUNT"]),
                 "plog": (B, nl + 1, RK["COUNT"]), "par": (B, RP["COUNT"]), "yref": (B, 18), "lbu": (B, 6), "ubu": (B, 6),
                 "lbx": (B, 6), "ubx": (B, 6), "sqp_stat": (B, self.dims.max_iter + 2, RQ["COUNT"])}[field]
        out = self.torch.empty(shape, dtype=self.torch.float64, device=self.tdev)
        self._check(self.lib.rmpc_get(self.h, field.encode(), int(stage), C.c_void_p(out.data_ptr()), self._stream()),
                    f"rmpc_get({field})")
        return out

    @property
    def steps_done(self):
        return self.lib.rmpc_steps_done(self.h)

    @property
    def launch_count(self):
        return self.lib.rmpc_launch_count(self.h)

    @property
    def device_bytes(self):
        return self.lib.rmpc_device_bytes(self.h)
