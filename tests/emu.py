"""Python binding of tests/simt_emu/librmpc_emu.so -- the device code of robotic-mpc_b200/csrc/rmpc_device.cuh compiled
for the host warp emulator.  TEST HARNESS ONLY (see tests/simt_emu/simt_emu.h)."""
import ctypes as C
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
SO = HERE / "simt_emu" / "librmpc_emu.so"


class Dims(C.Structure):
    _fields_ = [("B", C.c_int), ("N", C.c_int), ("solver_type", C.c_int), ("max_iter", C.c_int), ("qp_iter_max", C.c_int),
                ("tol", C.c_double * 4), ("qp_tol", C.c_double), ("dt", C.c_double), ("plant_dt", C.c_double)]


def closed_loop(par, z0, N, solver_type, nsteps, dt=0.01, seed=1, max_iter=100, qp_iter_max=50, tol=1e-6, qp_tol=1e-8,
                term=None):
    """term: optional [B][146] terminal-set blocks [P(144), lh, uh] (RE_* in csrc/rmpc_types.h)."""
    lib = C.CDLL(str(SO))
    B = par.shape[0]
    D = Dims(B, N, solver_type, max_iter, qp_iter_max, (C.c_double * 4)(tol, tol, tol, tol), qp_tol, dt, dt)
    par = np.ascontiguousarray(par, np.float64)
    z0 = np.ascontiguousarray(z0, np.float64)
    out = dict(z=np.zeros((B, nsteps + 1, 12)), u=np.zeros((B, nsteps, 6)), slog=np.zeros((B, nsteps, 20)),
               summary=np.zeros((B, 20)), flags=np.zeros((B, 4), np.int32))
    dp = C.POINTER(C.c_double)
    pe = None
    if term is not None:
        term = np.ascontiguousarray(term, np.float64)
        assert term.shape == (B, 146)
        pe = term.ctypes.data_as(dp)
    lib.rmpc_emu_closed_loop_term(C.byref(D), par.ctypes.data_as(dp), pe, z0.ctypes.data_as(dp), int(nsteps),
                             out["z"].ctypes.data_as(dp), out["u"].ctypes.data_as(dp), out["slog"].ctypes.data_as(dp),
                             out["summary"].ctypes.data_as(dp), out["flags"].ctypes.data_as(C.POINTER(C.c_int)),
                             C.c_ulonglong(seed))
    return out


def set_chain(xyzR=None, axis=None):
    """Process-global chain of the emulated device code (same table as rmpc_set_chain); None restores the UR10."""
    lib = C.CDLL(str(SO))
    if xyzR is None:
        lib.rmpc_emu_set_chain(None, None)
        return
    t = np.ascontiguousarray(np.asarray(xyzR, np.float64).reshape(8, 12))
    a = np.ascontiguousarray(axis, np.int32)
    assert lib.rmpc_emu_set_chain(t.ctypes.data_as(C.POINTER(C.c_double)), a.ctypes.data_as(C.POINTER(C.c_int))) == 0


def plant_record(par, z):
    """(rec [28] = RK_* record, y5 [5] task outputs) of one state through the emulated device FK."""
    lib = C.CDLL(str(SO))
    par, z = np.ascontiguousarray(par, np.float64), np.ascontiguousarray(z, np.float64)
    rec, y5 = np.zeros(28), np.zeros(5)
    dp = C.POINTER(C.c_double)
    lib.rmpc_emu_plant_record(par.ctypes.data_as(dp), z.ctypes.data_as(dp), rec.ctypes.data_as(dp), y5.ctypes.data_as(dp))
    return rec, y5


def set_shift(on):
    """Process-global 'shift' option of the emulated closed loop (rmpc_set_option(h, "shift", .))."""
    C.CDLL(str(SO)).rmpc_emu_set_shift(int(bool(on)))


def set_as_rti(level, iters=1):
    """Process-global AS-RTI level of the emulated closed loop (None / -1: standard RTI, 0..3 = A..D) and its iterations."""
    C.CDLL(str(SO)).rmpc_emu_set_as_rti(-1 if level is None else int(level), int(iters))


def set_cond_N(n2):
    """Process-global qp_solver_cond_N of the emulated solver (None / 0: no condensing)."""
    C.CDLL(str(SO)).rmpc_emu_set_cond_N(0 if n2 is None else int(n2))
