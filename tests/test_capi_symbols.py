"""The C-ABI shared library loads on a machine without a GPU, exports every symbol include/rmpc.h declares, and
refuses to work without a CUDA device (no host fallback).  No compute calls here."""
import ctypes as C
import re
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def test_exports_match_header(built):
    import robotic_mpc_b200.engine as eng

    hdr = (ROOT / "include" / "rmpc.h").read_text()
    declared = set(re.findall(r"\b(rmpc_[a-z_]+)\s*\(", hdr))
    declared -= {"rmpc_handle"}
    lib = C.CDLL(str(eng.LIB_PATH))
    for name in sorted(declared):
        assert hasattr(lib, name), f"librmpc.so does not export {name}"
    assert declared == set(eng.EXPORTED_SYMBOLS)


def test_dims_struct_layout(built):
    import robotic_mpc_b200.engine as eng
    from tests import emu

    assert C.sizeof(eng.RmpcDims) == C.sizeof(emu.Dims) == 5 * 4 + 4 + 7 * 8  # 5 ints (+pad) + 7 doubles


def test_create_without_gpu_fails_loudly(built):
    import torch

    import robotic_mpc_b200.engine as eng

    if torch.cuda.is_available():
        return
    lib = eng.load_library()
    d = eng.RmpcDims()
    d.B, d.N, d.solver_type, d.max_iter, d.qp_iter_max = 4, 10, 1, 100, 50
    h = C.c_void_p()
    rc = lib.rmpc_create(C.byref(d), 0, 0, C.byref(h))
    assert rc == -2 and not h  # RMPC_ENODEV
    assert b"no usable CUDA device" in lib.rmpc_last_error(None)
    try:
        eng.Engine(4, 10)
    except eng.RmpcError as e:
        assert "CUDA" in str(e)
    else:
        raise AssertionError("Engine() must raise without a GPU")


def test_bad_arguments(built):
    import robotic_mpc_b200.engine as eng

    lib = eng.load_library()
    h = C.c_void_p()
    d = eng.RmpcDims()
    d.B, d.N, d.solver_type = 0, 10, 1
    assert lib.rmpc_create(C.byref(d), 0, 0, C.byref(h)) == -1
    d.B, d.solver_type = 2, 7
    assert lib.rmpc_create(C.byref(d), 0, 0, C.byref(h)) == -1
    assert lib.rmpc_create(None, 0, 0, C.byref(h)) == -1


def test_field_layouts_match_public_header():
    """robotic-mpc_b200/params.py mirrors the enums of include/rmpc_fields.h (the layouts a C caller fills)."""
    import robotic_mpc_b200.params as pm

    hdr = (ROOT / "include" / "rmpc_fields.h").read_text()
    vals = {m.group(1): int(m.group(2)) for m in re.finditer(r"\b(R[PTLKSEQ]_[A-Z0-9_]+)\s*=\s*(\d+)", hdr)}
    for prefix, table in (("RP", pm.RP), ("RT", pm.RT), ("RL", pm.RL), ("RK", pm.RK), ("RS", pm.RS), ("RE", pm.RE), ("RQ", pm.RQ)):
        for key, v in table.items():
            assert vals[f"{prefix}_{key}"] == v, (prefix, key)
    assert pm.pack_params(3, 0.01).shape == (3, vals["RP_COUNT"])


def test_solve_static_argument_checks_need_no_gpu(built):
    """rmpc_solve_static validates before it touches the device: null buffers and batches beyond the fixed scratch array are
    RMPC_EINVAL with the numbers in the message; a valid request without a CUDA device is RMPC_ENODEV (no host fallback)."""
    import torch

    import robotic_mpc_b200.engine as eng

    lib = eng.load_library()
    d = eng.RmpcDims()
    d.B, d.N, d.solver_type, d.max_iter, d.qp_iter_max = 4, 10, 1, 100, 50
    one = C.c_void_p(8)  # never dereferenced on these paths
    assert lib.rmpc_solve_static(C.byref(d), 0, None, one, one, None, None) == eng.RMPC_EINVAL
    assert lib.rmpc_solve_static(None, 0, one, one, one, None, None) == eng.RMPC_EINVAL
    d.B, d.N = 99, 100  # 98 instances fit at N = 100
    assert lib.rmpc_solve_static(C.byref(d), 0, one, one, one, None, None) == eng.RMPC_EINVAL
    msg = lib.rmpc_last_error(None).decode()
    assert "at most 98 instances" in msg and str(64 << 20) in msg, msg
    d.B, d.N, d.solver_type = 4, 10, 5
    assert lib.rmpc_solve_static(C.byref(d), 0, one, one, one, None, None) == eng.RMPC_EINVAL
    if not torch.cuda.is_available():
        d.solver_type = 1
        assert lib.rmpc_solve_static(C.byref(d), 0, one, one, one, None, None) == eng.RMPC_ENODEV
        assert b"no usable CUDA device" in lib.rmpc_last_error(None)
