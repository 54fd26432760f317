"""rmpc_solve_static -- the one entry point of librmpc.so that allocates nothing on the device.

It is there for two kinds of host that cannot call cudaMalloc: a controller thread with a hard per-cycle deadline (allocation
in the CUDA driver has no bounded run time and can wait for the whole GPU), and a plug-in inside a process whose device
memory already belongs to another allocator's pool.  For them the solver's work set sits in a fixed array inside the library
(RMPC_STATIC_SCRATCH_BYTES, include/rmpc.h) instead of in memory rmpc_create would allocate.

The checker here is the library's own handle path (create, set "par", set "z", reset, solve, get "u" / "stats"), which the rest
of the suite ties to the reference: both run the same device code on the same inputs, so the controls, iteration counts,
costs and residuals must agree to rounding (1e-12 relative, far inside the suite's 1e-6).  Two calls in a row with different
inputs show that nothing of the first call survives in the shared array.  Everything goes through the C ABI with ctypes;
torch only provides the caller's input and output buffers.

The file sorts last on purpose: pytest -x stops at the first failure, and a failure here should not hide the parity tests.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
RTOL = 1e-12


def _inputs(o, B, seed):
    rng = np.random.default_rng(seed)
    surf = np.array([-0.1, 0.1, -0.01, 0.01, 0.01, 0.0]) + 0.01 * rng.standard_normal((B, 6))
    z0 = np.concatenate([o.README_Q0 + rng.uniform(-0.05, 0.05, (B, 6)), o.README_QD0 + rng.uniform(-0.1, 0.1, (B, 6))], 1)
    return surf, z0


def _dims(eng, B, N, solver_type):
    d = eng.RmpcDims()
    d.B, d.N, d.solver_type, d.max_iter, d.qp_iter_max = B, N, solver_type, 100, 50
    d.tol[:] = [1e-6] * 4
    d.qp_tol, d.dt, d.plant_dt = 1e-8, 0.01, 0.01
    return d


def _handle_path(lib, eng, torch, d, par, x0):
    """The existing entry points: a handle with its own (allocated) work set, one cold-started solve."""
    B = d.B
    h = C.c_void_p()
    assert lib.rmpc_create(C.byref(d), 0, 1, C.byref(h)) == 0, lib.rmpc_last_error(None)
    u = torch.empty(B, 6, dtype=torch.float64, device="cuda")
    st = torch.empty(B, eng.RL["COUNT"], dtype=torch.float64, device="cuda")
    try:
        for rc in (lib.rmpc_set(h, b"par", 0, par.data_ptr(), None), lib.rmpc_set(h, b"z", 0, x0.data_ptr(), None),
                   lib.rmpc_reset(h, None), lib.rmpc_solve(h, None), lib.rmpc_get(h, b"u", 0, u.data_ptr(), None),
                   lib.rmpc_get(h, b"stats", 0, st.data_ptr(), None)):
            assert rc == 0, lib.rmpc_last_error(h)
        torch.cuda.synchronize()
    finally:
        lib.rmpc_destroy(h)
    return u.cpu().numpy(), st.cpu().numpy()


@pytest.mark.parametrize("solver", ["SQP_RTI", "SQP"])
def test_solve_static_matches_handle_path_twice_in_a_row(built, oracle, solver):
    import torch

    import robotic_mpc_b200 as rm
    import robotic_mpc_b200.engine as eng

    lib = eng.load_library()
    RL = eng.RL
    st_type = eng.SQP_RTI if solver == "SQP_RTI" else eng.SQP
    # second call: fewer instances, another horizon, other surfaces and states -- it lands on memory the first call filled
    for B, N, seed in ((24, 20, 101), (9, 12, 202)):
        surf, z0 = _inputs(oracle, B, seed)
        par = torch.as_tensor(rm.pack_params(B, 0.01, surf=surf), dtype=torch.float64).cuda().contiguous()
        x0 = torch.as_tensor(z0, dtype=torch.float64).cuda().contiguous()
        d = _dims(eng, B, N, st_type)
        u_ref, st_ref = _handle_path(lib, eng, torch, d, par, x0)
        u = torch.full((B, 6), float("nan"), dtype=torch.float64, device="cuda")
        st = torch.full((B, RL["COUNT"]), float("nan"), dtype=torch.float64, device="cuda")
        x0_before = x0.clone()
        torch.cuda.synchronize()
        rc = lib.rmpc_solve_static(C.byref(d), 0, par.data_ptr(), x0.data_ptr(), u.data_ptr(), st.data_ptr(), None)
        assert rc == 0, lib.rmpc_last_error(None)
        torch.cuda.synchronize()
        u, st = u.cpu().numpy(), st.cpu().numpy()
        assert torch.equal(x0, x0_before)  # inputs are read only
        assert np.isfinite(u).all() and np.abs(u_ref).max() > 1e-3
        err = np.abs(u - u_ref).max() / np.abs(u_ref).max()
        assert err <= RTOL, (solver, B, N, err)
        for key in ("STATUS", "SQP_ITER", "QP_ITER", "QP_STATUS"):
            assert np.array_equal(st[:, RL[key]], st_ref[:, RL[key]]), key
        assert np.allclose(st[:, RL["COST"]], st_ref[:, RL["COST"]], rtol=RTOL, atol=0)
        assert np.allclose(st[:, RL["RES"]:RL["RES"] + 4], st_ref[:, RL["RES"]:RL["RES"] + 4], rtol=1e-9, atol=1e-14)
        if solver == "SQP":
            assert (st[:, RL["SQP_ITER"]] > 1).all()  # a real multi-iteration solve, not an early exit


def test_solve_static_without_step_record_and_over_the_limit(built, oracle):
    import torch

    import robotic_mpc_b200 as rm
    import robotic_mpc_b200.engine as eng

    lib = eng.load_library()
    B, N = 5, 10
    surf, z0 = _inputs(oracle, B, 303)
    par = torch.as_tensor(rm.pack_params(B, 0.01, surf=surf), dtype=torch.float64).cuda().contiguous()
    x0 = torch.as_tensor(z0, dtype=torch.float64).cuda().contiguous()
    d = _dims(eng, B, N, eng.SQP_RTI)
    u_ref, _ = _handle_path(lib, eng, torch, d, par, x0)
    u = torch.zeros(B, 6, dtype=torch.float64, device="cuda")
    assert lib.rmpc_solve_static(C.byref(d), 0, par.data_ptr(), x0.data_ptr(), u.data_ptr(), None, None) == 0
    torch.cuda.synchronize()
    assert np.abs(u.cpu().numpy() - u_ref).max() <= RTOL * np.abs(u_ref).max()
    # 99 instances at N = 100 are one more than the array holds: refused with the numbers, nothing launched or allocated
    free_before = torch.cuda.mem_get_info()[0]
    big = _dims(eng, 99, 100, eng.SQP_RTI)
    assert lib.rmpc_solve_static(C.byref(big), 0, par.data_ptr(), x0.data_ptr(), u.data_ptr(), None, None) == eng.RMPC_EINVAL
    msg = lib.rmpc_last_error(None).decode()
    assert "at most 98 instances" in msg and "static array" in msg, msg
    assert lib.rmpc_solve_static(C.byref(d), 0, None, x0.data_ptr(), u.data_ptr(), None, None) == eng.RMPC_EINVAL
    assert lib.rmpc_solve_static(C.byref(d), 0, par.data_ptr(), x0.data_ptr(), u.data_ptr(), None, None) == 0
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free_before  # the calls took no device memory


def test_python_wrapper_solve_static(built, oracle):
    """robotic_mpc_b200.solve_static, the Python face of the same entry point: caller-owned output tensors, torch's current
    stream, the same numbers as an Engine driven the usual way; the size limit surfaces as RmpcError."""
    import torch

    import robotic_mpc_b200 as rm

    B, N = 7, 16
    surf, z0 = _inputs(oracle, B, 404)
    par = torch.as_tensor(rm.pack_params(B, 0.01, surf=surf), dtype=torch.float64).cuda().contiguous()
    x0 = torch.as_tensor(z0, dtype=torch.float64).cuda().contiguous()
    u0 = torch.empty(B, 6, dtype=torch.float64, device="cuda")
    st = torch.empty(B, rm.RL["COUNT"], dtype=torch.float64, device="cuda")
    e = rm.Engine(B, N, "SQP", dt=0.01, nlog=1)
    e.set_params(par)
    e.set_state(x0)
    e.reset()
    e.solve()
    want_u, want_st = e.get("u", 0).cpu().numpy(), e.get("stats").cpu().numpy()
    e.close()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        got_u, got_st = rm.solve_static(par, x0, N, "SQP", u0=u0, stats=st)
    side.synchronize()
    assert got_u is u0 and got_st is st
    assert np.abs(u0.cpu().numpy() - want_u).max() <= RTOL * np.abs(want_u).max()
    assert np.array_equal(st.cpu().numpy()[:, :3], want_st[:, :3])  # status, SQP and QP iteration counts
    with pytest.raises(rm.RmpcError) as ei:
        rm.solve_static(par.repeat(100, 1), x0.repeat(100, 1), 100)  # 700 instances at N = 100
    assert ei.value.code == -1
    with pytest.raises(ValueError):
        rm.solve_static(par.cpu(), x0, N)
