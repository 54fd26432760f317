"""GPU parity at the STATED shapes of the BASELINE.json configs that tests/test_gpu_parity.py only covers at reduced sizes
("What's weak" #1): configs[3] proper (N = 200, full SQP, terminal ellipsoid, inactive AND active), the
README grid tuples at N = 40 incl. the (w_u = 0.001, SQP_RTI) corner, and interior-point iteration counts asserted EQUAL
(not +-1) with an explicit, counted exception list.  Everything goes through the C ABI; the checker is the compiled
reference (oracle/_ref) where it travelled, else the port.  Tolerance 1e-6 relative (BASELINE.json north_star)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
RTOL = 1e-6


def relerr(a, b):
    return float(np.abs(np.asarray(a) - np.asarray(b)).max() / max(np.abs(b).max(), 1e-300))


def _inputs(o, B, seed):
    rng = np.random.default_rng(seed)
    surf = np.array([-0.1, 0.1, -0.01, 0.01, 0.01, 0.0]) + 0.01 * rng.standard_normal((B, 6))
    z0 = np.concatenate([o.README_Q0 + rng.uniform(-0.05, 0.05, (B, 6)), o.README_QD0 + rng.uniform(-0.1, 0.1, (B, 6))], 1)
    return surf, z0


def _leg(o):
    return o.RefLeg() if o.REF_SO.exists() else o.PortLeg()


def _pool_closed_loops(o, jobs):
    """Reference closed loops of `jobs` = [(problem kwargs, z0, nsteps)] on the host cores (fork pool; each worker loads the
    checker library itself)."""
    import multiprocessing as mp
    import os

    n = min(len(jobs), len(os.sched_getaffinity(0)))
    with mp.get_context("fork").Pool(n) as pool:
        return pool.map(_one_closed_loop, jobs)


def _one_closed_loop(job):
    from oracle import pyoracle as o

    kw, z0, ns = job
    g = _leg(o).closed_loop(o.make_problem(**kw), z0, ns, full_traj=True)
    return {k: g[k] for k in ("u", "z", "status", "sqp_iter", "qp_iter", "cost", "res", "xtraj", "utraj")}


def test_config4_shape_n200_full_sqp_terminal_ellipsoid(oracle):
    """BASELINE configs[3]: N = 200, full SQP (<= 100 iterations, tol 1e-6), terminal ellipsoid x_N' P x_N <= alpha with
    (P, alpha) = terminal_ellipsoid_from_lqr, 64 instances x 3 solves.  Half of the batch keeps the LQR alpha (row inactive),
    half gets alpha = s x0' P x0, s in 0.5..0.95 (row ACTIVE at the solution).

    Asserted for every instance: applied controls and plant states <= 1e-6 relative.  For the inactive half the status, SQP
    and QP iteration sequences are asserted EQUAL.  For the active half they are not: Gauss-Newton SQP on an active
    quadratic row converges linearly and creeps along the 1e-6 tolerance for 40-100 iterations, so the iteration at which the
    test passes -- or whether max_iter = 100 is hit first (status 2 instead of 0) -- is decided by rounding noise: the
    reference's own port (which restates HPIPM's LQ route line by line) differs from the compiled reference by the same
    amount as the device does (profiles/r2a_config4_sqp_counts.txt: e.g. ref 89 / port 100 / device 87 iterations).  There
    the counts must stay within 25 of each other, a status may differ only as 0 vs 2 with the larger count at max_iter,
    and the number of differing instances is printed."""
    import robotic_mpc_b200 as rm
    import robotic_mpc_b200.terminal_constraint as tc

    o = oracle
    B, N, ns = 64, 200, 3
    surf, z0 = _inputs(o, B, 400 + N)
    P, alpha0 = tc.default_terminal_set(rho=100.0)
    scale = [None if i % 2 == 0 else (0.5, 0.6, 0.7, 0.8, 0.9, 0.95)[(i // 2) % 6] for i in range(B)]
    alpha = np.array([alpha0 if s is None else s * float(z @ P @ z) for s, z in zip(scale, z0)])
    eng = rm.Engine(B, N, "SQP", dt=0.01, nlog=ns)
    eng.set_params(rm.pack_params(B, 0.01, surf=surf))
    eng.set_terminal(P, alpha, -alpha)
    eng.set_state(z0)
    eng.reset()
    eng.closed_loop(ns)
    ulog, zlog, slog = (eng.get(k).cpu().numpy() for k in ("ulog", "zlog", "slog"))
    xN = eng.get("x", N).cpu().numpy()
    eng.close()
    ref = _pool_closed_loops(o, [(dict(N=N, solver_type=o.SQP, surf=surf[i], term_P=P, term_alpha=alpha[i]), z0[i], ns)
                                 for i in range(B)])
    worst, n_sqp_diff, n_status_diff, n_active_bind = 0.0, 0, 0, 0
    for i, g in enumerate(ref):
        worst = max(worst, relerr(ulog[i], g["u"]), relerr(zlog[i], g["z"]))
        assert relerr(ulog[i], g["u"]) < RTOL and relerr(zlog[i], g["z"]) < RTOL, (i, scale[i])
        st, sq, qp = slog[i, :, 0].astype(int), slog[i, :, 1].astype(int), slog[i, :, 2].astype(int)
        if scale[i] is None:
            assert np.array_equal(st, g["status"]), (i, st, g["status"])
            assert np.array_equal(sq, g["sqp_iter"]) and np.array_equal(qp, g["qp_iter"]), (i, sq, g["sqp_iter"], qp, g["qp_iter"])
            assert float(xN[i] @ P @ xN[i]) < alpha[i]
        else:
            n_sqp_diff += int(not np.array_equal(sq, g["sqp_iter"]))
            assert np.abs(sq - g["sqp_iter"]).max() <= 25, (i, sq, g["sqp_iter"])
            for k in range(ns):
                if st[k] != g["status"][k]:  # only "converged at iteration 8x-99" against "stopped at max_iter"
                    n_status_diff += 1
                    assert {int(st[k]), int(g["status"][k])} == {0, 2} and max(sq[k], g["sqp_iter"][k]) == 100, (i, k, st, g["status"])
            if st[-1] == 0:
                h = float(xN[i] @ P @ xN[i])
                assert h <= alpha[i] * (1 + 1e-6)
                n_active_bind += int(h > alpha[i] * (1 - 1e-4))
    assert n_active_bind >= B // 8  # the tightened rows really bind (instances that stopped at max_iter are not counted)
    print(f"config4 shape: worst rel err {worst:.2e}; active instances with a differing SQP count: {n_sqp_diff} of {B // 2}, "
          f"solves with status 0 vs 2 at max_iter: {n_status_diff} of {ns * B // 2}")


@pytest.mark.parametrize("solver,w_u,w_qddot", [("SQP", 0.001, 0.05), ("SQP_RTI", 0.001, 0.02), ("SQP_RTI", 0.01, 0.05),
                                                 ("SQP", 0.01, 0.02)])
def test_readme_grid_tuples_at_n40(oracle, solver, w_u, w_qddot):
    """BASELINE configs[1] at its largest horizon: README grid tuples N = 40 x (w_u, w_qddot) x solver type, 16 random
    surfaces each, 60 closed-loop steps; iteration counts (SQP and interior point) EQUAL to the reference's."""
    import robotic_mpc_b200 as rm

    o = oracle
    B, N, ns = 16, 40, 60
    surf, z0 = _inputs(o, B, 4000 + int(1000 * w_u) + int(100 * w_qddot))
    z0[:] = np.concatenate([o.README_Q0, o.README_QD0])  # the README grid varies the surface, not the initial state
    eng = rm.Engine(B, N, solver, dt=0.01, nlog=ns)
    eng.set_params(rm.pack_params(B, 0.01, surf=surf, w_u=w_u, w_qddot=w_qddot))
    eng.set_state(z0)
    eng.reset()
    eng.closed_loop(ns)
    ulog, zlog, slog = (eng.get(k).cpu().numpy() for k in ("ulog", "zlog", "slog"))
    eng.close()
    st = o.SQP_RTI if solver == "SQP_RTI" else o.SQP
    ref = _pool_closed_loops(o, [(dict(N=N, solver_type=st, surf=surf[i], w_u=w_u, w_qddot=w_qddot), z0[i], ns) for i in range(B)])
    for i, g in enumerate(ref):
        assert relerr(ulog[i], g["u"]) < RTOL and relerr(zlog[i], g["z"]) < RTOL
        assert np.array_equal(slog[i, :, 0].astype(int), g["status"])
        assert np.array_equal(slog[i, :, 1].astype(int), g["sqp_iter"])
        assert np.array_equal(slog[i, :, 2].astype(int), g["qp_iter"]), (i, slog[i, :, 2].astype(int) - g["qp_iter"])
        assert relerr(slog[i, :, 3], g["cost"]) < RTOL


def test_qp_iteration_counts_equal_to_reference(oracle):
    """Interior-point iteration counts of the device EQUAL the reference's, step by step, on the bench workload's own
    distribution (N = 100 SQP_RTI, 64 instances x 40 steps incl. the start-up transient) and on a full-SQP batch (N = 20,
    32 x 20).  Exceptions are counted, not tolerated silently: none are expected; the test fails on the first."""
    import robotic_mpc_b200 as rm

    o = oracle
    n_exc, n_tot = 0, 0
    for solver, B, N, ns in (("SQP_RTI", 64, 100, 40), ("SQP", 32, 20, 20)):
        surf, z0 = _inputs(o, B, 44)
        eng = rm.Engine(B, N, solver, dt=0.01, nlog=ns)
        eng.set_params(rm.pack_params(B, 0.01, surf=surf))
        eng.set_state(z0)
        eng.reset()
        eng.closed_loop(ns)
        ulog, slog = eng.get("ulog").cpu().numpy(), eng.get("slog").cpu().numpy()
        eng.close()
        st = o.SQP_RTI if solver == "SQP_RTI" else o.SQP
        ref = _pool_closed_loops(o, [(dict(N=N, solver_type=st, surf=surf[i]), z0[i], ns) for i in range(B)])
        for i, g in enumerate(ref):
            assert relerr(ulog[i], g["u"]) < RTOL
            assert np.array_equal(slog[i, :, 1].astype(int), g["sqp_iter"])
            n_exc += int((slog[i, :, 2].astype(int) != g["qp_iter"]).sum())
            n_tot += ns
    print(f"qp_iter: {n_exc} exceptions in {n_tot} solves")
    assert n_exc == 0


def test_bench_window_full_batch_sample_against_reference(oracle):
    """The window the driver's bench times -- bench.py's own workload (make_workload(4096, 44 + rank), rank 0), 4096 instances x
    N = 100 x SQP_RTI, the first 20 closed-loop steps from the initial state -- run as ONE batch through rmpc_run_host-sized
    geometry, with every 16th instance (256 of them, every CTA position and both halves of a warp) checked step by step
    against the compiled reference on the host cores: controls and states at 1e-6, status / SQP / QP iteration sequences equal."""
    import bench
    import robotic_mpc_b200 as rm

    o = oracle
    B, N, ns = 4096, 100, 20
    surf, z0 = bench.make_workload(B, 44)
    eng = rm.Engine(B, N, "SQP_RTI", dt=0.01, nlog=ns)
    eng.set_params(rm.pack_params(B, 0.01, surf=surf))
    eng.set_state(z0)
    eng.reset()
    eng.closed_loop(ns)
    ulog, zlog, slog = (eng.get(k).cpu().numpy() for k in ("ulog", "zlog", "slog"))
    summ = eng.get("summary").cpu().numpy()
    eng.close()
    assert int(summ[:, rm.RS["NFAIL"]].sum()) == 0
    pick = list(range(5, B, 16))
    ref = _pool_closed_loops(o, [(dict(N=N, solver_type=o.SQP_RTI, surf=surf[i]), z0[i], ns) for i in pick])
    worst = 0.0
    for i, g in zip(pick, ref):
        worst = max(worst, relerr(ulog[i], g["u"]), relerr(zlog[i], g["z"]))
        assert relerr(ulog[i], g["u"]) < RTOL and relerr(zlog[i], g["z"]) < RTOL, i
        assert np.array_equal(slog[i, :, 0].astype(int), g["status"]), i
        assert np.array_equal(slog[i, :, 1].astype(int), g["sqp_iter"]), i
        assert np.array_equal(slog[i, :, 2].astype(int), g["qp_iter"]), (i, slog[i, :, 2].astype(int) - g["qp_iter"])
        assert relerr(slog[i, :, 3], g["cost"]) < RTOL, i
    print(f"bench window, {len(pick)} of {B} instances x {ns} steps: worst relative deviation {worst:.2e}")


@pytest.mark.parametrize("N,N2,solver", [(40, 20, "SQP_RTI"), (40, 8, "SQP"), (40, 4, "SQP_RTI"), (100, 10, "SQP_RTI"), (100, 50, "SQP")])
def test_partial_condensing_cond_n(oracle, N, N2, solver):
    """qp_solver_cond_N in {N/2, N/5, N/10} through the C ABI option 'qp_cond_N': 16 instances with a
    tight joint-angle box (so that the condensed problem's interior-point path differs from the un-condensed one), closed
    loop against the compiled acados at the same cond_N: controls / states at 1e-6, status, SQP and QP iteration sequences
    equal; the number of instances where acados' own cond_N = N run takes a different QP iteration sequence is printed."""
    import robotic_mpc_b200 as rm

    o = oracle
    if not o.REF_SO.exists():
        pytest.skip("oracle/_ref did not travel (the port has no condensing)")
    B, ns = 16, 5 if solver == "SQP" else 10
    surf, z0 = _inputs(o, B, 900 + N + N2)
    rng = np.random.default_rng(N2)
    w = rng.uniform(0.06, 0.14, (B, 1))
    qmin, qmax = z0[:, :6] - w, z0[:, :6] + w
    eng = rm.Engine(B, N, solver, dt=0.01, nlog=ns)
    eng.set_params(rm.pack_params(B, 0.01, surf=surf, q_min=qmin, q_max=qmax))
    eng.set_option("qp_cond_N", N2)
    eng.set_state(z0)
    eng.reset()
    eng.closed_loop(ns)
    ulog, zlog, slog = (eng.get(k).cpu().numpy() for k in ("ulog", "zlog", "slog"))
    eng.close()
    st = o.SQP_RTI if solver == "SQP_RTI" else o.SQP
    ref = _pool_closed_loops(o, [(dict(N=N, solver_type=st, surf=surf[i], q_min=qmin[i], q_max=qmax[i], cond_N=N2), z0[i], ns) for i in range(B)])
    ref0 = _pool_closed_loops(o, [(dict(N=N, solver_type=st, surf=surf[i], q_min=qmin[i], q_max=qmax[i]), z0[i], ns) for i in range(B)])
    n_other = 0
    for i, g in enumerate(ref):
        assert relerr(ulog[i], g["u"]) < RTOL and relerr(zlog[i], g["z"]) < RTOL, i
        assert np.array_equal(slog[i, :, 0].astype(int), g["status"])
        assert np.array_equal(slog[i, :, 1].astype(int), g["sqp_iter"])
        assert np.array_equal(slog[i, :, 2].astype(int), g["qp_iter"]), (i, slog[i, :, 2], g["qp_iter"])
        n_other += int(not np.array_equal(ref0[i]["qp_iter"], g["qp_iter"]))
    print(f"cond_N {N2} of {N}: acados' cond_N = N run differs in its QP iteration sequence for {n_other} of {B} instances")
