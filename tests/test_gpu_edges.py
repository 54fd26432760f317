"""Edge cases of the CUDA path through the C ABI: smallest horizon, ragged batch sizes, failing instances beside healthy
ones, log capacity, determinism and independence of batch composition, error behaviour of the entry points.  The reference
ships no tests of its own; the cases follow what its API allows (R/simulator.py:199-241)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
RTOL = 1e-6


def _inputs(o, B, seed):
    rng = np.random.default_rng(seed)
    surf = np.array([-0.1, 0.1, -0.01, 0.01, 0.01, 0.0]) + 0.01 * rng.standard_normal((B, 6))
    z0 = np.concatenate([o.README_Q0 + rng.uniform(-0.05, 0.05, (B, 6)), o.README_QD0 + rng.uniform(-0.1, 0.1, (B, 6))], 1)
    return surf, z0


def _closed_loop(rm, B, N, solver, par, z0, ns, nlog=None):
    eng = rm.Engine(B, N, solver, dt=0.01, nlog=ns if nlog is None else nlog)
    eng.set_params(par)
    eng.set_state(z0)
    eng.reset()
    eng.closed_loop(ns)
    out = {k: eng.get(k).cpu().numpy() for k in ("ulog", "zlog", "slog", "summary", "z")}
    eng.close()
    return out


@pytest.mark.parametrize("solver", ["SQP_RTI", "SQP"])
@pytest.mark.parametrize("B", [1, 3, 9])
def test_minimum_horizon_and_ragged_batches(oracle, solver, B):
    """N = 2 (the smallest the reference's OCP layout allows: one interior stage) with batch sizes that do not fill a CTA."""
    import robotic_mpc_b200 as rm

    o = oracle
    N, ns = 2, 12
    surf, z0 = _inputs(o, B, 300 + B)
    r = _closed_loop(rm, B, N, solver, rm.pack_params(B, 0.01, surf=surf), z0, ns)
    leg = o.RefLeg() if o.REF_SO.exists() else o.PortLeg()
    st = o.SQP_RTI if solver == "SQP_RTI" else o.SQP
    for i in range(B):
        g = leg.closed_loop(o.make_problem(N=N, solver_type=st, surf=surf[i]), z0[i], ns)
        assert np.abs(r["ulog"][i] - g["u"]).max() / np.abs(g["u"]).max() < RTOL
        assert np.abs(r["zlog"][i] - g["z"]).max() / np.abs(g["z"]).max() < RTOL
        assert np.array_equal(r["slog"][i, :, 0].astype(int), g["status"])
        assert np.array_equal(r["slog"][i, :, 1].astype(int), g["sqp_iter"])


@pytest.mark.parametrize("solver", ["SQP_RTI", "SQP"])
def test_failing_instance_beside_healthy_ones(oracle, solver):
    """An instance whose initial joint angle lies outside the joint box: the reference's QP fails (status 4, zero control) at
    every step; the device reports the same and the other instances of the batch are untouched, bit for bit."""
    import robotic_mpc_b200 as rm

    o = oracle
    B, N, ns = 6, 10, 8
    surf, z0 = _inputs(o, B, 77)
    par = rm.pack_params(B, 0.01, surf=surf)
    clean = _closed_loop(rm, B, N, solver, par, z0, ns)
    zbad = z0.copy()
    zbad[2, 1] = 2 * np.pi + 0.3
    r = _closed_loop(rm, B, N, solver, par, zbad, ns)
    leg = o.RefLeg() if o.REF_SO.exists() else o.PortLeg()
    g = leg.closed_loop(o.make_problem(N=N, solver_type=o.SQP_RTI if solver == "SQP_RTI" else o.SQP, surf=surf[2]), zbad[2], ns)
    assert np.all(g["status"] != 0)
    assert np.array_equal(r["slog"][2, :, 0].astype(int), g["status"])
    assert np.abs(r["ulog"][2] - g["u"]).max() <= 1e-12 and np.abs(r["zlog"][2] - g["z"]).max() <= 1e-9
    assert r["summary"][2, rm.RS["NFAIL"]] == ns
    keep = [0, 1, 3, 4, 5]
    assert np.array_equal(r["ulog"][keep], clean["ulog"][keep]) and np.array_equal(r["zlog"][keep], clean["zlog"][keep])
    assert np.all(r["summary"][keep, rm.RS["NFAIL"]] == 0)


def test_non_finite_input_is_contained(built):
    """A NaN in one instance's state ends in a failure status for that instance only; nothing hangs, neighbours are bit-identical."""
    import oracle.pyoracle as o
    import robotic_mpc_b200 as rm

    B, N, ns = 5, 10, 5
    surf, z0 = _inputs(o, B, 78)
    par = rm.pack_params(B, 0.01, surf=surf)
    clean = _closed_loop(rm, B, N, "SQP_RTI", par, z0, ns)
    zbad = z0.copy()
    zbad[4, 3] = np.nan
    r = _closed_loop(rm, B, N, "SQP_RTI", par, zbad, ns)
    assert np.all(r["slog"][4, :, 0] != 0) and r["summary"][4, rm.RS["NFAIL"]] == ns
    assert np.array_equal(r["ulog"][:4], clean["ulog"][:4]) and np.array_equal(r["zlog"][:4], clean["zlog"][:4])


def test_deterministic_and_independent_of_batch_composition(built):
    """Same inputs -> same bits, run to run; an instance's results do not depend on its neighbours, its position in the batch
    or the batch size (a group of 16 lanes owns an instance, nothing is shared)."""
    import oracle.pyoracle as o
    import robotic_mpc_b200 as rm

    B, N, ns = 11, 20, 10
    surf, z0 = _inputs(o, B, 79)
    par = rm.pack_params(B, 0.01, surf=surf)
    a = _closed_loop(rm, B, N, "SQP", par, z0, ns)
    b = _closed_loop(rm, B, N, "SQP", par, z0, ns)
    assert all(np.array_equal(a[k][..., :10], b[k][..., :10]) for k in ("ulog", "zlog")) and np.array_equal(a["slog"][:, :, :10], b["slog"][:, :, :10])
    perm = np.random.default_rng(1).permutation(B)
    c = _closed_loop(rm, B, N, "SQP", par[perm], z0[perm], ns)
    assert np.array_equal(c["ulog"], a["ulog"][perm]) and np.array_equal(c["zlog"], a["zlog"][perm])
    d = _closed_loop(rm, 1, N, "SQP", par[7:8], z0[7:8], ns)
    assert np.array_equal(d["ulog"][0], a["ulog"][7]) and np.array_equal(d["summary"][0, :10], a["summary"][7, :10])


def test_log_capacity_and_step_counter(built):
    """More closed-loop steps than the logs can hold: the first nlog steps are logged, the summaries count every step, and the
    run is the same as one with room for everything; repeated closed_loop calls continue where the last one stopped."""
    import oracle.pyoracle as o
    import robotic_mpc_b200 as rm

    B, N = 4, 10
    surf, z0 = _inputs(o, B, 80)
    par = rm.pack_params(B, 0.01, surf=surf)
    full = _closed_loop(rm, B, N, "SQP_RTI", par, z0, 12)
    short = _closed_loop(rm, B, N, "SQP_RTI", par, z0, 12, nlog=5)
    assert short["ulog"].shape == (B, 5, 6) and np.array_equal(short["ulog"], full["ulog"][:, :5])
    assert np.array_equal(short["zlog"], full["zlog"][:, :6]) and np.array_equal(short["z"], full["z"])
    assert np.array_equal(short["summary"][:, :10], full["summary"][:, :10])
    assert np.all(short["summary"][:, rm.RS["NSAMPLES"]] == 13)
    eng = rm.Engine(B, N, "SQP_RTI", dt=0.01, nlog=12)
    eng.set_params(par)
    eng.set_state(z0)
    eng.reset()
    for n in (0, 5, 1, 6):
        eng.closed_loop(n)
    assert eng.lib.rmpc_steps_done(eng.h) == 12
    assert np.array_equal(eng.get("ulog").cpu().numpy(), full["ulog"]) and np.array_equal(eng.get("z").cpu().numpy(), full["z"])
    eng.reset()  # reset starts the same closed loop again from the current plant state
    eng.set_state(z0)
    eng.reset()
    eng.closed_loop(12)
    assert np.array_equal(eng.get("ulog").cpu().numpy(), full["ulog"])
    eng.close()


def test_error_behaviour_of_the_entry_points(built):
    import robotic_mpc_b200 as rm

    with pytest.raises(rm.RmpcError):
        rm.Engine(4, 1, "SQP_RTI")          # N < 2
    with pytest.raises(rm.RmpcError):
        rm.Engine(0, 10, "SQP_RTI")         # empty batch
    with pytest.raises(rm.RmpcError):
        rm.Engine(4, 10, "SQP_RTI", device=99)
    with pytest.raises(KeyError):
        rm.Engine(4, 10, "DDP")
    eng = rm.Engine(4, 10, "SQP", nlog=0)
    with pytest.raises(rm.RmpcError):
        eng.solve(rti_phase=1)              # rti_phase needs SQP_RTI
    with pytest.raises(rm.RmpcError):
        eng.get("x", 11)
    with pytest.raises(rm.RmpcError):
        eng.get("u", 10)
    with pytest.raises(rm.RmpcError):
        eng.get("lam", 10)                  # no terminal row on this handle
    with pytest.raises(rm.RmpcError):
        eng.closed_loop(-1)
    with pytest.raises(ValueError):
        eng.set_state(np.zeros((3, 12)))    # wrong batch size
    with pytest.raises(ValueError):
        eng.set_params(np.zeros((4, 5)))
    with pytest.raises(rm.RmpcError):
        eng.get("stats")                    # per-step records need nlog > 0
    eng.closed_loop(0)                      # no-op
    assert eng.lib.rmpc_steps_done(eng.h) == 0
    eng.close()
    eng.close()                             # idempotent


def test_checkpoint_resume_is_bit_identical(built, tmp_path):
    """rmpc_store_state / rmpc_load_state: a closed loop interrupted, saved to a file, and resumed in a NEW handle continues
    bit for bit (iterate, multipliers, merit weights, QP memory, plant state, summaries, step counter travel in the blob)."""
    import oracle.pyoracle as o
    import robotic_mpc_b200 as rm

    B, N = 6, 12
    surf, z0 = _inputs(o, B, 81)
    par = rm.pack_params(B, 0.01, surf=surf)
    for solver in ("SQP_RTI", "SQP"):
        full = _closed_loop(rm, B, N, solver, par, z0, 16)
        with rm.Engine(B, N, solver, dt=0.01, nlog=16) as eng:
            eng.set_params(par)
            eng.set_state(z0)
            eng.reset()
            eng.closed_loop(7)
            blob = eng.store_state()
            np.save(tmp_path / "ckpt.npy", blob.cpu().numpy())
            eng.closed_loop(9)
            assert np.array_equal(eng.get("ulog").cpu().numpy(), full["ulog"])
        with rm.Engine(B, N, solver, dt=0.01, nlog=16) as eng2:
            eng2.set_params(par)  # parameters belong to the problem, not to the state
            eng2.load_state(np.load(tmp_path / "ckpt.npy"))
            assert eng2.lib.rmpc_steps_done(eng2.h) == 7
            eng2.closed_loop(9)
            assert np.array_equal(eng2.get("ulog").cpu().numpy()[:, 7:], full["ulog"][:, 7:])
            assert np.array_equal(eng2.get("z").cpu().numpy(), full["z"])
            assert np.array_equal(eng2.get("summary").cpu().numpy()[:, :10], full["summary"][:, :10])
            assert np.all(eng2.get("summary").cpu().numpy()[:, rm.RS["NSAMPLES"]] == 17)
        with rm.Engine(B, N + 1, solver, dt=0.01) as other:  # same slab size (stage slots are padded to even), other N
            with pytest.raises((ValueError, rm.RmpcError)):
                other.load_state(blob)
        with rm.Engine(B + 2, N, solver, dt=0.01) as other:
            bad = np.zeros(int(other.lib.rmpc_state_size(other.h)))
            with pytest.raises(rm.RmpcError):
                other.load_state(bad)
