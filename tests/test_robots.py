"""Kinematic chains other than the UR10: chain tables, the URDF reader, the axis-pattern folding of
rmpc_set_chain, and closed-loop parity of the device code (host lane emulator) against the oracle with the UR5 chain.
The reference picks the chain by robot_name (R/loader.py:24-36, R/prediction_model.py:28-31)."""
import numpy as np
import pytest

import robotic_mpc_b200 as rm
import robotic_mpc_b200.robots as rb

pytestmark = pytest.mark.timeout(600)

URDF5 = """<robot name="mini">
  <link name="base_link"/><link name="l0"/><link name="l1"/><link name="l2"/><link name="l3"/><link name="l4"/><link name="l5"/>
  <link name="l6"/><link name="flange"/><link name="tool0"/>
  <joint name="fix0" type="fixed"><parent link="base_link"/><child link="l0"/><origin rpy="0 0 0.5" xyz="0.1 0 0"/></joint>
  <joint name="j1" type="revolute"><parent link="l0"/><child link="l1"/><origin rpy="0 0 3.141592653589793" xyz="0 0 0.089159"/><axis xyz="0 0 1"/></joint>
  <joint name="j2" type="revolute"><parent link="l1"/><child link="l2"/><origin rpy="1.570796327 0 0" xyz="0 0 0"/><axis xyz="1 0 0"/></joint>
  <joint name="j3" type="continuous"><parent link="l2"/><child link="l3"/><origin rpy="0 0 0" xyz="-0.425 0 0"/><axis xyz="0 1 0"/></joint>
  <joint name="j4" type="revolute"><parent link="l3"/><child link="l4"/><origin rpy="0 0.3 0" xyz="-0.39225 0 0.10915"/><axis xyz="0 0 1"/></joint>
  <joint name="j5" type="revolute"><parent link="l4"/><child link="l5"/><origin rpy="1.570796327 0 0" xyz="0 -0.09465 0"/><axis xyz="1 0 0"/></joint>
  <joint name="j6" type="revolute"><parent link="l5"/><child link="l6"/><origin rpy="0.2 0.1 0" xyz="0 0.0823 0"/><axis xyz="0 0 1"/></joint>
  <joint name="f1" type="fixed"><parent link="l6"/><child link="flange"/><origin rpy="0 -1.5707963267948966 -1.5707963267948966" xyz="0 0 0.01"/></joint>
  <joint name="f2" type="fixed"><parent link="flange"/><child link="tool0"/><origin rpy="1.5707963267948966 0 1.5707963267948966" xyz="0 0.02 0"/></joint>
</robot>"""


def _fk_pattern(Rfix, pfix, q):
    """FK the way the kernels do it: fixed axis pattern z y y y z y over a folded table."""
    R, p = np.eye(3), np.zeros(3)
    for j in range(6):
        p = p + R @ pfix[j]
        Rj = R @ Rfix[j].reshape(3, 3)
        c, s = np.cos(q[j]), np.sin(q[j])
        Rq = np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]]) if rb.PATTERN[j] == 2 else np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]])
        R = Rj @ Rq
    return p + R @ pfix[6], R @ Rfix[6].reshape(3, 3)


def test_ur5_known_answer_and_ur10_table():
    """UR5 at q = 0: tool0 at (a2 + a3, d4 + d6, d1 - d5) of the DH table in the REP-103 base_link frame =
    (0.81725, 0.19145, -0.005491) (R/ur_description/urdf/ur5.urdf:284-357).  UR10 at q = 0: ee_link from the joint origins."""
    t, a = rb.chain_table("ur5")
    p, R = rb.forward_kinematics(t, a, np.zeros(6))
    assert np.allclose(p, [0.425 + 0.39225, 0.10915 + 0.0823, 0.089159 - 0.09465], atol=1e-9)
    assert np.allclose(R @ R.T, np.eye(3), atol=1e-12) and np.isclose(np.linalg.det(R), 1.0)
    assert list(a) == [2] * 6
    t, a = rb.chain_table("ur10")
    assert list(a) == list(rb.PATTERN)
    p, _ = rb.forward_kinematics(t, a, np.zeros(6))
    # the two rpy = (0, pi/2, 0) origins turn the arm along +x: x = 0.612 + 0.5723, z = 0.1273 - 0.1157, y = sum of y offsets
    assert np.allclose(p, [0.612 + 0.5723, 0.220941 - 0.1719 + 0.1149 + 0.0922, 0.1273 - 0.1157], atol=1e-8)
    with pytest.raises(ValueError):
        rb.chain_table("ur3")


def test_pattern_folding_is_exact_for_any_axes():
    rng = np.random.default_rng(5)
    for name in ("ur5", "ur10"):
        t, a = rb.chain_table(name)
        Rf, pf = rb.device_chain(t, a)
        for _ in range(4):
            q = rng.uniform(-3, 3, 6)
            p1, R1 = rb.forward_kinematics(t, a, q)
            p2, R2 = _fk_pattern(Rf, pf, q)
            assert np.abs(p1 - p2).max() < 1e-14 and np.abs(R1 - R2).max() < 1e-14
    for _ in range(20):
        a = rng.integers(0, 3, 6)
        t = np.array([rb._row(rng.normal(size=3), rng.normal(size=3)) for _ in range(8)])
        Rf, pf = rb.device_chain(t, a)
        q = rng.uniform(-3, 3, 6)
        p1, R1 = rb.forward_kinematics(t, a, q)
        p2, R2 = _fk_pattern(Rf, pf, q)
        assert np.abs(p1 - p2).max() < 1e-13 and np.abs(R1 - R2).max() < 1e-13


def test_urdf_reader(tmp_path):
    f = tmp_path / "mini.urdf"
    f.write_text(URDF5)
    t, a = rb.chain_from_urdf(f, "tool0")
    assert list(a) == [2, 0, 1, 2, 0, 2]
    assert np.allclose(t[0], rb._row((0.1, 0, 0), (0, 0, 0.5)))            # fixed joint in front of joint 1 -> root row
    assert np.allclose(t[4], rb._row((-0.39225, 0, 0.10915), (0, 0.3, 0)))
    assert np.allclose(t[7], rb._compose([((0, 0, 0.01), (0, -rb.HALF_PI, -rb.HALF_PI)), ((0, 0.02, 0), (rb.HALF_PI, 0, rb.HALF_PI))]))
    ld = rm.UrdfLoader("mini", urdf_path=f, ee_frame="tool0")
    m = rm.SixDofRobot(urdf_loader=ld, Ts=0.01, Wcv=np.full(6, 200.0))
    assert m.robot_name == "mini" and m.chain[0].shape == (8, 12)
    p, R = m.forward_kinematics(np.zeros(6))
    assert np.allclose(R @ R.T, np.eye(3), atol=1e-12)
    with pytest.raises(ValueError):
        rb.chain_from_urdf(f, "l3")  # only 3 revolute joints up to there
    bad = tmp_path / "bad.urdf"
    bad.write_text(URDF5.replace('<axis xyz="0 1 0"/>', '<axis xyz="0 0.6 0.8"/>'))
    with pytest.raises(ValueError):
        rb.chain_from_urdf(bad, "tool0")


def test_emulated_device_fk_matches_table_and_oracle(built, tmp_path):
    """rmpc_chain_from_table (the C side of rmpc_set_chain) + the device FK vs numpy FK of the table and vs the oracle's
    generalised ur10_fk, for the UR5 and for a chain with mixed axes."""
    from tests import emu
    import oracle.pyoracle as po

    f = tmp_path / "mini.urdf"
    f.write_text(URDF5)
    port = po.PortLeg()
    model = po.Model(port.lib)
    rng = np.random.default_rng(11)
    par = rm.pack_params(1, 0.01)[0]
    try:
        for t, a in (rb.chain_table("ur5"), rb.chain_from_urdf(f, "tool0")):
            emu.set_chain(t, a)
            port.set_chain(t, a)
            for _ in range(3):
                z = np.concatenate([rng.uniform(-3, 3, 6), rng.uniform(-1, 1, 6)])
                rec, y5 = emu.plant_record(par, z)
                p, R = rb.forward_kinematics(t, a, z[:6])
                assert np.abs(rec[rm.RK["POSE"]:rm.RK["POSE"] + 3] - p).max() < 1e-14
                assert np.abs(rec[rm.RK["POSE"] + 3:rm.RK["POSE"] + 12].reshape(3, 3) - R).max() < 1e-14
                pose, rpy, tw = model.plant_logs(z)
                assert np.abs(pose - rec[rm.RK["POSE"]:rm.RK["POSE"] + 12]).max() < 1e-14
                assert np.abs(tw - rec[rm.RK["TWIST"]:rm.RK["TWIST"] + 6]).max() < 1e-13
                P = po.make_problem(N=10).model
                y = model.cost_y(P, z, np.zeros(6))
                assert np.abs(y[:5] - y5).max() < 1e-13
    finally:
        emu.set_chain(None)
        port.set_chain(None)


@pytest.mark.parametrize("solver", ["SQP_RTI", "SQP"])
def test_ur5_closed_loop_matches_oracle(built, solver):
    """The device closed loop (host lane emulator) with the UR5 chain against the plain-C oracle and, when it is built, the
    compiled acados reference -- same tolerances as the UR10 golden cases."""
    from tests import emu
    import oracle.pyoracle as po

    t, a = rb.chain_table("ur5")
    z0 = np.array([0.0, -1.8, 1.9, -1.7, -1.57, 0.0, 0.2, 0.1, -0.1, 0.0, 0.0, 0.0])
    par = rm.pack_params(1, 0.01)
    st = po.SQP_RTI if solver == "SQP_RTI" else po.SQP
    prob = po.make_problem(N=10, solver_type=st)
    legs = [po.PortLeg()]
    if po.REF_SO.exists():
        legs.append(po.RefLeg())
    try:
        emu.set_chain(t, a)
        e = emu.closed_loop(par, z0[None], 10, 1 if solver == "SQP_RTI" else 0, 12)
        assert np.abs(e["z"][0, -1] - z0).max() > 0.05  # the arm moves
        for leg in legs:
            leg.set_chain(t, a)
            o = leg.closed_loop(prob, z0, 12)
            assert np.abs(e["z"][0] - o["z"]).max() < 1e-9 and np.abs(e["u"][0] - o["u"]).max() < 1e-8
            assert np.array_equal(e["slog"][0, :, 0].astype(int), o["status"])
            assert np.array_equal(e["slog"][0, :, 1].astype(int), o["sqp_iter"])
        # and it is a different problem from the UR10 one
        emu.set_chain(None)
        e10 = emu.closed_loop(par, z0[None], 10, 1 if solver == "SQP_RTI" else 0, 12)
        assert np.abs(e10["u"][0] - e["u"][0]).max() > 1e-3
    finally:
        emu.set_chain(None)
        for leg in legs:
            leg.set_chain(None)


ROBOT_GOLD = __import__("pathlib").Path(__file__).resolve().parent / "golden" / "robots"
ROBOT_CASES = sorted(p.stem for p in ROBOT_GOLD.glob("*.npz"))


def _gold_problem(po, g):
    term = {}
    if "term_P" in g.files:
        term = dict(term_P=g["term_P"], term_alpha=float(g["term_alpha"]), term_lh=float(g["term_lh"]))
    return po.make_problem(N=int(g["N"]), solver_type=int(g["solver_type"]), surf=g["surf"], **term)


@pytest.mark.parametrize("case", ROBOT_CASES)
def test_ur5_golden_of_the_reference(built, case):
    """Vectors of the compiled acados reference on the UR5 chain (tests/golden/make_robot_golden.py): the built-in table equals
    the one read from the reference's URDF; the plain-C oracle and the emulated device code reproduce the closed loop."""
    from tests import emu
    import oracle.pyoracle as po

    g = np.load(ROBOT_GOLD / f"{case}.npz")
    t, a = rb.chain_table("ur5")
    assert np.array_equal(a, g["chain_axis"]) and np.abs(t - g["chain_xyzR"]).max() == 0.0
    ns = min(int(g["nsteps"]), 30)
    port = po.PortLeg()
    term = None
    if "term_P" in g.files:
        term = np.concatenate([g["term_P"].ravel(), [float(g["term_lh"]), float(g["term_alpha"])]])[None]
    par = rm.pack_params(1, 0.01, surf=g["surf"][None])
    try:
        port.set_chain(t, a)
        emu.set_chain(t, a)
        o = port.closed_loop(_gold_problem(po, g), g["z0"], ns)
        e = emu.closed_loop(par, g["z0"][None], int(g["N"]), 1 if int(g["solver_type"]) == po.SQP_RTI else 0, ns, term=term)
    finally:
        port.set_chain(None)
        emu.set_chain(None)
    for r_u, r_z, r_st, r_it in ((o["u"], o["z"], o["status"], o["sqp_iter"]),
                                 (e["u"][0], e["z"][0], e["slog"][0, :, 0].astype(int), e["slog"][0, :, 1].astype(int))):
        assert np.abs(r_u - g["u"][:ns]).max() / np.abs(g["u"][:ns]).max() < 1e-6
        assert np.abs(r_z - g["z"][:ns + 1]).max() / np.abs(g["z"][:ns + 1]).max() < 1e-6
        assert np.array_equal(r_st, g["status"][:ns]) and np.array_equal(r_it, g["sqp_iter"][:ns])
