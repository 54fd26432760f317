"""Second, independent restatement of the reference's model functions in PyTorch FP64 with autograd (SURVEY.md 7 step 2,
8(c) mitigation 1): the hand-written C of oracle/ur10_model.c -- the one part of the parity chain that is not the
reference's own compiled code -- against a restatement that shares nothing with it but the URDF numbers.

  - kinematics: 4x4 homogeneous transforms built from the <origin xyz rpy> / <axis> entries of
    R/ur_description/urdf/ur10.urdf:50-55,78-83,106-111,134-139,162-167,190-195,218-222 with URDF's fixed-axis roll-pitch-yaw
    (R = Rz(yaw) Ry(pitch) Rx(roll)) and Rodrigues' formula for the joints; the C code goes through quaternions and
    per-axis rotations.  What pinocchio gives the reference at R/prediction_model.py:117-132 (oMf of frame ee_link).
  - WORLD frame Jacobian (R/prediction_model.py:163-164): obtained from autograd derivatives of the pose -- angular part
    from dR/dq R', linear part d p/dq + p x a, i.e. the velocity of the body point passing through the world origin -- where
    the C code writes the columns [p_j x a_j; a_j] directly.
  - outputs and their Jacobian: the expressions of R/prediction_model.py:256-326 (task frame, v_task), R/surface.py:243-262
    (unit normal), R/trajectory_optimizer.py:88-155 (g1..g5, u, qddot, exp(-2 sqrt(det(J J') + eps))) written out in torch and
    differentiated with torch.autograd.functional.jacobian, where the C code carries a hand-derived analytic Jacobian.
Agreement: <= 1e-12 (absolute on O(1) quantities)."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

TOL = 1e-12
# <origin xyz>, <origin rpy>, <axis xyz> of the six revolute joints, then the fixed joint wrist_3_link -> ee_link
XYZ = [(0.0, 0.0, 0.1273), (0.0, 0.220941, 0.0), (0.0, -0.1719, 0.612), (0.0, 0.0, 0.5723), (0.0, 0.1149, 0.0), (0.0, 0.0, 0.1157),
       (0.0, 0.0922, 0.0)]
RPY = [(0.0, 0.0, 0.0), (0.0, 1.570796325, 0.0), (0.0, 0.0, 0.0), (0.0, 1.570796325, 0.0), (0.0, 0.0, 0.0), (0.0, 0.0, 0.0),
       (-1.5707963267948966, 0.0, 0.0)]
AXIS = [(0, 0, 1), (0, 1, 0), (0, 1, 0), (0, 1, 0), (0, 0, 1), (0, 1, 0)]
f64 = torch.float64


def _rpy(r, p, y):
    cr, sr, cp, sp, cy, sy = np.cos(r), np.sin(r), np.cos(p), np.sin(p), np.cos(y), np.sin(y)
    Rx = np.array([[1, 0, 0], [0, cr, -sr], [0, sr, cr]])
    Ry = np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
    Rz = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def _hom(R, t):
    T = torch.eye(4, dtype=f64)
    T = T.clone()
    T[:3, :3] = R
    T[:3, 3] = t
    return T


def _skew(a):
    z = torch.zeros((), dtype=f64)
    return torch.stack([torch.stack([z, -a[2], a[1]]), torch.stack([a[2], z, -a[0]]), torch.stack([-a[1], a[0], z])])


def fk(q):
    """Pose of frame ee_link: (p [3], R [3, 3])."""
    T = torch.eye(4, dtype=f64)
    for j in range(6):
        T = T @ _hom(torch.tensor(_rpy(*RPY[j]), dtype=f64), torch.tensor(XYZ[j], dtype=f64))
        K = _skew(torch.tensor(AXIS[j], dtype=f64))
        Rj = torch.eye(3, dtype=f64) + torch.sin(q[j]) * K + (1.0 - torch.cos(q[j])) * (K @ K)  # Rodrigues
        T = T @ _hom(Rj, torch.zeros(3, dtype=f64))
    T = T @ _hom(torch.tensor(_rpy(*RPY[6]), dtype=f64), torch.tensor(XYZ[6], dtype=f64))
    return T[:3, 3], T[:3, :3]


def jacobian_world(q, create_graph=False):
    """pinocchio computeFrameJacobian(.., WORLD): rows [linear at the world origin (3); angular (3)], from the pose derivatives."""
    dp, dR = torch.autograd.functional.jacobian(fk, q, create_graph=create_graph)  # [3, 6], [3, 3, 6]
    p, R = fk(q)
    cols = []
    for j in range(6):
        Wj = dR[:, :, j] @ R.T  # skew(a_j)
        a = torch.stack([Wj[2, 1], Wj[0, 2], Wj[1, 0]])
        cols.append(torch.cat([dp[:, j] + torch.linalg.cross(p, a), a]))
    return torch.stack(cols, dim=1)


def cost_y(v, surf, wcv, Ts, t_ee, create_graph=False):
    """The 18 outputs of the NONLINEAR_LS cost as functions of the stage variable v = [u; q; qdot]."""
    u, q, qd = v[:6], v[6:12], v[12:]
    p_ee, R_ee = fk(q)
    J = jacobian_world(q, create_graph=True)
    vee = J @ qd
    v_w, w_w = vee[:3], vee[3:]
    tw = R_ee @ t_ee  # task frame = EE frame shifted by t_ee (R/prediction_model.py:256-292: R_ee_t is the identity)
    p_t = p_ee + tw
    v_t = R_ee.T @ (v_w + torch.linalg.cross(w_w, tw))
    a, b, c, d, e, f = surf
    S = a * p_t[0] ** 2 + b * p_t[1] ** 2 + c * p_t[0] * p_t[1] + d * p_t[0] + e * p_t[1] + f
    n = torch.stack([2 * a * p_t[0] + c * p_t[1] + d, 2 * b * p_t[1] + c * p_t[0] + e, torch.tensor(-1.0, dtype=f64)])
    n = n / torch.sqrt((n * n).sum())
    g = torch.stack([S - p_t[2], (n * R_ee[:, 2]).sum(), R_ee[0, 1], p_t[0], v_t[1]])
    a22 = torch.exp(-wcv * Ts)
    qddot = ((a22 * qd + (1.0 - a22) * u) - qd) / Ts
    manip = torch.sqrt(torch.linalg.det(J @ J.T) + 0.01)
    return torch.cat([g, u, qddot, torch.exp(-2.0 * manip).reshape(1)])


@pytest.fixture(scope="module")
def model(oracle):
    return oracle.Model(oracle.PortLeg().lib)


def _points(n, seed):
    rng = np.random.default_rng(seed)
    return [(rng.uniform(-2, 2, 6), rng.uniform(-1, 1, 6), rng.uniform(-1, 1, 6)) for _ in range(n)]


def test_fk_and_world_jacobian_match_torch(oracle, model):
    pts = _points(6, 11) + [(oracle.README_Q0, oracle.README_QD0, np.zeros(6)), (np.zeros(6), np.ones(6), np.zeros(6))]
    for q, _, _ in pts:
        _, _, R, p = model.fk(q)
        tq = torch.tensor(q, dtype=f64)
        pt, Rt = fk(tq)
        assert np.abs(pt.numpy() - p).max() <= TOL and np.abs(Rt.numpy() - R).max() <= TOL
        assert np.abs(jacobian_world(tq).numpy() - model.jacobian_world(q)).max() <= TOL


def test_cost_outputs_and_jacobian_match_torch_autograd(oracle, model):
    worst_y = worst_J = 0.0
    for k, (q, qd, u) in enumerate(_points(5, 12) + [(oracle.README_Q0, oracle.README_QD0, 0.3 * np.ones(6))]):
        rng = np.random.default_rng(100 + k)
        surf = np.array([-0.1, 0.1, -0.01, 0.01, 0.01, 0.0]) + 0.02 * rng.standard_normal(6)
        P = oracle.make_problem(surf=surf, wcv=150.0 + 20.0 * k).model
        x = np.concatenate([q, qd])
        y, Jt = model.cost_y_jac(P, x, u)  # Jt[r, c] = d y_c / d v_r, v = [u; x]
        assert np.array_equal(y, model.cost_y(P, x, u))
        args = (torch.tensor(np.array(P.surf), dtype=f64), torch.tensor(np.array(P.wcv), dtype=f64), float(P.Ts),
                torch.tensor(np.array(P.t_ee), dtype=f64))
        v = torch.tensor(np.concatenate([u, x]), dtype=f64)
        yt = cost_y(v, *args).detach().numpy()
        Jy = torch.autograd.functional.jacobian(lambda vv: cost_y(vv, *args), v).numpy()  # [c, r]
        # column 17, the manipulability output: the C code returns its derivative as ZERO on purpose (oracle/ur10_model.h:69-72:
        # weight 0 in W, so it contributes exactly 0.0 to gradient and Gauss-Newton Hessian); the value y[17] is compared
        assert not Jt[:, 17].any() and np.abs(Jy[17]).max() > 1e-3
        worst_y, worst_J = max(worst_y, np.abs(yt - y).max()), max(worst_J, np.abs(Jy.T - Jt)[:, :17].max())
    print(f'C vs torch autograd: outputs {worst_y:.2e}, Jacobian {worst_J:.2e}')
    assert worst_y <= TOL, worst_y
    assert worst_J <= TOL, worst_J


def test_discrete_dynamics_match_torch(oracle, model):
    """x+ = Ad x + Bd u with the exact-ZOH blocks of R/prediction_model.py:87-115, and its constant Jacobian [B A]'."""
    P = oracle.make_problem(wcv=173.0).model
    w, Ts = torch.tensor(np.array(P.wcv), dtype=f64), float(P.Ts)

    def phi(v):
        u, q, qd = v[:6], v[6:12], v[12:]
        a22 = torch.exp(-w * Ts)
        a12 = (1.0 - a22) / w
        return torch.cat([q + a12 * qd + (Ts - a12) * u, a22 * qd + (1.0 - a22) * u])

    for q, qd, u in _points(3, 13):
        v = torch.tensor(np.concatenate([u, q, qd]), dtype=f64)
        assert np.abs(phi(v).numpy() - model.dyn(P, np.concatenate([q, qd]), u)).max() <= TOL
        Jphi = torch.autograd.functional.jacobian(phi, v).numpy()  # [12, 18]
        assert np.abs(Jphi.T - model.BAbt(P)).max() <= TOL
