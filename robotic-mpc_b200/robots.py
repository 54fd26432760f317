"""Kinematic chains of the supported arms as constant tables -- the host-side "URDF -> constant table" step that replaces
pinocchio's model building (R/loader.py:24-36, R/prediction_model.py:28-31) for the device kernels.

A chain is ``(xyzR, axis)``: ``xyzR`` [8, 12] = rigid transforms ``[xyz(3); R row-major(9)]`` of
  row 0      root: fixed joints between the base link and the parent of joint 1,
  rows 1..6  origins of the six revolute joints (URDF ``<origin>``, rpy -> rotation as urdfdom does),
  row 7      joint 6 -> end-effector frame (fixed joints folded),
and ``axis`` [6] the joint axes (0 x, 1 y, 2 z).

The kernels are written for the UR10 axis pattern (z y y y z y, compile-time): ``device_chain`` folds constant rotations
into the fixed transforms so that ANY axis assignment runs through the same code, R_actual(q) = C R_pattern(q) C'.
"""
from __future__ import annotations

import functools
import xml.etree.ElementTree as ET

import numpy as np

PATTERN = (2, 1, 1, 1, 2, 1)  # joint axes the device code assumes (robotic-mpc_b200/csrc/rmpc_device.cuh, ur10_task)


def rpy_to_rot(rpy):
    """URDF rpy -> rotation matrix through a normalised quaternion, as urdfdom / pinocchio's URDF parser do."""
    r, p, y = (0.5 * float(v) for v in rpy)
    x = np.sin(r) * np.cos(p) * np.cos(y) - np.cos(r) * np.sin(p) * np.sin(y)
    yy = np.cos(r) * np.sin(p) * np.cos(y) + np.sin(r) * np.cos(p) * np.sin(y)
    z = np.cos(r) * np.cos(p) * np.sin(y) - np.sin(r) * np.sin(p) * np.cos(y)
    w = np.cos(r) * np.cos(p) * np.cos(y) + np.sin(r) * np.sin(p) * np.sin(y)
    n = np.sqrt(x * x + yy * yy + z * z + w * w)
    x, yy, z, w = x / n, yy / n, z / n, w / n
    return np.array([[1 - 2 * (yy * yy + z * z), 2 * (x * yy - z * w), 2 * (x * z + yy * w)],
                     [2 * (x * yy + z * w), 1 - 2 * (x * x + z * z), 2 * (yy * z - x * w)],
                     [2 * (x * z - yy * w), 2 * (yy * z + x * w), 1 - 2 * (x * x + yy * yy)]])


def _row(xyz, rpy):
    return np.concatenate([np.asarray(xyz, float), rpy_to_rot(rpy).ravel()])


def _compose(rows):
    """Product of rigid transforms given as (xyz, rpy) pairs -> one [12] row."""
    p, R = np.zeros(3), np.eye(3)
    for xyz, rpy in rows:
        p = p + R @ np.asarray(xyz, float)
        R = R @ rpy_to_rot(rpy)
    return np.concatenate([p, R.ravel()])


HALF_PI = 1.5707963267948966

# R/ur_description/urdf/ur10.urdf:50-55,78-83,106-111,134-139,162-167,190-195 (joints), :218-222 (ee_fixed_joint -> ee_link)
_UR10 = dict(
    root=[((0, 0, 0), (0, 0, 0))],
    joints=[((0.0, 0.0, 0.1273), (0, 0, 0), 2), ((0.0, 0.220941, 0.0), (0.0, 1.570796325, 0.0), 1),
            ((0.0, -0.1719, 0.612), (0, 0, 0), 1), ((0.0, 0.0, 0.5723), (0.0, 1.570796325, 0.0), 1),
            ((0.0, 0.1149, 0.0), (0, 0, 0), 2), ((0.0, 0.0, 0.1157), (0, 0, 0), 1)],
    ee=[((0.0, 0.0922, 0.0), (-HALF_PI, 0.0, 0.0))], frame="ee_link")

# R/ur_description/urdf/ur5.urdf:274-282 (base_link -> base_link_inertia), :284-328 (joints, all about z),
# :346-357 (wrist_3-flange, flange-tool0); EE frame tool0 (R/prediction_model.py:28-29)
_UR5 = dict(
    root=[((0, 0, 0), (0, 0, 0))],
    joints=[((0, 0, 0.089159), (0, 0, 3.141592653589793), 2), ((0, 0, 0), (1.570796327, 0, 0), 2), ((-0.425, 0, 0), (0, 0, 0), 2),
            ((-0.39225, 0, 0.10915), (0, 0, 0), 2), ((0, -0.09465, -1.941303950897609e-11), (1.570796327, 0, 0), 2),
            ((0, 0.0823, -1.688001216681175e-11), (1.570796326589793, 3.141592653589793, 3.141592653589793), 2)],
    ee=[((0, 0, 0), (0, -HALF_PI, -HALF_PI)), ((0, 0, 0), (HALF_PI, 0, HALF_PI))], frame="tool0")

_ROBOTS = {"ur10": _UR10, "ur5": _UR5}


def chain_table(robot_name):
    """(xyzR [8,12], axis [6]) of a built-in robot ('ur10' frame ee_link, 'ur5' frame tool0 -- R/prediction_model.py:28-31).
    The arrays are built once per robot and shared (read-only): a grid search constructs thousands of models."""
    return _chain_table(str(robot_name).lower())


@functools.lru_cache(maxsize=None)
def _chain_table(name):
    try:
        r = _ROBOTS[name]
    except KeyError:
        raise ValueError(f"unknown robot {name!r}: built-in chains are {sorted(_ROBOTS)}; use chain_from_urdf() for others")
    rows = [_compose(r["root"])] + [_row(xyz, rpy) for xyz, rpy, _ in r["joints"]] + [_compose(r["ee"])]
    xyzR, axis = np.stack(rows), np.array([a for _, _, a in r["joints"]], dtype=np.int32)
    xyzR.setflags(write=False)
    axis.setflags(write=False)
    return xyzR, axis


def chain_from_urdf(path, ee_frame, base_link=None):
    """Read a serial 6R chain ending at link ``ee_frame`` from a URDF file: fixed joints are folded into the neighbouring
    transforms, revolute / continuous joints must have an axis along +x, +y or +z.  Returns (xyzR [8,12], axis [6])."""
    root = ET.parse(str(path)).getroot()
    by_child = {}
    for j in root.findall("joint"):
        o = j.find("origin")
        xyz = [float(v) for v in (o.get("xyz", "0 0 0") if o is not None else "0 0 0").split()]
        rpy = [float(v) for v in (o.get("rpy", "0 0 0") if o is not None else "0 0 0").split()]
        ax = j.find("axis")
        axis = [float(v) for v in ax.get("xyz").split()] if ax is not None else [1.0, 0.0, 0.0]
        by_child[j.find("child").get("link")] = dict(type=j.get("type"), parent=j.find("parent").get("link"), xyz=xyz, rpy=rpy, axis=axis)
    path_joints, link = [], ee_frame
    while link in by_child and link != base_link:
        path_joints.append(by_child[link])
        link = by_child[link]["parent"]
    path_joints.reverse()
    segs, axes, cur = [], [], []
    for j in path_joints:
        if j["type"] in ("revolute", "continuous"):
            a = np.asarray(j["axis"])
            k = int(np.argmax(np.abs(a)))
            if not np.allclose(a, np.eye(3)[k]):
                raise ValueError(f"joint axis {j['axis']} is not +x, +y or +z")
            segs.append(("joint", cur, (j["xyz"], j["rpy"])))
            axes.append(k)
            cur = []
        elif j["type"] == "fixed":
            cur.append((j["xyz"], j["rpy"]))
        else:
            raise ValueError(f"unsupported joint type {j['type']!r}")
    if len(axes) != 6:
        raise ValueError(f"expected 6 revolute joints between the base and {ee_frame!r}, found {len(axes)}")
    rows = [_compose(segs[0][1]) if segs[0][1] else _compose([((0, 0, 0), (0, 0, 0))])]
    for k, (_, pre, own) in enumerate(segs):
        rows.append(_row(*own) if (k == 0 or not pre) else _compose(pre + [own]))
    rows.append(_compose(cur) if cur else _compose([((0, 0, 0), (0, 0, 0))]))
    return np.stack(rows), np.array(axes, dtype=np.int32)


def _axis_map(actual, pattern):
    """Rotation C with C e_pattern = e_actual (identity when they agree)."""
    if actual == pattern:
        return np.eye(3)
    C = np.zeros((3, 3))
    third = 3 - actual - pattern
    C[actual, pattern] = 1.0
    C[pattern, actual] = -1.0
    C[third, third] = 1.0  # a transposition with one sign flipped: det = +1
    return C


def device_chain(xyzR, axis):
    """(Rfix [7,9], pfix [7,3]) for ``RmpcChain``: the table expressed in the kernels' fixed axis pattern.
    With C_j e_pattern_j = e_axis_j:  Rfix'_j = C_{j-1}' R_j C_j,  pfix'_j = C_{j-1}' t_j  (C_-1 = root rotation folded into
    joint 1), end-effector row = C_5' (R_ee, t_ee)."""
    xyzR = np.asarray(xyzR, float).reshape(8, 12)
    t = [xyzR[k, :3] for k in range(8)]
    R = [xyzR[k, 3:].reshape(3, 3) for k in range(8)]
    C = [_axis_map(int(axis[j]), PATTERN[j]) for j in range(6)]
    Rfix, pfix = np.zeros((7, 9)), np.zeros((7, 3))
    prev_R, prev_t = R[0], t[0]  # root transform, applied in front of joint 1
    for j in range(6):
        Rfix[j] = (prev_R @ R[1 + j] @ C[j]).ravel()
        pfix[j] = prev_t + prev_R @ t[1 + j] if j == 0 else prev_R @ t[1 + j]
        prev_R, prev_t = C[j].T, np.zeros(3)
    Rfix[6] = (prev_R @ R[7]).ravel()
    pfix[6] = prev_R @ t[7]
    return Rfix, pfix


def forward_kinematics(xyzR, axis, q):
    """Reference FK of a chain table in plain numpy (host-side checks): returns (p_ee [3], R_ee [3,3])."""
    xyzR = np.asarray(xyzR, float).reshape(8, 12)
    p, R = xyzR[0, :3].copy(), xyzR[0, 3:].reshape(3, 3).copy()
    for j in range(6):
        p = p + R @ xyzR[1 + j, :3]
        R = R @ xyzR[1 + j, 3:].reshape(3, 3)
        c, s = np.cos(q[j]), np.sin(q[j])
        k = int(axis[j])
        Rq = np.eye(3)
        a, b = [(1, 2), (2, 0), (0, 1)][k]
        Rq[a, a], Rq[a, b], Rq[b, a], Rq[b, b] = c, -s, s, c
        R = R @ Rq
    return p + R @ xyzR[7, :3], R @ xyzR[7, 3:].reshape(3, 3)
