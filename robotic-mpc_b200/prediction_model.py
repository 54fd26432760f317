"""Prediction-model constants -- host-side mirror of the data ``R/prediction_model.py`` (class ``SixDofRobot``) hands to
the OCP: sample time, joint bandwidths, the exact-ZOH matrices ``Ad, Bd`` (``R/prediction_model.py:87-115``) and the
task-frame offset (``:9``).  The symbolic FK / Jacobian / output functions of the reference are hand-written device
code (``csrc/rmpc_device.cuh: ur10_task``) over a constant chain table: 'ur10' (frame ``ee_link``) and 'ur5' (frame ``tool0``)
are built in (``R/prediction_model.py:28-31``), any other serial 6R arm comes in through ``UrdfLoader`` / ``robots.chain_from_urdf``."""
from __future__ import annotations

import math

import numpy as np

from . import robots


class UrdfLoader:
    """Mirror of ``R/loader.py:5-36`` without pinocchio: resolves the chain table of a robot.  ``UrdfLoader('ur5')`` /
    ``('ur10')`` use the built-in tables; ``UrdfLoader(name, urdf_path=..., ee_frame=...)`` reads a serial 6R chain from a URDF
    file (axes along +x / +y / +z).  Meshes, collision and visual models are presentation and are not loaded."""

    def __init__(self, robot_name, urdf_path=None, ee_frame=None, base_link=None):
        self.name = robot_name
        if urdf_path is None:
            self.chain = robots.chain_table(robot_name)
            self.ee_frame = {"ur5": "tool0", "ur10": "ee_link"}[str(robot_name).lower()]
        else:
            if ee_frame is None:
                raise ValueError("UrdfLoader: ee_frame is needed with urdf_path")
            self.chain = robots.chain_from_urdf(urdf_path, ee_frame, base_link=base_link)
            self.ee_frame = ee_frame


class SixDofRobot:
    def __init__(self, urdf_loader=None, Ts=0.01, Wcv=None, translation_ee_t=(0.0, 0.0, 0.1), robot_name="ur10"):
        if urdf_loader is None:
            urdf_loader = UrdfLoader(robot_name)  # ValueError for a name without a built-in table
        self.robot_name = urdf_loader.name
        xyzR, axis = urdf_loader.chain
        self.chain = (np.ascontiguousarray(xyzR, dtype=float).reshape(8, 12), np.ascontiguousarray(axis, dtype=np.int32).reshape(6))
        self._chain_key = None
        self.n_dof = 6
        self.Ts = float(Ts)
        self.Wcv = np.full(6, 200.0) if Wcv is None else np.asarray(Wcv, dtype=float).reshape(6)
        self.translation = np.asarray(translation_ee_t, dtype=float).reshape(3)
        self._update_discrete_lti_matrices()

    def _update_discrete_lti_matrices(self):
        w, Ts = self.Wcv, self.Ts
        a22 = np.array([math.exp(-x * Ts) for x in w])
        a12 = (1.0 - a22) / w
        k = np.arange(6)
        Ad, Bd = np.zeros((12, 12)), np.zeros((12, 6))
        Ad[k, k] = 1.0
        Ad[k, 6 + k] = a12
        Ad[6 + k, 6 + k] = a22
        Bd[k, k] = Ts - a12
        Bd[6 + k, k] = 1.0 - a22
        self.Ad, self.Bd = Ad, Bd

    def chain_key(self):
        """Hashable identity of the kinematic chain (instances of one batched launch share it)."""
        if self._chain_key is None:
            self._chain_key = (self.chain[0].tobytes(), self.chain[1].tobytes())
        return self._chain_key

    def forward_kinematics(self, q):
        """(p, R) of the end-effector frame at joint angles q -- host-side check of the table (R/prediction_model.py:126-132)."""
        return robots.forward_kinematics(self.chain[0], self.chain[1], np.asarray(q, float))
