"""Prediction-model constants -- host-side mirror of the data ``R/prediction_model.py`` (class ``SixDofRobot``) hands to
the OCP: sample time, joint bandwidths, the exact-ZOH matrices ``Ad, Bd`` (``R/prediction_model.py:87-115``) and the
task-frame offset (``:9``).  The symbolic FK / Jacobian / output functions of the reference are hand-written device
code (``csrc/rmpc_device.cuh: ur10_task``); only ``robot_name == 'ur10'`` (frame ``ee_link``) is built in."""
from __future__ import annotations

import math

import numpy as np


class SixDofRobot:
    def __init__(self, urdf_loader=None, Ts=0.01, Wcv=None, translation_ee_t=(0.0, 0.0, 0.1), robot_name="ur10"):
        name = getattr(urdf_loader, "name", robot_name) if urdf_loader is not None else robot_name
        if str(name).lower() != "ur10":
            raise NotImplementedError(f"robot {name!r}: only the UR10 chain is compiled into the kernels (SURVEY.md 8(f) #4)")
        self.n_dof = 6
        self.Ts = float(Ts)
        self.Wcv = np.full(6, 200.0) if Wcv is None else np.asarray(Wcv, dtype=float).reshape(6)
        self.translation = np.asarray(translation_ee_t, dtype=float).reshape(3)
        self._update_discrete_lti_matrices()

    def _update_discrete_lti_matrices(self):
        w, Ts = self.Wcv, self.Ts
        a22 = np.array([math.exp(-x * Ts) for x in w])
        a12 = (1.0 - a22) / w
        I, Z = np.eye(6), np.zeros((6, 6))
        self.Ad = np.block([[I, np.diag(a12)], [Z, np.diag(a22)]])
        self.Bd = np.vstack([Ts * I - np.diag(a12), np.diag(1.0 - a22)])
