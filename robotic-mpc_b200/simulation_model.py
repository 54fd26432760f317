"""Plant logs of one instance -- host-side mirror of the read surface of ``R/simulation_model.py`` (class ``Robot``).

The plant itself (``zdot = [qd; -W qd + W u]``, RK4, ``R/simulation_model.py:79-117``) is integrated on the device
inside the closed-loop kernel; this class only holds the per-instance views of the device logs in the reference's
array layout (time along the LAST axis, ``R/simulation_model.py:24-29``)."""
from __future__ import annotations

import numpy as np


class Robot:
    def __init__(self, z0, u0, Nsim, dt, wcv, integration_method="RK4"):
        if integration_method != "RK4":
            raise NotImplementedError("the device plant integrates with RK4 (the method the reference's Simulator uses)")
        self._dt, self._N = float(dt), int(Nsim)
        self._z0, self._u0 = np.asarray(z0, float), np.asarray(u0, float)
        self.wcv = np.diag(np.asarray(wcv, float))
        self._logs = None  # allocated on first access: a grid search builds thousands of these and most only get summaries
        self._integration_method_name = integration_method

    def _alloc(self):
        if self._logs is None:
            n = self._N + 1
            self._logs = {"z": np.zeros((12, n)), "u": np.zeros((6, n)), "_ee_rpy_log": np.zeros((3, n)),
                          "_ee_pose_log": np.zeros((12, n)), "_ee_velocity_log": np.zeros((6, n))}
            self._logs["z"][:, 0] = self._z0
            self._logs["u"][:, 0] = self._u0
        return self._logs

    def _log_get(name):  # noqa: N805 -- property factory
        return property(lambda self: self._alloc()[name], lambda self, v: self._alloc().__setitem__(name, v))

    z = _log_get("z")
    u = _log_get("u")
    _ee_rpy_log = _log_get("_ee_rpy_log")
    _ee_pose_log = _log_get("_ee_pose_log")
    _ee_velocity_log = _log_get("_ee_velocity_log")
    del _log_get

    def _load(self, zlog, ulog, plog, RK):
        """zlog [Nsim+1,12], ulog [Nsim,6], plog [Nsim+1,RK_COUNT] of this instance."""
        self.z = np.ascontiguousarray(zlog.T)
        self.u[:, 1:] = ulog.T  # u[:, t+1] = control applied at step t (R/simulation_model.py:88)
        self._ee_pose_log = np.ascontiguousarray(plog[:, RK["POSE"]:RK["POSE"] + 12].T)
        self._ee_rpy_log = np.ascontiguousarray(plog[:, RK["RPY"]:RK["RPY"] + 3].T)
        self._ee_velocity_log = np.ascontiguousarray(plog[:, RK["TWIST"]:RK["TWIST"] + 6].T)

    dt = property(lambda self: self._dt)
    N = property(lambda self: self._N)

    def simulate_output(self):
        return self._ee_pose_log, self._ee_velocity_log

    def get_state(self):
        return self.z[:, -1]

    def get_control_input(self):
        return self.u[:, -1]

    def ee_position(self, t):
        return self._ee_pose_log[:3, t]

    def ee_orientation_euler(self, t):
        return self._ee_rpy_log[:, t]

    def ee_orientation_rot(self, t):
        return self._ee_pose_log[3:12, t]

    def ee_velocity(self, t):
        return self._ee_velocity_log[:, t]

    def joint_angles(self, t):
        return self.z[:6, t]

    def joint_velocities(self, t):
        return self.z[6:, t]

    def state(self, t):
        return self.z[:, t]

    def ee_pose(self, t):
        return self._ee_pose_log[:, t]
