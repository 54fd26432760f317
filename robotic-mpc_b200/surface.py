"""Quadratic target surface -- host-side mirror of the part of ``R/surface.py`` the OCP uses.

``S(x, y) = a x^2 + b y^2 + c x y + d x + e y + f`` (defaults ``R/surface.py:14-17``) and its unit normal
``n = [S_x, S_y, -1] / |.|`` (``R/surface.py:243-262``).  Position / orientation / limits are stored because the
reference's ``Simulator`` passes them, but -- as in the reference (``R/trajectory_optimizer.py:104-120``) -- they do
not enter the OCP.  The device evaluates the same polynomial from the six coefficients (``RP_SURF`` parameter slot).
"""
from __future__ import annotations

import numpy as np

COEFF_KEYS = ("a", "b", "c", "d", "e", "f")
DEFAULT_COEFFS = {"a": -0.15, "b": 0.15, "c": -0.01, "d": 0.01, "e": 0.01, "f": 0.0}


class Surface:
    def __init__(self, position=None, orientation_rpy=None, limits=None, coefficients=None):
        self.position = None if position is None else np.asarray(position, dtype=float)
        self.orientation_rpy = None if orientation_rpy is None else np.asarray(orientation_rpy, dtype=float)
        self.limits = limits
        self.coeffs = dict(DEFAULT_COEFFS)
        if coefficients:
            unknown = set(coefficients) - set(COEFF_KEYS)
            if unknown:
                raise KeyError(f"unknown surface coefficient(s) {sorted(unknown)}; expected a subset of {COEFF_KEYS}")
            self.coeffs.update({k: float(v) for k, v in coefficients.items()})

    # same accessor names as the reference
    def get_position(self):
        return self.position

    def get_orientation_rpy(self):
        return self.orientation_rpy

    def get_limits(self):
        return self.limits

    def coeff_vector(self) -> np.ndarray:
        """[a, b, c, d, e, f] in the order of the engine's parameter block."""
        return np.array([self.coeffs[k] for k in COEFF_KEYS], dtype=np.float64)

    def get_surface_function(self):
        a, b, c, d, e, f = self.coeff_vector()
        return lambda x, y: a * x * x + b * y * y + c * x * y + d * x + e * y + f

    def get_normal_vector(self, x, y):
        a, b, c, d, e, _ = self.coeff_vector()
        n = np.array([2 * a * x + c * y + d, 2 * b * y + c * x + e, -1.0])
        return n / np.linalg.norm(n)

    def get_normal_vector_casadi(self):  # name kept for drop-in compatibility; returns a plain callable
        return self.get_normal_vector
