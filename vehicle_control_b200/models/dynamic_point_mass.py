"""PointMass model (mirror of vehicle_control/models/dynamic_point_mass.py:10-188): curvilinear point mass with
explicit-Euler one-step maps; the NMPC kernels use it for the second phase of the cascaded horizon."""
from __future__ import annotations

import numpy as np

from ..utils.fancy_vector import FancyVector
from .dynamic_car import _DM, RacingCar


class DynamicPointMassAction(FancyVector):
    _keys = ["Fx", "Fy"]
    _labels = [r"$F_x$", r"$F_y$"]

    def __init__(self, Fx=0.0, Fy=0.0):
        super().__init__(Fx, Fy)


class DynamicPointMassState(FancyVector):
    _keys = ["V", "s", "ey", "epsi", "t"]
    _labels = [r"$V$", r"$s$", r"$e_y$", r"$e_\psi$", r"$t$"]

    def __init__(self, V=0.0, s=0.0, ey=0.0, epsi=0.0, t=0.0):
        super().__init__(V, s, ey, epsi, t)


class DynamicPointMass(RacingCar):
    @classmethod
    def create_state(cls, *args, **kwargs):
        return DynamicPointMassState(*args, **kwargs)

    @classmethod
    def create_action(cls, *args, **kwargs):
        return DynamicPointMassAction(*args, **kwargs)

    def spatial_transition(self, x, u, curvature, ds):
        """dynamic_point_mass.py:93-103 (explicit Euler in s)."""
        e = self.engine
        X = e.rollout_pm(e.tensor(np.asarray(x, dtype=float).reshape(1, 5)),
                         e.tensor(np.asarray(u, dtype=float).reshape(1, 1, 2)),
                         e.tensor(np.array([[float(curvature)]])), e.tensor(np.array([[float(ds)]])))
        return _DM(X.cpu().numpy()[0, 1])

    def drive(self, input):
        raise NotImplementedError("the PointMass is a planning model only; the plant is the SingleTrack (racing.py:43-49)")
