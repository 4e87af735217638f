"""Track geometry on the host + the spline tables the device kernels evaluate.

Mirror of vehicle_control/environment/track.py (Track :83-100, rel2glob :102-107, get_curvature
:109-119, get_orientation :121-129, _construct_obstacles :131-138, _precompute_curvatures :156-167,
_construct_spline :206-252, _construct_path :254-296).  Construction is one-time host work (as in the
reference); per-tick lookups -- k(s) along every horizon, rel2glob of every prediction -- run on the
GPU from the coefficient tables returned by ``device_tables()`` (csrc/vc_kernels.cu: k_horizon_params,
k_plant_step, k_rel2glob).  CasADi's cubic "bspline" interpolant is the not-a-knot interpolating
spline, built here with scipy.
"""
from __future__ import annotations

from typing import List

import numpy as np
from scipy.integrate import trapezoid
from scipy.interpolate import make_interp_spline

from ..utils.common_utils import as_config, wrap


class Obstacle:
    """Circular obstacle, stored in curvilinear (s, ey) and world (cx, cy) coordinates (track.py:55-80)."""

    def __init__(self, cx, cy, s, ey, radius):
        self.cx, self.cy, self.s, self.ey, self.radius = float(cx), float(cy), float(s), float(ey), float(radius)

    def __repr__(self) -> str:
        return f"Obstacle(cx={self.cx}, cy={self.cy}, radius={self.radius})"


class Track:
    CURVATURE_STEP = 0.05   # track.py:157

    def __init__(self, config):
        config = as_config(config)
        self.name = config.name
        self.width = config.width
        self.resolution = config.resolution
        self.smoothing = int(config.smoothing)
        wx, wy = self._smoothed_waypoints(config.corners)
        self.waypoints_xy = np.stack([wx, wy], 1)
        self.n_waypoints = len(wx)
        self._fit_centerline(wx, wy)
        self._fit_curvature()
        self.obstacles: List[Obstacle] = []
        for s, ey, radius in (config.get("obstacle_data") or []):
            x, y, _ = self.rel2glob(s, ey, 0.0)
            self.obstacles.append(Obstacle(x, y, s, ey, radius))

    # ---------------------------------------------------------------- construction
    def _smoothed_waypoints(self, corners):
        """Corner polygon -> waypoints every `resolution` m -> centred moving average of half-width
        `smoothing` away from the two ends (track.py:254-296; the closing point is dropped again by
        _construct_waypoints :316)."""
        xs, ys = [], []
        for p, q in zip(corners[:-1], corners[1:]):
            count = int(np.hypot(q[0] - p[0], q[1] - p[1]) / self.resolution)
            xs.append(np.linspace(p[0], q[0], count, endpoint=False))
            ys.append(np.linspace(p[1], q[1], count, endpoint=False))
        xs, ys = np.concatenate(xs), np.concatenate(ys)
        w, n = self.smoothing, len(xs)
        out = []
        for v in (xs, ys):
            sm = v.copy()
            # same summation as np.mean over each window (pairwise sums differ in the last bits from
            # a cumulative-sum trick, and the golden trajectories resolve those bits on the curves)
            for i in range(w, n - w):
                sm[i] = np.mean(v[i - w:i + w + 1])
            out.append(sm)
        return out[0], out[1]

    def _fit_centerline(self, wx, wy):
        u = np.arange(self.n_waypoints, dtype=float)
        # the reference fits an interpolating spline, samples it AT the integer sites (which returns
        # the waypoints up to rounding) and interpolates those samples again (track.py:216-229)
        from scipy.interpolate import InterpolatedUnivariateSpline
        xv = InterpolatedUnivariateSpline(u, wx, k=3, ext=3)(u)
        yv = InterpolatedUnivariateSpline(u, wy, k=3, ext=3)(u)
        self._x_of_u = make_interp_spline(u, xv, k=3)
        self._y_of_u = make_interp_spline(u, yv, k=3)
        speed = np.sqrt(self._x_of_u(u, 1) ** 2 + self._y_of_u(u, 1) ** 2)   # not hypot: last-bit parity of `length`
        self.length = float(trapezoid(speed, u))
        self._u_per_s = self.n_waypoints / self.length

    def _fit_curvature(self):
        self.ds = self.CURVATURE_STEP
        self.s_table = np.arange(0, self.length - 0.1, self.ds)
        self.curvature_table = self.get_curvature(self.s_table)
        self._k_spline = make_interp_spline(self.s_table, self.curvature_table, k=3)

    # ---------------------------------------------------------------- host evaluation (setup / single queries)
    def _xy_deriv(self, s, order):
        u = np.asarray(s, dtype=float) * self._u_per_s
        f = self._u_per_s ** order
        return self._x_of_u(u, order, extrapolate=True) * f, self._y_of_u(u, order, extrapolate=True) * f

    def x(self, s):
        return self._xy_deriv(s, 0)[0]

    def y(self, s):
        return self._xy_deriv(s, 0)[1]

    def get_curvature(self, s):
        s = np.fmod(np.asarray(s, dtype=float), self.length)
        dx, dy = self._xy_deriv(s, 1)
        ddx, ddy = self._xy_deriv(s, 2)
        return np.abs(dx * ddy - ddx * dy) / np.power(dx * dx + dy * dy, 1.5)

    def get_orientation(self, s):
        s = np.fmod(np.asarray(s, dtype=float), self.length)
        dx, dy = self._xy_deriv(s, 1)
        mag = np.sqrt(dx * dx + dy * dy)
        return np.arctan2(dy / mag, dx / mag)

    def k(self, s):
        """Curvature spline; like the reference it is NOT wrapped with fmod (cascaded_mpc.py:330)."""
        out = self._k_spline(np.asarray(s, dtype=float), extrapolate=True)
        return out if out.ndim else float(out)

    def rel2glob(self, s, ey, epsi):
        th = self.get_orientation(s)
        return self.x(s) - np.sin(th) * ey, self.y(s) + np.cos(th) * ey, wrap(th + epsi)

    # ---------------------------------------------------------------- device export
    def device_tables(self):
        """Coefficient tables for vc_set_track (include/vc_api.h)."""
        return {
            "k_coefs": np.ascontiguousarray(self._k_spline.c, dtype=np.float64),
            "k_x0": float(self.s_table[0]),
            "k_step": float(self.ds),
            "x_coefs": np.ascontiguousarray(self._x_of_u.c, dtype=np.float64),
            "y_coefs": np.ascontiguousarray(self._y_of_u.c, dtype=np.float64),
            "length": self.length,
        }

    def obstacle_array(self):
        """[n_obs,3] (s, ey, radius)."""
        return np.array([[o.s, o.ey, o.radius] for o in self.obstacles], dtype=np.float64).reshape(-1, 3)
