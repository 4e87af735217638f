"""ORACLE (test infrastructure only -- never imported by the product path).

CPU restatement of the reference's track geometry with numpy/scipy, following
vehicle_control/environment/track.py:
  _construct_path :254-296, _construct_spline :206-252, get_curvature :109-119,
  get_orientation :121-129, _precompute_curvatures :156-167, rel2glob :102-107.
CasADi's ``interpolant("bspline")`` (cubic, not-a-knot) is restated with
``scipy.interpolate.make_interp_spline(k=3)``.

Parity status: PINNED -- kappa(s) through the golden plant steps and
rel2glob through the golden ``*_preds.npy`` (tests/test_oracle_golden.py).
"""
from __future__ import annotations

import numpy as np
from scipy.integrate import trapezoid
from scipy.interpolate import InterpolatedUnivariateSpline, make_interp_spline


def _wrap(a):  # common_utils.py:29-37
    a = np.asarray(a, dtype=float)
    return np.where(a < -np.pi, 2 * np.pi + a, np.where(a > np.pi, a - 2 * np.pi, a))


class OracleTrack:
    def __init__(self, cfg):
        self.name = cfg["name"]
        self.width = cfg["width"]
        res, smooth = cfg["resolution"], int(cfg["smoothing"])
        # --- _construct_path :254-296
        wx, wy = [], []
        corners = cfg["corners"]
        for i in range(len(corners) - 1):
            s0, e0 = corners[i], corners[i + 1]
            dist = np.sqrt((e0[0] - s0[0]) ** 2 + (e0[1] - s0[1]) ** 2)
            n = int(dist / res)
            wx.extend(np.linspace(s0[0], e0[0], n, endpoint=False).tolist())
            wy.extend(np.linspace(s0[1], e0[1], n, endpoint=False).tolist())
        xs, ys = [], []
        for i in range(len(wx)):
            if i < smooth or i > len(wx) - smooth - 1:
                xs.append(wx[i]); ys.append(wy[i])
            else:
                xs.append(np.mean(wx[i - smooth:i + smooth + 1]))
                ys.append(np.mean(wy[i - smooth:i + smooth + 1]))
        # closing point is appended then dropped by _construct_waypoints (:316)
        self.n_waypoints = len(xs)
        u = np.arange(self.n_waypoints)
        # --- _construct_spline :206-252
        xv = InterpolatedUnivariateSpline(u, xs, k=3, ext=3)(u)
        yv = InterpolatedUnivariateSpline(u, ys, k=3, ext=3)(u)
        self._xs = make_interp_spline(u, xv, k=3)
        self._ys = make_interp_spline(u, yv, k=3)
        seg = np.sqrt(self._xs(u, 1) ** 2 + self._ys(u, 1) ** 2)
        self.length = float(trapezoid(seg, u, dx=0.1))
        self._scale = self.n_waypoints / self.length
        # --- _precompute_curvatures :156-167
        self.ds = 0.05
        self.s_table = np.arange(0, self.length - 0.1, self.ds)
        self.kappa_table = self.get_curvature(self.s_table)
        self._k = make_interp_spline(self.s_table, self.kappa_table, k=3)
        self.obstacles = []
        for (s, ey, rad) in cfg.get("obstacle_data", []) or []:
            x, y, _ = self.rel2glob(s, ey, 0.0)
            self.obstacles.append((float(s), float(ey), float(rad), float(x), float(y)))

    # x(s) = x_spline(s/L*n) and its s-derivatives (chain rule) :241-251
    def _d(self, spl, s, nu):
        return spl(np.asarray(s, dtype=float) * self._scale, nu, extrapolate=True) * self._scale ** nu

    def x(self, s):
        return self._d(self._xs, s, 0)

    def y(self, s):
        return self._d(self._ys, s, 0)

    def get_curvature(self, s):
        s = np.fmod(s, self.length)
        dx, dy = self._d(self._xs, s, 1), self._d(self._ys, s, 1)
        ddx, ddy = self._d(self._xs, s, 2), self._d(self._ys, s, 2)
        return np.abs(dx * ddy - ddx * dy) / np.power(dx ** 2 + dy ** 2, 1.5)

    def get_orientation(self, s):
        s = np.fmod(s, self.length)
        dx, dy = self._d(self._xs, s, 1), self._d(self._ys, s, 1)
        mag = (dx ** 2 + dy ** 2) ** 0.5
        return np.arctan2(dy / mag, dx / mag)

    def k(self, s):
        """kappa spline, no fmod (cascaded_mpc.py:330, racing_car.py:38)."""
        return self._k(np.asarray(s, dtype=float), extrapolate=True)

    def rel2glob(self, s, ey, epsi):
        o = self.get_orientation(s)
        return self.x(s) - np.sin(o) * ey, self.y(s) + np.cos(o) * ey, _wrap(o + epsi)
