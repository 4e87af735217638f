"""SingleTrack model on the B200 engine (mirror of vehicle_control/models/dynamic_car.py:10-290 and
models/racing_car.py:15-52): named state/action vectors, ``drive`` (closed-loop plant step), ``transition`` /
``spatial_transition`` one-step maps and the tyre helper functions, all evaluated by the CUDA kernels."""
from __future__ import annotations

import numpy as np

from ..utils.common_utils import as_config
from ..utils.fancy_vector import FancyVector


class DynamicCarAction(FancyVector):
    _keys = ["Fx", "w"]
    _labels = [r"$F_x$", r"$\omega$"]

    def __init__(self, Fx=0.0, w=0.0):
        super().__init__(Fx, w)


class DynamicCarState(FancyVector):
    _keys = ["Ux", "Uy", "r", "delta", "s", "ey", "epsi", "t"]
    _labels = [r"$U_x$", r"$U_y$", r"$r$", r"$\delta$", r"$s$", r"$e_y$", r"$e_\psi$", r"$t$"]

    def __init__(self, Ux=0.0, Uy=0.0, r=0.0, delta=0.0, s=0.0, ey=0.0, epsi=0.0, t=0.0):
        super().__init__(Ux, Uy, r, delta, s, ey, epsi, t)


class _DM:
    """Stand-in for the CasADi DM the reference's Functions return: ``.full()`` gives a column."""

    def __init__(self, a):
        self._a = np.asarray(a, dtype=float)

    def full(self):
        return self._a.reshape(-1, 1)

    def __float__(self):
        return float(self._a.reshape(-1)[0])

    def __array__(self, dtype=None, copy=None):
        return self._a if dtype is None else self._a.astype(dtype)

    def __repr__(self):
        return f"DM({self._a.reshape(-1)})"


class RacingCar:
    def __init__(self, config, track, engine=None):
        from ..engine import Engine
        self.config = as_config(config)
        self.length = self.config.car.l
        self.track = track
        self.dt = self.config.dt
        self.engine = engine if engine is not None else Engine(self.config, None, track)
        self.state = self.create_state()
        self.input = self.create_action()

    def drive(self, input):
        """racing_car.py:34-46: kappa = k(s); next = RK4_t(state, input, kappa, dt) on the device."""
        e = self.engine
        x = e.tensor(self.state.values[None, :8].copy())
        e.plant_step(x, e.tensor(np.asarray(input.values, dtype=float)[None].copy()), self.dt)
        self.state = self.create_state(*x.cpu().numpy()[0])
        self.input = input
        return self.state

    def rel2glob(self, state):
        s, ey, epsi = (state[self.state.index(k)] for k in ("s", "ey", "epsi"))
        return self.track.rel2glob(s, ey, epsi)


class DynamicCar(RacingCar):
    @classmethod
    def create_state(cls, *args, **kwargs):
        return DynamicCarState(*args, **kwargs)

    @classmethod
    def create_action(cls, *args, **kwargs):
        return DynamicCarAction(*args, **kwargs)

    # ---- one-step maps (dynamic_car.py:166-167, 188-199)
    def _step(self, x, u, curvature, h, spatial):
        e = self.engine
        x = np.asarray(x, dtype=float).reshape(1, 8)
        u = np.asarray(u, dtype=float).reshape(1, 1, 2)
        if spatial:
            X = e.rollout_st(e.tensor(x), e.tensor(u), e.tensor(np.array([[float(curvature)]])),
                             e.tensor(np.array([[float(h)]])))
            return _DM(X.cpu().numpy()[0, 1])
        # temporal: the plant kernel evaluates k(s) itself; a foreign curvature needs the generic path
        k_here = float(self.track.k(x[0, 4]))
        if abs(k_here - float(curvature)) > 0.0:
            raise ValueError("transition(): curvature must be track.k(s) (that is what the plant kernel evaluates)")
        xt = e.tensor(x.copy())
        e.plant_step(xt, e.tensor(u.reshape(1, 2)), float(h))
        return _DM(xt.cpu().numpy()[0])

    def transition(self, x, u, curvature, dt):
        return self._step(x, u, curvature, dt, False)

    def spatial_transition(self, x, u, curvature, ds):
        return self._step(x, u, curvature, ds, True)

    # ---- tyre helpers (dynamic_car.py:79-142)
    def _tyres(self, Ux, Uy, r, delta, Fx):
        e = self.engine
        x = np.zeros((1, 8)); x[0, :4] = [Ux, Uy, r, delta]
        return e.tyre_forces(e.tensor(x), e.tensor(np.array([[float(Fx), 0.0]]))).cpu().numpy()[0]

    def Fx_f(self, Fx):
        return _DM(self._tyres(10.0, 0, 0, 0, Fx)[0])

    def Fx_r(self, Fx):
        return _DM(self._tyres(10.0, 0, 0, 0, Fx)[1])

    def Fy_f(self, Ux, Uy, r, delta, Fx):
        return _DM(self._tyres(Ux, Uy, r, delta, Fx)[2])

    def Fy_r(self, Ux, Uy, r, delta, Fx):
        return _DM(self._tyres(Ux, Uy, r, delta, Fx)[3])

    def alpha_f(self, Ux, Uy, r, delta):
        return _DM(self._tyres(Ux, Uy, r, delta, 0.0)[4])

    def alpha_r(self, Ux, Uy, r, delta):
        return _DM(self._tyres(Ux, Uy, r, delta, 0.0)[5])

    def print(self, state, input):
        Ux, Uy, r, delta = (float(v) for v in list(state)[:4])
        f = self._tyres(Ux, Uy, r, delta, float(list(input)[0]))
        for n, v in zip(("Fx_f", "Fx_r", "Fy_f", "Fy_r"), f[:4]):
            print(f"{n}: {v:.6g}")
