"""ORACLE (test infrastructure only -- never imported by the product path).

CPU restatement of the reference's vehicle models, generic over the array
backend ``xp`` (``numpy`` for values, ``torch`` for exact fp64 autodiff
Jacobians/Hessians).  Follows, formula by formula:

* SingleTrack tyres / loads / ODE  : vehicle_control/models/dynamic_car.py:78-163
* SingleTrack spatial ODE          : vehicle_control/models/dynamic_car.py:170-187
* PointMass ODE (temporal/spatial) : vehicle_control/models/dynamic_point_mass.py:81-103
* RK4 / Euler one-step maps        : vehicle_control/utils/integrators.py:15-37
* parameters                       : config/models/dynamic_car.yaml

Parity status: the temporal SingleTrack RK4 plant step is PINNED by the 44
golden closed-loop trajectories of the reference (tests/golden, see
tests/test_oracle_golden.py); Jacobians are pinned only by torch autodiff of
this same restatement (no reference artefact stores them).
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np

G_GRAV = 9.88  # dynamic_car.py:61 (sic)


@dataclass(frozen=True)
class CarParams:
    """config/models/dynamic_car.yaml, flattened."""
    dt: float = 0.05
    l: float = 3.0
    m: float = 1700.0
    Izz: float = 3000.0
    a: float = 1.1
    b: float = 1.4
    h: float = 0.55
    eps: float = 0.9
    Peng: float = 170000.0
    Xdf: float = 1.0
    Xdr: float = 0.0
    Xbf: float = 0.78
    Xbr: float = 0.22
    Caf: float = 234000.0
    Car: float = 390000.0
    Cd: float = 0.4243
    muf: float = 0.9
    mur: float = 0.95
    theta: float = 0.0
    phi: float = 0.0
    Av2: float = 0.0
    Frr: float = 220.0

    @staticmethod
    def from_config(cfg) -> "CarParams":
        car, env = cfg["car"], cfg["env"]
        return CarParams(
            dt=float(cfg["dt"]), l=float(car["l"]), m=float(car["m"]), Izz=float(car["Izz"]),
            a=float(car["a"]), b=float(car["b"]), h=float(car["h"]), eps=float(car["eps"]),
            Peng=float(car["Peng"]), Xdf=float(car["Xd"]["f"]), Xdr=float(car["Xd"]["r"]),
            Xbf=float(car["Xb"]["f"]), Xbr=float(car["Xb"]["r"]),
            Caf=float(car["C_alpha"]["f"]), Car=float(car["C_alpha"]["r"]),
            Cd=float(env["Cd"]), muf=float(env["mu"]["f"]), mur=float(env["mu"]["r"]),
            theta=float(env["theta"]), phi=float(env["phi"]), Av2=float(env["Av2"]),
            Frr=float(env["Frr"]))


DEFAULT = CarParams()


# --------------------------------------------------------------------------- tyres / loads
def Xf(xp, p, Fx):      # dynamic_car.py:78
    return (p.Xdf - p.Xbf) / 2 * xp.tanh(2 * (Fx / 1000 + 0.5)) + (p.Xdf + p.Xbf) / 2


def Xr(xp, p, Fx):      # dynamic_car.py:83
    return (p.Xbr - p.Xdr) / 2 * xp.tanh(-2 * (Fx / 1000 + 0.5)) + (p.Xdr + p.Xbr) / 2


def Fx_f(xp, p, Fx):    # :80
    return Fx * Xf(xp, p, Fx)


def Fx_r(xp, p, Fx):    # :85
    return Fx * Xr(xp, p, Fx)


def Fz_f(xp, p, Ux, Fx):  # :98
    return (p.b / p.l) * p.m * (G_GRAV * math.cos(p.theta) * math.cos(p.phi) + p.Av2 * Ux ** 2) - p.h * Fx / p.l


def Fz_r(xp, p, Ux, Fx):  # :101
    return (p.a / p.l) * p.m * (G_GRAV * math.cos(p.theta) * math.cos(p.phi) + p.Av2 * Ux ** 2) + p.h * Fx / p.l


def Fymax_f(xp, p, Ux, Fx):  # :107
    return ((p.muf * Fz_f(xp, p, Ux, Fx)) ** 2 - (0.99 * Fx_f(xp, p, Fx)) ** 2) ** 0.5


def Fymax_r(xp, p, Ux, Fx):  # :108
    return ((p.mur * Fz_r(xp, p, Ux, Fx)) ** 2 - (0.99 * Fx_r(xp, p, Fx)) ** 2) ** 0.5


def alpha_f(xp, p, Ux, Uy, r, delta):  # :111
    return xp.arctan((Uy + p.a * r) / Ux) - delta


def alpha_r(xp, p, Ux, Uy, r, delta):  # :114
    return xp.arctan((Uy - p.b * r) / Ux)


def alphamod_f(xp, p, Ux, Fx):  # :119
    return xp.arctan(3 * Fymax_f(xp, p, Ux, Fx) * p.eps / p.Caf)


def alphamod_r(xp, p, Ux, Fx):  # :132
    return xp.arctan(3 * Fymax_r(xp, p, Ux, Fx) * p.eps / p.Car)


def _fiala(xp, C, eps, alpha, amod, Fymax):  # :121-128 / :134-141
    ta = xp.tan(alpha)
    inner = (-C * ta + C ** 2 * xp.abs(ta) * ta / (3 * Fymax)
             - (C ** 3 * ta ** 3) / (27 * Fymax ** 2))
    outer = -C * (1 - 2 * eps + eps ** 2) * ta - Fymax * (3 * eps ** 2 - 2 * eps ** 3) * xp.sign(alpha)
    return xp.where(xp.abs(alpha) <= amod, inner, outer)


def Fy_f(xp, p, Ux, Uy, r, delta, Fx):
    return _fiala(xp, p.Caf, p.eps, alpha_f(xp, p, Ux, Uy, r, delta),
                  alphamod_f(xp, p, Ux, Fx), Fymax_f(xp, p, Ux, Fx))


def Fy_r(xp, p, Ux, Uy, r, delta, Fx):
    return _fiala(xp, p.Car, p.eps, alpha_r(xp, p, Ux, Uy, r, delta),
                  alphamod_r(xp, p, Ux, Fx), Fymax_r(xp, p, Ux, Fx))


# --------------------------------------------------------------------------- SingleTrack ODE
def st_dot(xp, p, x, u, kappa):
    """Temporal ODE, dynamic_car.py:144-163.  x[...,8], u[...,2] -> tuple of 8."""
    Ux, Uy, r, delta, s, ey, epsi, t = (x[..., i] for i in range(8))
    Fx, w = u[..., 0], u[..., 1]
    fxf, fxr = Fx_f(xp, p, Fx), Fx_r(xp, p, Fx)
    fyf, fyr = Fy_f(xp, p, Ux, Uy, r, delta, Fx), Fy_r(xp, p, Ux, Uy, r, delta, Fx)
    Fb = 0.0
    Fd = p.Frr + p.Cd * Ux ** 2
    cd, sd = xp.cos(delta), xp.sin(delta)
    Ux_dot = (fxf * cd - fyf * sd + fxr - Fd) / p.m + r * Uy
    Uy_dot = (fyf * cd + fxf * sd + fyr + Fb) / p.m - r * Ux
    r_dot = (p.a * (fyf * cd + fxf * sd) - p.b * fyr) / p.Izz
    s_dot = (Ux * xp.cos(epsi) - Uy * xp.sin(epsi)) / (1 - kappa * ey)
    ey_dot = Ux * xp.sin(epsi) + Uy * xp.cos(epsi)
    epsi_dot = r - kappa * s_dot
    one = s_dot * 0 + 1
    return (Ux_dot, Uy_dot, r_dot, w + 0 * one, s_dot, ey_dot, epsi_dot, one)


def st_f_temporal(xp, p, x, u, kappa):
    return xp.stack(st_dot(xp, p, x, u, kappa), -1)


def st_f_spatial(xp, p, x, u, kappa):
    """dynamic_car.py:170-187."""
    d = st_dot(xp, p, x, u, kappa)
    sd = d[4]
    return xp.stack((d[0] / sd, d[1] / sd, d[2] / sd, d[3] / sd, d[7], d[5] / sd, d[6] / sd, 1 / sd), -1)


def rk4(f, x, h):
    """integrators.py:30-34 (same operation order)."""
    k1 = f(x)
    k2 = f(x + 0.5 * h * k1)
    k3 = f(x + 0.5 * h * k2)
    k4 = f(x + h * k3)
    return x + h * (1 / 6) * (k1 + 2 * k2 + 2 * k3 + k4)


def _col(h):
    return h[..., None] if hasattr(h, "shape") and getattr(h, "ndim", 0) > 0 else h


def st_step_temporal(xp, p, x, u, kappa, dt):
    return rk4(lambda z: st_f_temporal(xp, p, z, u, kappa), x, _col(dt))


def st_step_spatial(xp, p, x, u, kappa, ds):
    return rk4(lambda z: st_f_spatial(xp, p, z, u, kappa), x, _col(ds))


# --------------------------------------------------------------------------- PointMass
def pm_dot(xp, p, x, u, kappa):
    """dynamic_point_mass.py:81-87."""
    V, s, ey, epsi, t = (x[..., i] for i in range(5))
    Fx, Fy = u[..., 0], u[..., 1]
    Fd = p.Frr + p.Cd * V ** 2
    V_dot = (Fx - Fd) / p.m
    s_dot = (V * xp.cos(epsi)) / (1 - kappa * ey)
    ey_dot = V * xp.sin(epsi)
    epsi_dot = (Fy + 0.0) / (p.m * V) - kappa * s_dot
    one = s_dot * 0 + 1
    return (V_dot, s_dot, ey_dot, epsi_dot, one)


def pm_step_temporal(xp, p, x, u, kappa, dt):
    return x + _col(dt) * xp.stack(pm_dot(xp, p, x, u, kappa), -1)


def pm_step_spatial(xp, p, x, u, kappa, ds):
    """dynamic_point_mass.py:93-103 + integrators.py:17-23."""
    d = pm_dot(xp, p, x, u, kappa)
    sd = d[1]
    f = xp.stack((d[0] / sd, d[4], d[2] / sd, d[3] / sd, d[4] / sd), -1)
    return x + _col(ds) * f


# --------------------------------------------------------------------------- rollouts
def st_rollout_spatial(p, x0, U, kappa, ds):
    """numpy: x0[B,8], U[B,N,2], kappa[B,N], ds[B,N] -> X[B,N+1,8]."""
    B, N = U.shape[0], U.shape[1]
    X = np.empty((B, N + 1, 8))
    X[:, 0] = x0
    for n in range(N):
        X[:, n + 1] = st_step_spatial(np, p, X[:, n], U[:, n], kappa[:, n], ds[:, n])
    return X


def pm_rollout_spatial(p, x0, U, kappa, ds):
    B, M = U.shape[0], U.shape[1]
    X = np.empty((B, M + 1, 5))
    X[:, 0] = x0
    for n in range(M):
        X[:, n + 1] = pm_step_spatial(np, p, X[:, n], U[:, n], kappa[:, n], ds[:, n])
    return X


def st_step_jac(p, x, u, kappa, ds):
    """[A|B] (B,8,10) of the spatial RK4 step by torch forward-mode autodiff (fp64)."""
    import torch
    from torch.func import jacfwd, vmap
    xt = torch.as_tensor(x, dtype=torch.float64)
    ut = torch.as_tensor(u, dtype=torch.float64)
    kt = torch.as_tensor(kappa, dtype=torch.float64)
    ht = torch.as_tensor(ds, dtype=torch.float64)

    def one(z, k, h):
        return st_step_spatial(torch, p, z[:8], z[8:], k, h)
    J = vmap(jacfwd(one))(torch.cat([xt, ut], -1), kt, ht)
    return J.numpy()


def pm_step_jac(p, x, u, kappa, ds):
    import torch
    from torch.func import jacfwd, vmap
    xt = torch.as_tensor(x, dtype=torch.float64)
    ut = torch.as_tensor(u, dtype=torch.float64)

    def one(z, k, h):
        return pm_step_spatial(torch, p, z[:5], z[5:], k, h)
    J = vmap(jacfwd(one))(torch.cat([xt, ut], -1), torch.as_tensor(kappa, dtype=torch.float64),
                          torch.as_tensor(ds, dtype=torch.float64))
    return J.numpy()
