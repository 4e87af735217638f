"""ORACLE (test infrastructure only -- never imported by the product path).

CPU restatement of the NLP that the reference's CascadedMPC builds with CasADi Opti
(vehicle_control/controllers/mpc/cascaded_mpc.py): decision variables, objective, equality and
inequality constraints, written with torch fp64 so that exact first and second derivatives come
from torch.func autodiff (the role CasADi's AD plays in the reference, cascaded_mpc.py:65).

  variables      _init_variables        :68-89   (PointMass columns keep only their 5 used rows)
  objective      _stage_cost            :130-179, _pm_stage_cost :204-239, _switching_cost :241-254,
                 _terminal_cost         :279-304
  equalities     initial state :26-28, dynamics :113-120 / :194-202, _switching_constraints :256-277
  inequalities   _stage_constraints     :91-128, _pm_stage_constraints :181-192
  parameters     _init_horizon          :316-338

Parity status: the transcription is PINNED by the cold-start known answers of the reference's saved
runs (tick 0 -> action_traj[1]; tests/test_oracle_nmpc.py); objective values are unpinned by any
reference artefact (none stores them).
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np
import torch

from . import models as om

torch.set_default_dtype(torch.float64)


@dataclass
class MpcCfg:
    """config/controllers/*.yaml flattened."""
    N: int = 20
    M: int = 35
    mpc_dt: float = 0.03
    ds_pm: float = 3.0
    obstacles: bool = False
    w_time: float = 5.0
    w_speed: float = 1.0
    w_ey: float = 1.0
    w_epsi: float = 2.0
    w_w: float = 4.0
    w_Fx: float = 1e-6
    w_Fy: float = 1e-6
    w_dev_st: float = 1e-3
    w_dev_pm: float = 1e-3
    w_boundary: float = 10.0
    w_slip: float = 1500.0
    w_obstacles: float = 10.0
    w_switch_F: float = 1e-5
    w_min: float = -0.4
    w_max: float = 0.4
    Ux_min: float = 3.0
    max_speed: float = 20.0
    delta_min: float = -0.45
    delta_max: float = 0.45
    ey_min: float = -3.0
    ey_max: float = 3.0
    V_min: float = 3.0
    ey_min_pm: float = -3.0
    ey_max_pm: float = 3.0

    @staticmethod
    def from_config(c) -> "MpcCfg":
        cw, ic, sc = c["cost_weights"], c["input_constraints"], c["state_constraints"]
        pc = c.get("state_pm_constraints") or {}
        return MpcCfg(
            N=int(c["horizon"]), M=int(c["horizon_pm"]), mpc_dt=float(c["mpc_dt"]), ds_pm=float(c["ds_pm"]),
            obstacles=bool(c.get("obstacles", False)),
            w_time=float(cw["time"]), w_speed=float(cw["speed"]), w_ey=float(cw["ey"]), w_epsi=float(cw["epsi"]),
            w_w=float(cw["w"]), w_Fx=float(cw["Fx"]), w_Fy=float(cw.get("Fy", 0.0)),
            w_dev_st=float(cw["deviation_st"]), w_dev_pm=float(cw.get("deviation_pm", 0.0)),
            w_boundary=float(cw["boundary"]), w_slip=float(cw["slip"]), w_obstacles=float(cw.get("obstacles", 0.0)),
            w_switch_F=float(cw.get("switch_F", 0.0)),
            w_min=float(ic["w_min"]), w_max=float(ic["w_max"]), Ux_min=float(sc["Ux_min"]),
            max_speed=float(sc["max_speed"]), delta_min=float(sc["delta_min"]), delta_max=float(sc["delta_max"]),
            ey_min=float(sc["ey_min"]), ey_max=float(sc["ey_max"]), V_min=float(pc.get("V_min", sc["Ux_min"])),
            ey_min_pm=float(pc.get("ey_min", sc["ey_min"])), ey_max_pm=float(pc.get("ey_max", sc["ey_max"])))

    @property
    def H(self):
        return self.N + self.M


def horizon_params(cfg: MpcCfg, track, x0, Ux_prev):
    """_init_horizon :316-338 -> (ds[H], kappa[H])."""
    N, M = cfg.N, cfg.M
    ds_st = cfg.mpc_dt * np.asarray(Ux_prev[:N], dtype=float)
    s_st = np.cumsum(ds_st) - ds_st[0] + x0[4]
    ds_pm = np.full(M, cfg.ds_pm)
    s_pm = np.cumsum(ds_pm) - ds_st[-1] + s_st[-1]
    return np.concatenate([ds_st, ds_pm]), np.concatenate([track.k(s_st), track.k(s_pm)])


def cold_start(cfg: MpcCfg, rng_uniform_2xH):
    """_init_variables :72-76: states = 1 (Ux = 4 on the SingleTrack columns), actions = 1 + U[0,1)."""
    X = np.ones((8, cfg.H))
    X[0, :cfg.N] += 3
    return X, np.ones((2, cfg.H)) + rng_uniform_2xH


class NLP:
    """One NMPC instance.  w = [Xst(8xN) row-major | Xpm(5xM) row-major | U(2xH) row-major]."""

    def __init__(self, cfg: MpcCfg, car: om.CarParams, x0, kappa, ds, obstacles=None, bound_relax=1e-8):
        self.cfg, self.p = cfg, car
        self.N, self.M, self.H = cfg.N, cfg.M, cfg.H
        self.x0 = torch.as_tensor(np.asarray(x0, dtype=float))
        self.kappa = torch.as_tensor(np.asarray(kappa, dtype=float))
        self.ds = torch.as_tensor(np.asarray(ds, dtype=float))
        self.obst = None
        if cfg.obstacles and obstacles is not None and len(obstacles):
            self.obst = torch.as_tensor(np.asarray(obstacles, dtype=float).reshape(-1, 3))
        self.relax = bound_relax
        self.n = 8 * self.N + 5 * self.M + 2 * self.H

    # ------------------------------------------------------------ packing
    def pack(self, X, U):
        """X[8,H], U[2,H] (reference layout) -> w."""
        X, U = np.asarray(X, dtype=float), np.asarray(U, dtype=float)
        return np.concatenate([X[:, :self.N].ravel(), X[:5, self.N:].ravel(), U.ravel()])

    def unpack(self, w):
        N, M, H = self.N, self.M, self.H
        w = np.asarray(w)
        X = np.ones((8, H))
        X[:, :N] = w[:8 * N].reshape(8, N)
        X[:5, N:] = w[8 * N:8 * N + 5 * M].reshape(5, M)
        return X, w[8 * N + 5 * M:].reshape(2, H)

    def _split(self, w):
        N, M, H = self.N, self.M, self.H
        return w[:8 * N].reshape(8, N), w[8 * N:8 * N + 5 * M].reshape(5, M), w[8 * N + 5 * M:].reshape(2, H)

    # ------------------------------------------------------------ objective
    def _obstacle_cost(self, s, ey, ds):
        if self.obst is None:
            return 0.0
        so, eo, ro = self.obst[:, 0:1], self.obst[:, 1:2], self.obst[:, 2:3]
        dist = torch.sqrt((s[None] - so) ** 2 + (ey[None] - eo) ** 2)
        gap = dist - (ro + 0.1)
        # inside the pole the term is negative and unbounded: treated as an evaluation error (NaN -> backtrack)
        gap = torch.where(gap > 0, gap, torch.full_like(gap, float("nan")))
        return (self.cfg.w_obstacles * ds[None] / gap).sum()

    def f(self, w):
        c, p, N, M, H = self.cfg, self.p, self.N, self.M, self.H
        Xs, Xp, U = self._split(w)
        ds = self.ds
        Ux, Uy, r, delta, s, ey, epsi, t = Xs
        Fx, wst = U[0, :N], U[1, :N]
        d = ds[:N]
        zero = torch.zeros(())
        J = (torch.where(ey < c.ey_min, c.w_boundary * d * (ey - c.ey_min) ** 2, zero).sum()
             + torch.where(ey > c.ey_max, c.w_boundary * d * (ey - c.ey_max) ** 2, zero).sum()
             + (c.w_dev_st * d * ey ** 2).sum() + (c.w_w * wst ** 2).sum())
        taf = torch.abs(torch.tan(om.alpha_f(torch, p, Ux, Uy, r, delta)))
        tmf = torch.tan(om.alphamod_f(torch, p, Ux, Fx))
        J = J + torch.where(taf >= tmf, c.w_slip * (taf - tmf) ** 2, zero).sum()
        tar = torch.abs(torch.tan(om.alpha_r(torch, p, Ux, Uy, r, delta)))
        tmr = torch.tan(om.alphamod_r(torch, p, Ux, Fx))
        J = J + torch.where(tar >= tmr, c.w_slip * (tar - tmr) ** 2, zero).sum()
        if N > 1:
            J = J + ((c.w_Fx / d[:-1]) * (Fx[1:] - Fx[:-1]) ** 2).sum()
        J = J + self._obstacle_cost(s, ey, d)
        if M > 0:
            V, sp, eyp, epp, tp = Xp
            Fxp, Fyp = U[0, N:], U[1, N:]
            dp = ds[N:]
            J = J + (torch.where(eyp < c.ey_min_pm, c.w_boundary * dp * (eyp - c.ey_min_pm) ** 2, zero).sum()
                     + torch.where(eyp > c.ey_max_pm, c.w_boundary * dp * (eyp - c.ey_max_pm) ** 2, zero).sum()
                     + (c.w_dev_pm * dp * eyp ** 2).sum())
            if M > 1:
                J = J + ((1.0 / dp[:-1]) * (c.w_Fy * (Fyp[1:] - Fyp[:-1]) ** 2
                                            + c.w_Fx * (Fxp[1:] - Fxp[:-1]) ** 2)).sum()
            J = J + self._obstacle_cost(sp, eyp, dp)
            # switching cost :241-254
            fyf = om.Fy_f(torch, p, Ux[-1], Uy[-1], r[-1], delta[-1], Fx[-1])
            fyr = om.Fy_r(torch, p, Ux[-1], Uy[-1], r[-1], delta[-1], Fx[-1])
            J = J + c.w_switch_F * (1.0 / ds[N - 1]) * ((Fxp[0] - Fx[-1]) ** 2 + (Fyp[0] - fyf - fyr) ** 2)
            vT, tT, eyT, epT = V[-1], tp[-1], eyp[-1], epp[-1]
        else:
            vT, tT, eyT, epT = Ux[-1], t[-1], ey[-1], epsi[-1]
        J = J + torch.where(vT >= c.max_speed, c.w_speed * (vT - c.max_speed) ** 2, zero)
        J = J + c.w_time * tT + c.w_ey * eyT ** 2 + c.w_epsi * epT ** 2
        return J

    # ------------------------------------------------------------ equalities g(w) = 0
    def g(self, w):
        p, N, M = self.p, self.N, self.M
        Xs, Xp, U = self._split(w)
        out = [Xs[:, 0] - self.x0]
        if N > 1:
            nxt = om.st_step_spatial(torch, p, Xs[:, :-1].T, U[:, :N - 1].T, self.kappa[:N - 1], self.ds[:N - 1])
            out.append((Xs[:, 1:].T - nxt).reshape(-1))
        if M > 0:
            Ux, Uy, s, ey, epsi, t = Xs[0, -1], Xs[1, -1], Xs[4, -1], Xs[5, -1], Xs[6, -1], Xs[7, -1]
            out.append(torch.stack([Xp[0, 0] - (Ux ** 2 + Uy ** 2) ** 0.5, Xp[2, 0] - ey,
                                    Xp[3, 0] - (torch.arctan(Uy / Ux) + epsi), Xp[1, 0] - s, Xp[4, 0] - t]))
            if M > 1:
                nxt = om.pm_step_spatial(torch, p, Xp[:, :-1].T, U[:, N:N + M - 1].T, self.kappa[N:N + M - 1],
                                         self.ds[N:N + M - 1])
                out.append((Xp[:, 1:].T - nxt).reshape(-1))
        return torch.cat(out)

    # ------------------------------------------------------------ inequalities c(w) >= 0
    def c(self, w):
        c, p, N, M = self.cfg, self.p, self.N, self.M
        Xs, Xp, U = self._split(w)
        Ux, Uy, r, delta = Xs[0], Xs[1], Xs[2], Xs[3]
        Fx, wst = U[0, :N], U[1, :N]
        rl = self.relax

        def lo(b):  # relaxed lower / upper bounds (IPOPT bound_relax_factor)
            return b - rl * max(1.0, abs(b))

        def hi(b):
            return b + rl * max(1.0, abs(b))
        bf = p.muf * om.Fz_f(torch, p, Ux, Fx) * torch.cos(om.alpha_f(torch, p, Ux, Uy, r, delta))
        br = p.mur * om.Fz_r(torch, p, Ux, Fx) * torch.cos(om.alpha_r(torch, p, Ux, Uy, r, delta))
        fxf, fxr = om.Fx_f(torch, p, Fx), om.Fx_r(torch, p, Fx)
        rows = [Ux - lo(c.Ux_min), delta - lo(c.delta_min), hi(c.delta_max) - delta, p.Peng / Ux - Fx + rl,
                wst - lo(c.w_min), hi(c.w_max) - wst, bf - fxf + rl, fxf + bf + rl, br - fxr + rl, fxr + br + rl]
        out = [torch.stack(rows, 1).reshape(-1)]
        if M > 0:
            V, Fxp = Xp[0], U[0, N:]
            out.append(torch.stack([V - lo(c.V_min), p.Peng / V - Fxp + rl], 1).reshape(-1))
        return torch.cat(out)


    # ------------------------------------------------------------ restoration substitute
    def repair(self, w_np):
        """Guess-independent, dynamically feasible re-initialisation used where IPOPT would enter its
        restoration phase: hold the current speed (Fx = drag), no steering rate, PointMass lateral force
        that follows the centre-line curvature; states re-generated by rolling the models out from x0."""
        p, N, M = self.p, self.N, self.M
        X, U = self.unpack(w_np)
        X, U = X.copy(), U.copy()
        kap, ds = self.kappa.numpy(), self.ds.numpy()
        X[:, 0] = self.x0.numpy()
        for n in range(N):
            U[0, n] = p.Frr + p.Cd * X[0, n] ** 2
            U[1, n] = 0.0
            if n < N - 1:
                X[:, n + 1] = om.st_step_spatial(np, p, X[:, n][None], U[:, n][None], kap[n], ds[n])[0]
        if M > 0:
            xl = X[:, N - 1]
            X[:5, N] = [np.hypot(xl[0], xl[1]), xl[4], xl[5], np.arctan(xl[1] / xl[0]) + xl[6], xl[7]]
            for m in range(N, N + M):
                U[0, m] = p.Frr + p.Cd * X[0, m] ** 2
                U[1, m] = p.m * X[0, m] ** 2 * kap[m]
                if m < N + M - 1:
                    X[:5, m + 1] = om.pm_step_spatial(np, p, X[:5, m][None], U[:, m][None], kap[m], ds[m])[0]
        # positions guessed inside an obstacle (pole of the reciprocal term; plus a margin) are moved sideways out
        if self.obst is not None:
            for k in range(1, N + M):
                si, ei = (4, 5) if k < N else (1, 2)
                for so, eo, ro in self.obst.numpy():
                    R, a, e = ro + 0.1 + 0.5, X[si, k] - so, X[ei, k] - eo
                    if a * a + e * e < R * R:
                        X[ei, k] = eo + (1.0 if e > 1e-9 else -1.0) * np.sqrt(R * R - a * a)   # |e| <= 1e-9: dead centre
        return self.pack(X, U)

    # ------------------------------------------------------------ derivatives (exact, autodiff)
    def eval_all(self, w_np, lam_g, lam_c, obj_scale=1.0):
        """Returns f, grad f, g, Jg, c, Jc, W = hess(obj_scale*f + lam_g.g - lam_c.c)."""
        from torch.func import grad, hessian, jacfwd
        w = torch.as_tensor(np.asarray(w_np, dtype=float))
        lg, lc = torch.as_tensor(lam_g), torch.as_tensor(lam_c)
        fv = self.f(w)
        gf = grad(self.f)(w)
        gv, Jg = self.g(w), jacfwd(self.g)(w)
        cv, Jc = self.c(w), jacfwd(self.c)(w)

        def lag(v):
            return obj_scale * self.f(v) + (lg * self.g(v)).sum() - (lc * self.c(v)).sum()
        W = hessian(lag)(w)
        return (float(fv), gf.numpy(), gv.numpy(), Jg.numpy(), cv.numpy(), Jc.numpy(), W.numpy())

    def values(self, w_np):
        w = torch.as_tensor(np.asarray(w_np, dtype=float))
        with torch.no_grad():
            return float(self.f(w)), self.g(w).numpy(), self.c(w).numpy()

    def grad_f(self, w_np):
        from torch.func import grad
        return grad(self.f)(torch.as_tensor(np.asarray(w_np, dtype=float))).numpy()
