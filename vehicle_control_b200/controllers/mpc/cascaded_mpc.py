"""NMPC controllers on the B200 engine.

``CascadedMPC`` keeps the surface of the reference class (vehicle_control/controllers/mpc/cascaded_mpc.py:16-381:
constructor (car, point_mass, config), ``command(state) -> DynamicCarAction``, ``state_prediction (8,H)``,
``action_prediction (2,H)``, ``get_state_prediction() -> (H,3)``); ``horizon_pm: 0`` gives the plain
singletrack controller (config/controllers/singletrack.yaml).  ``BatchedCascadedMPC`` is the same controller
for B independent vehicles / problem instances with everything resident on the device.

Per tick (reference :306-338): parameters ds, kappa from the previous prediction (k_horizon_params), warm start
= previous solution un-shifted, then one batched interior-point solve (vc_nmpc_solve).  No CasADi, no IPOPT.
"""
from __future__ import annotations

import os

import numpy as np
import torch

from ...engine import STATUS_NAMES, Engine
from ...models.dynamic_car import DynamicCarAction
from ...utils.common_utils import as_config
from ..controller import Controller

np.random.seed(31)   # cascaded_mpc.py:13 -- the cold guess of the actions is 1 + U[0,1) drawn at construction


def cold_guess(N, H, rng_draw=None):
    """_init_variables :72-76."""
    X = np.ones((8, H))
    X[0, :N] += 3
    U = np.ones((2, H)) + (np.random.random((2, H)) if rng_draw is None else rng_draw)
    return X, U


class BatchedCascadedMPC(Controller):
    def __init__(self, car, point_mass, config, batch: int, engine: Engine = None, obstacles=None):
        self.config = as_config(config)
        self.car, self.point_mass = car, point_mass
        self.N, self.M = int(self.config.horizon), int(self.config.horizon_pm)
        self.H = self.N + self.M
        self.B = int(batch)
        self.warm_rungs = int(os.environ.get("VC_WARM_RUNGS", self.config.get("warm_rungs", 0) or 0))
        # IPOPT's feasibility-restoration phase is not restated.  An instance whose filter line
        # search fails is re-solved from the previous solution shifted by one stage, then from the cold guess
        # (`retry_failed: false` in the controller config = one attempt, failure reported as the reference would raise)
        self.retry_failed = bool(int(os.environ.get("VC_RETRY_FAILED", int(self.config.get("retry_failed", True)))))
        self.retries = {"shifted": 0, "cold": 0, "unrecovered": 0}
        self.track_id = None
        self._ticks = 0
        self.engine = engine if engine is not None else Engine(car.config, self.config, car.track)
        if engine is not None:
            self.engine.set_config(self.config)
        e = self.engine
        X, U = cold_guess(self.N, self.H)
        self._cold = (e.tensor(X[None]), e.tensor(U[None]))
        self.state_prediction = e.tensor(np.repeat(X[None], self.B, 0))      # [B,8,H]
        self.action_prediction = e.tensor(np.repeat(U[None], self.B, 0))     # [B,2,H]
        self.ds, self.kappa = e.empty(self.B, self.H), e.empty(self.B, self.H)
        self.info = {"obj": e.empty(self.B), "status": e.empty(self.B, dtype=torch.int32),
                     "iters": e.empty(self.B, dtype=torch.int32), "kkt": e.empty(self.B)}
        self.obstacles = None
        if self.config.get("obstacles"):
            ob = car.track.obstacle_array() if obstacles is None else np.asarray(obstacles, dtype=float)
            if ob.ndim == 2:
                ob = np.repeat(ob[None], self.B, 0)
            self.obstacles = e.tensor(ob) if ob.shape[1] else None

    def set_tracks(self, tracks, track_id):
        """Per-instance tracks: instance b drives on tracks[track_id[b]]."""
        self.engine.set_track_atlas(tracks)
        self.track_id = torch.as_tensor(np.asarray(track_id), dtype=torch.int32).to(self.engine.device).contiguous()

    def command(self, states: torch.Tensor, active: torch.Tensor = None) -> torch.Tensor:
        """states [B,8] (device) -> first actions [B,2] (device view of action_prediction[:, :, 0]).
        `active` [B] bool: the instances to solve (the lock-step simulator passes the vehicles still racing); the others
        keep their prediction and are not touched -- the active ones are compacted into a sub-batch, solved and scattered
        back, so that finished vehicles cost nothing and cannot hold the lock-step passes."""
        e = self.engine
        # first tick: cold guess, barrier ladder from mu_init; later ticks are warm starts and may enter the ladder
        # `warm_rungs` rungs down (controller config key, default 0 = the reference's IPOPT behaviour)
        rungs = self.warm_rungs if self._ticks > 0 else 0
        if e.cfg.warm_rungs != rungs:
            e.cfg.warm_rungs = rungs
            e.set_config(e.cfg)
        self._ticks += 1
        if active is None or bool(active.all()):
            self._solve(states, self.state_prediction, self.action_prediction, self.ds, self.kappa, self.info,
                        self.obstacles, self.track_id)
        else:
            idx = active.nonzero().squeeze(1)
            if idx.numel() == 0:
                return self.action_prediction[:, :, 0].contiguous()
            sub = lambda t: None if t is None else t[idx].contiguous()
            X, U, ds, kap = sub(self.state_prediction), sub(self.action_prediction), sub(self.ds), sub(self.kappa)
            info = {k: v[idx].contiguous() for k, v in self.info.items()}
            self._solve(sub(states), X, U, ds, kap, info, sub(self.obstacles), sub(self.track_id))
            self.state_prediction[idx], self.action_prediction[idx] = X, U
            self.ds[idx], self.kappa[idx] = ds, kap
            for k in self.info:
                self.info[k][idx] = info[k]
        return self.action_prediction[:, :, 0].contiguous()

    def _solve(self, states, X, U, ds, kap, info, obstacles, track_id):
        """_init_horizon + solve (+ the retries of failed instances) of one batch, in place in the given tensors."""
        e = self.engine
        if self.retry_failed:
            Xw, Uw = X.clone(), U.clone()
        e.horizon_params(states, X, ds, kap, track_id=track_id)
        e.solve(states, kap, ds, X, U, obstacles=obstacles, out=info)
        if self.retry_failed and bool((info["status"] > 1).any()):
            self._retry(states, Xw, Uw, X, U, ds, kap, info, obstacles, track_id)

    def _retry(self, states, Xw, Uw, Xs, Us, dss, kaps, info, obstacles, track_id):
        """Second and third attempt for the instances that failed, as a compacted sub-batch."""
        e, N = self.engine, self.N
        for kind in ("shifted", "cold"):
            idx = (info["status"] > 1).nonzero().squeeze(1)
            if idx.numel() == 0:
                return
            if kind == "shifted":    # SingleTrack columns advance by one stage; the PointMass tail stays
                X, U = Xw[idx].clone(), Uw[idx].clone()
                X[:, :, :N - 1] = Xw[idx][:, :, 1:N]
                U[:, :, :N - 1] = Uw[idx][:, :, 1:N]
            else:
                X = self._cold[0].repeat(idx.numel(), 1, 1)
                U = self._cold[1].repeat(idx.numel(), 1, 1)
            x = states[idx].contiguous()
            ds, kap = e.horizon_params(x, X, track_id=None if track_id is None else track_id[idx].contiguous())
            ob = obstacles[idx].contiguous() if obstacles is not None else None
            out = e.solve(x, kap, ds, X, U, obstacles=ob)
            ok = out["status"] <= 1
            self.retries[kind] += int(ok.sum())
            good = idx[ok]
            Xs[good], Us[good] = X[ok], U[ok]
            dss[good], kaps[good] = ds[ok], kap[ok]
            for k in info:
                info[k][good] = out[k][ok]
        bad = info["status"] > 1
        self.retries["unrecovered"] += int(bad.sum())
        # nothing converged for these: keep the warm start they had (not the failed iterate) for the next tick
        Xs[bad], Us[bad] = Xw[bad], Uw[bad]

    def get_state_prediction(self) -> torch.Tensor:
        """[B,H,3] global (x, y, psi) of every predicted state (:340-352)."""
        X, N = self.state_prediction, self.N
        s = torch.cat([X[:, 4, :N], X[:, 1, N:]], 1).contiguous()
        ey = torch.cat([X[:, 5, :N], X[:, 2, N:]], 1).contiguous()
        ep = torch.cat([X[:, 6, :N], X[:, 3, N:]], 1).contiguous()
        return self.engine.rel2glob(s, ey, ep)


class CascadedMPC(Controller):
    """Single-vehicle drop-in (B = 1) with the reference's attribute surface."""

    def __init__(self, car, point_mass, config, engine: Engine = None):
        self._batched = BatchedCascadedMPC(car, point_mass, config, 1, engine=engine)
        self.config = self._batched.config
        self.car, self.point_mass = car, point_mass
        self.N, self.M, self.H = self._batched.N, self._batched.M, self._batched.H
        self.ns, self.na, self.ns_pm, self.na_pm = 8, 2, 5, 2
        self.dt = self.config.mpc_dt
        self.last_info = None

    @property
    def state_prediction(self) -> np.ndarray:
        return self._batched.state_prediction[0].cpu().numpy()

    @state_prediction.setter
    def state_prediction(self, v):
        self._batched.state_prediction.copy_(self._batched.engine.tensor(np.asarray(v, dtype=float)[None]))

    @property
    def action_prediction(self) -> np.ndarray:
        return self._batched.action_prediction[0].cpu().numpy()

    @action_prediction.setter
    def action_prediction(self, v):
        self._batched.action_prediction.copy_(self._batched.engine.tensor(np.asarray(v, dtype=float)[None]))

    def command(self, state):
        x = np.asarray(getattr(state, "values", state), dtype=float).reshape(1, 8)
        u = self._batched.command(self._batched.engine.tensor(x)).cpu().numpy()[0]
        info = {k: v.cpu().numpy()[0] for k, v in self._batched.info.items()}
        self.last_info = info
        if int(info["status"]) > 1:   # the reference raises (CasADi RuntimeError) when IPOPT fails
            raise RuntimeError(f"NMPC solve failed: {STATUS_NAMES.get(int(info['status']), info['status'])}")
        return DynamicCarAction(Fx=float(u[0]), w=float(u[1]))

    def get_state_prediction(self) -> np.ndarray:
        return self._batched.get_state_prediction()[0].cpu().numpy()
