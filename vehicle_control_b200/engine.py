"""Batched device engine: thin torch-tensor front end of the C-ABI (include/vc_api.h).

torch is used only as the owner of device memory / streams; every computation happens in the
hand-written CUDA kernels of libvc_b200.so.  There is no CPU path: constructing an Engine without a
CUDA device raises.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Optional

import numpy as np
import torch

from . import _capi
from ._capi import MpcConfig, ModelParams, check
from .utils.common_utils import as_config

STATUS_NAMES = {0: "solved", 1: "acceptable", 2: "max_iter", 3: "eval_error", 4: "linesearch_fail"}


def model_params_from(cfg) -> ModelParams:
    """config/models/dynamic_car.yaml -> vc_model_params (dynamic_car.py:61-102 naming)."""
    cfg = as_config(cfg)
    car, env = cfg.car, cfg.env
    p = ModelParams()
    p.dt = float(cfg.dt)
    for k in ("l", "m", "Izz", "a", "b", "h", "eps", "Peng"):
        setattr(p, k, float(car[k]))
    p.Xdf, p.Xdr, p.Xbf, p.Xbr = float(car.Xd.f), float(car.Xd.r), float(car.Xb.f), float(car.Xb.r)
    p.Caf, p.Car = float(car.C_alpha.f), float(car.C_alpha.r)
    p.Cd, p.muf, p.mur = float(env.Cd), float(env.mu.f), float(env.mu.r)
    p.Av2, p.Frr = float(env.Av2), float(env.Frr)
    p.g_eff = 9.88 * math.cos(float(env.theta)) * math.cos(float(env.phi))
    return p


def mpc_config_from(cfg, **overrides) -> MpcConfig:
    """config/controllers/*.yaml -> vc_mpc_config (keys read by cascaded_mpc.py)."""
    cfg = as_config(cfg)
    cw, ic, sc = cfg.cost_weights, cfg.input_constraints, cfg.state_constraints
    pc = cfg.get("state_pm_constraints") or {}
    c = MpcConfig()
    _capi.load().vc_default_config(C.byref(c))   # IPOPT defaults of the reference (cascaded_mpc.py:55-65) + engine policies
    c.N, c.M = int(cfg.horizon), int(cfg.horizon_pm)
    c.obstacles = 1 if cfg.get("obstacles") else 0
    c.mpc_dt, c.ds_pm = float(cfg.mpc_dt), float(cfg.ds_pm)
    c.w_time, c.w_speed, c.w_ey, c.w_epsi = float(cw.time), float(cw.speed), float(cw.ey), float(cw.epsi)
    c.w_w, c.w_Fx, c.w_Fy = float(cw.w), float(cw.Fx), float(cw.get("Fy", 0.0))
    c.w_dev_st, c.w_dev_pm = float(cw.deviation_st), float(cw.get("deviation_pm", 0.0))
    c.w_boundary, c.w_slip = float(cw.boundary), float(cw.slip)
    c.w_obstacles, c.w_switch_F = float(cw.get("obstacles", 0.0)), float(cw.get("switch_F", 0.0))
    c.w_min, c.w_max = float(ic.w_min), float(ic.w_max)
    c.Ux_min, c.max_speed = float(sc.Ux_min), float(sc.max_speed)
    c.delta_min, c.delta_max = float(sc.delta_min), float(sc.delta_max)
    c.ey_min, c.ey_max = float(sc.ey_min), float(sc.ey_max)
    c.V_min = float(pc.get("V_min", sc.Ux_min))
    c.ey_min_pm, c.ey_max_pm = float(pc.get("ey_min", sc.ey_min)), float(pc.get("ey_max", sc.ey_max))
    # solver options may be overridden from the controller yaml ("solver: {max_iter: ..., max_soc: ...}") or by keyword
    for k, v in dict(cfg.get("solver") or {}).items():
        setattr(c, k, type(getattr(c, k))(v))
    for k, v in overrides.items():
        setattr(c, k, v)
    return c


class Engine:
    def __init__(self, car_config, mpc_config=None, track=None, device: Optional[int] = None, stream=None):
        self.lib = _capi.load()
        if not torch.cuda.is_available():
            raise RuntimeError("vehicle_control_b200.Engine needs a CUDA device: there is no CPU fallback")
        self.device_index = torch.cuda.current_device() if device is None else int(device)
        self.device = torch.device("cuda", self.device_index)
        self.params = car_config if isinstance(car_config, ModelParams) else model_params_from(car_config)
        self.cfg = None
        h = C.c_void_p()
        check(self.lib.vc_create(C.byref(self.params), None, self.device_index, stream, C.byref(h)), "vc_create")
        self._h = h
        self.track = None
        if mpc_config is not None:
            self.set_config(mpc_config)
        if track is not None:
            self.set_track(track)

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            try:
                self.lib.vc_destroy(h)
            except Exception:
                pass
            self._h = None

    # ------------------------------------------------------------------ setup
    def set_config(self, mpc_config, **overrides):
        self.cfg = mpc_config if isinstance(mpc_config, MpcConfig) else mpc_config_from(mpc_config, **overrides)
        check(self.lib.vc_set_config(self._h, C.byref(self.cfg)), "vc_set_config")

    def set_track(self, track):
        t = track.device_tables()
        kc, xc, yc = t["k_coefs"], t["x_coefs"], t["y_coefs"]
        check(self.lib.vc_set_track(self._h, kc.ctypes.data, len(kc), t["k_x0"], t["k_step"], xc.ctypes.data,
                                    yc.ctypes.data, len(xc), t["length"]), "vc_set_track")
        self.track = track

    def set_track_atlas(self, tracks):
        """Per-instance tracks: upload the curvature tables of several tracks; instance b then names its track by
        `track_id[b]` in horizon_params / plant_step (include/vc_api.h: vc_set_track_atlas)."""
        tabs = [t.device_tables() for t in tracks]
        coefs = np.ascontiguousarray(np.concatenate([t["k_coefs"] for t in tabs]), dtype=np.float64)
        off = np.ascontiguousarray(np.cumsum([0] + [len(t["k_coefs"]) for t in tabs]), dtype=np.int32)
        x0 = np.ascontiguousarray([t["k_x0"] for t in tabs], dtype=np.float64)
        step = np.ascontiguousarray([t["k_step"] for t in tabs], dtype=np.float64)
        check(self.lib.vc_set_track_atlas(self._h, len(tabs), coefs.ctypes.data, off.ctypes.data, x0.ctypes.data,
                                          step.ctypes.data), "vc_set_track_atlas")
        self.tracks = list(tracks)

    def build_tracks(self, track_configs):
        """Track construction ON THE DEVICE for a list of track configs (corners / resolution / smoothing, the keys of
        config/environment/*.yaml): waypoints, smoothing, centre-line splines, length, curvature table and Track.k of every
        track (include/vc_api.h: vc_build_tracks).  The built tracks become the atlas: track_id = position in the list.
        Returns one dict per track: length, n_waypoints, n_k."""
        cfgs = [as_config(c) for c in track_configs]
        T = len(cfgs)
        mc = max(len(c.corners) for c in cfgs)
        corners = np.zeros((T, mc, 2))
        for t, c in enumerate(cfgs):
            corners[t, :len(c.corners)] = np.asarray([[float(p[0]), float(p[1])] for p in c.corners])
        nc = np.ascontiguousarray([len(c.corners) for c in cfgs], dtype=np.int32)
        res = np.ascontiguousarray([float(c.resolution) for c in cfgs], dtype=np.float64)
        sm = np.ascontiguousarray([int(c.smoothing) for c in cfgs], dtype=np.int32)
        check(self.lib.vc_build_tracks(self._h, T, mc, corners.ctypes.data, nc.ctypes.data, res.ctypes.data, sm.ctypes.data),
              "vc_build_tracks")
        out = []
        for t in range(T):
            L, nw, nk = C.c_double(), C.c_int(), C.c_int()
            check(self.lib.vc_built_track_info(self._h, t, C.byref(L), C.byref(nw), C.byref(nk)), "vc_built_track_info")
            out.append({"length": L.value, "n_waypoints": nw.value, "n_k": nk.value})
        self.built_tracks = out
        return out

    BUILT_TABLES = {"k_coefs": 0, "waypoints_x": 1, "waypoints_y": 2, "x_coefs": 3, "y_coefs": 4, "curvature_table": 5}

    def built_track_table(self, t, name):
        info = self.built_tracks[t]
        n = info["n_k"] if name in ("k_coefs", "curvature_table") else info["n_waypoints"]
        out = np.empty(n)
        check(self.lib.vc_built_track_read(self._h, int(t), self.BUILT_TABLES[name], out.ctypes.data, n), "vc_built_track_read")
        return out

    def built_track_rel2glob(self, t, s, ey):
        """(s, ey) -> (cx, cy) on built track t (obstacle placement, track.py:131-138)."""
        out = self.empty(s.numel(), 2)
        check(self.lib.vc_built_track_rel2glob(self._h, int(t), s.numel(), self._p(s), self._p(ey), self._p(out)),
              "vc_built_track_rel2glob")
        return out

    @property
    def H(self):
        return self.cfg.N + self.cfg.M

    # ------------------------------------------------------------------ helpers
    def tensor(self, a, dtype=torch.float64):
        return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype).to(self.device).contiguous()

    def _p(self, t, shape=None, dtype=torch.float64):
        if t is None:
            return None
        if not (isinstance(t, torch.Tensor) and t.is_cuda and t.dtype == dtype and t.is_contiguous()
                and t.device.index == self.device_index):
            raise TypeError(f"expected a contiguous {dtype} tensor on {self.device}")
        if shape is not None and tuple(t.shape) != tuple(shape):
            raise ValueError(f"expected shape {tuple(shape)}, got {tuple(t.shape)}")
        return t.data_ptr()

    def empty(self, *shape, dtype=torch.float64):
        return torch.empty(*shape, dtype=dtype, device=self.device)

    # ------------------------------------------------------------------ dynamics kernels
    def rollout_st(self, x0, U, kappa, ds, jac=False):
        B, N = U.shape[0], U.shape[1]
        X = self.empty(B, N + 1, 8)
        AB = self.empty(B, N, 8, 10) if jac else None
        check(self.lib.vc_rollout_st(self._h, B, N, self._p(x0, (B, 8)), self._p(U, (B, N, 2)),
                                     self._p(kappa, (B, N)), self._p(ds, (B, N)), self._p(X), self._p(AB)),
              "vc_rollout_st")
        return (X, AB) if jac else X

    def rollout_st_into(self, x0, U, kappa, ds, X, AB=None):
        B, N = U.shape[0], U.shape[1]
        check(self.lib.vc_rollout_st(self._h, B, N, self._p(x0, (B, 8)), self._p(U, (B, N, 2)),
                                     self._p(kappa, (B, N)), self._p(ds, (B, N)), self._p(X, (B, N + 1, 8)),
                                     self._p(AB, (B, N, 8, 10)) if AB is not None else None), "vc_rollout_st")

    def rollout_pm(self, x0, U, kappa, ds, jac=False):
        B, M = U.shape[0], U.shape[1]
        X = self.empty(B, M + 1, 5)
        AB = self.empty(B, M, 5, 7) if jac else None
        check(self.lib.vc_rollout_pm(self._h, B, M, self._p(x0, (B, 5)), self._p(U, (B, M, 2)),
                                     self._p(kappa, (B, M)), self._p(ds, (B, M)), self._p(X), self._p(AB)),
              "vc_rollout_pm")
        return (X, AB) if jac else X

    def plant_step(self, x, u, dt=None, track_id=None):
        """In-place closed-loop plant step of every vehicle (racing_car.py:34-46); `track_id` [B] int32 selects the
        track of each vehicle from the atlas (set_track_atlas)."""
        B = x.shape[0]
        if track_id is not None:
            check(self.lib.vc_plant_step_tracks(self._h, B, self._p(track_id, (B,), dtype=torch.int32), self._p(x, (B, 8)),
                                                self._p(u, (B, 2)), float(self.params.dt if dt is None else dt)),
                  "vc_plant_step_tracks")
            return x
        check(self.lib.vc_plant_step(self._h, B, self._p(x, (B, 8)), self._p(u, (B, 2)),
                                     float(self.params.dt if dt is None else dt)), "vc_plant_step")
        return x

    def tyre_forces(self, x, u):
        B = x.shape[0]
        out = self.empty(B, 6)
        check(self.lib.vc_tyre_forces(self._h, B, self._p(x, (B, 8)), self._p(u, (B, 2)), self._p(out)),
              "vc_tyre_forces")
        return out

    # ------------------------------------------------------------------ track kernels
    def curvature(self, s):
        out = torch.empty_like(s)
        check(self.lib.vc_track_curvature(self._h, s.numel(), self._p(s), self._p(out)), "vc_track_curvature")
        return out

    def rel2glob(self, s, ey, epsi):
        n = s.numel()
        out = self.empty(*s.shape, 3)
        check(self.lib.vc_rel2glob(self._h, n, self._p(s), self._p(ey), self._p(epsi), self._p(out)), "vc_rel2glob")
        return out

    def horizon_params(self, x0, Xprev, ds=None, kappa=None, track_id=None):
        B, H = x0.shape[0], self.H
        ds = self.empty(B, H) if ds is None else ds
        kappa = self.empty(B, H) if kappa is None else kappa
        if track_id is not None:
            check(self.lib.vc_horizon_params_tracks(self._h, B, self._p(track_id, (B,), dtype=torch.int32),
                                                    self._p(x0, (B, 8)), self._p(Xprev, (B, 8, H)),
                                                    self._p(ds, (B, H)), self._p(kappa, (B, H))), "vc_horizon_params_tracks")
            return ds, kappa
        check(self.lib.vc_horizon_params(self._h, B, self._p(x0, (B, 8)), self._p(Xprev, (B, 8, H)),
                                         self._p(ds, (B, H)), self._p(kappa, (B, H))), "vc_horizon_params")
        return ds, kappa

    # ------------------------------------------------------------------ NMPC
    def solve(self, x0, kappa, ds, X, U, obstacles=None, out=None):
        """Batched NMPC solve (cascaded_mpc.py:306-314).  X[B,8,H], U[B,2,H] are warm start in / solution out.
        Returns dict(obj, status, iters, kkt) of device tensors."""
        B, H = x0.shape[0], self.H
        if out is None:
            out = {"obj": self.empty(B), "status": self.empty(B, dtype=torch.int32),
                   "iters": self.empty(B, dtype=torch.int32), "kkt": self.empty(B)}
        n_obs = 0 if obstacles is None else obstacles.shape[1]
        check(self.lib.vc_nmpc_solve(self._h, B, self._p(x0, (B, 8)), self._p(kappa, (B, H)), self._p(ds, (B, H)),
                                     self._p(obstacles, (B, n_obs, 3)) if n_obs else None, n_obs,
                                     self._p(X, (B, 8, H)), self._p(U, (B, 2, H)), self._p(out["obj"]),
                                     self._p(out["status"], dtype=torch.int32),
                                     self._p(out["iters"], dtype=torch.int32), self._p(out["kkt"])), "vc_nmpc_solve")
        return out

    def solve_static(self, x0, kappa, ds, X, U, out, obstacles=None):
        """`solve` through vc_nmpc_solve_static: nothing is allocated inside the call (the solver's scratch is a static
        16 MiB device array of the library, at most 63 instances at H = 55), so `out` (obj, status, iters, kkt) must be
        given.  No counterpart in the reference; calls that overlap on one GPU race (include/vc_api.h)."""
        B, H = x0.shape[0], self.H
        n_obs = 0 if obstacles is None else obstacles.shape[1]
        check(self.lib.vc_nmpc_solve_static(self._h, B, self._p(x0, (B, 8)), self._p(kappa, (B, H)), self._p(ds, (B, H)),
                                            self._p(obstacles, (B, n_obs, 3)) if n_obs else None, n_obs,
                                            self._p(X, (B, 8, H)), self._p(U, (B, 2, H)), self._p(out["obj"]),
                                            self._p(out["status"], dtype=torch.int32),
                                            self._p(out["iters"], dtype=torch.int32), self._p(out["kkt"])),
              "vc_nmpc_solve_static")
        return out

    def evaluate(self, x0, kappa, ds, X, U, obstacles=None):
        """[B,3] objective, max equality violation, min inequality margin of a given trajectory."""
        B, H = x0.shape[0], self.H
        out = self.empty(B, 3)
        n_obs = 0 if obstacles is None else obstacles.shape[1]
        check(self.lib.vc_nmpc_eval(self._h, B, self._p(x0, (B, 8)), self._p(kappa, (B, H)), self._p(ds, (B, H)),
                                    self._p(obstacles, (B, n_obs, 3)) if n_obs else None, n_obs,
                                    self._p(X, (B, 8, H)), self._p(U, (B, 2, H)), self._p(out)), "vc_nmpc_eval")
        return out

    def solve_host(self, x0, kappa, ds, X, U, obstacles=None):
        """Host-buffer call (numpy in / out; X and U are updated in place).  Copies are inside the call."""
        B, H = x0.shape[0], self.H
        for a in (x0, kappa, ds, X, U):
            assert a.dtype == np.float64 and a.flags.c_contiguous
        obj, kkt = np.empty(B), np.empty(B)
        status, iters = np.empty(B, np.int32), np.empty(B, np.int32)
        n_obs = 0 if obstacles is None else obstacles.shape[1]
        check(self.lib.vc_nmpc_solve_host(self._h, B, x0.ctypes.data, kappa.ctypes.data, ds.ctypes.data,
                                          obstacles.ctypes.data if n_obs else None, n_obs, X.ctypes.data,
                                          U.ctypes.data, obj.ctypes.data, status.ctypes.data, iters.ctypes.data,
                                          kkt.ctypes.data), "vc_nmpc_solve_host")
        return {"obj": obj, "status": status, "iters": iters, "kkt": kkt}

    def command_host(self, x0, X, U, obstacles=None):
        """`command()` for a batch from host buffers (numpy; X and U = previous prediction in, solution out): horizon
        parameters, solve and all copies inside the call (include/vc_api.h: vc_nmpc_command_host)."""
        B = x0.shape[0]
        for a in (x0, X, U):
            assert a.dtype == np.float64 and a.flags.c_contiguous
        obj, kkt = np.empty(B), np.empty(B)
        status, iters = np.empty(B, np.int32), np.empty(B, np.int32)
        n_obs = 0 if obstacles is None else obstacles.shape[1]
        check(self.lib.vc_nmpc_command_host(self._h, B, x0.ctypes.data, obstacles.ctypes.data if n_obs else None, n_obs,
                                            X.ctypes.data, U.ctypes.data, obj.ctypes.data, status.ctypes.data,
                                            iters.ctypes.data, kkt.ctypes.data), "vc_nmpc_command_host")
        return {"obj": obj, "status": status, "iters": iters, "kkt": kkt}

    def set_blocking_sync(self, on: bool):
        """Host waits of the solve sleep on a blocking-sync event instead of spinning (include/vc_api.h)."""
        check(self.lib.vc_set_blocking_sync(self._h, 1 if on else 0), "vc_set_blocking_sync")

    def set_profiling(self, on: bool):
        check(self.lib.vc_set_profiling(self._h, 1 if on else 0), "vc_set_profiling")

    def profile(self):
        """dict(ms=[k_eval_st, k_eval_pm, k_step], launches=[...], last_passes=int) since set_profiling(True)."""
        ms, ln, lp = (C.c_double * 3)(), (C.c_longlong * 3)(), C.c_int()
        check(self.lib.vc_get_profile(self._h, C.addressof(ms), C.addressof(ln), C.addressof(lp)), "vc_get_profile")
        return {"ms": list(ms), "launches": list(ln), "last_passes": lp.value,
                "instance_passes": int(self.lib.vc_get_profile_instance_passes(self._h))}

    def debug_read(self, b, array, n):
        out = np.empty(n)
        check(self.lib.vc_nmpc_debug_read(self._h, int(b), int(array), out.ctypes.data, int(n)), "vc_nmpc_debug_read")
        return out

    def workspace_bytes(self) -> int:
        N, M = self.cfg.N, self.cfg.M
        return int(self.lib.vc_workspace_bytes(self._h))

    def launch_count(self) -> int:
        return int(self.lib.vc_launch_count(self._h))
