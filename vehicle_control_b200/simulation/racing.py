"""Headless closed-loop racing simulation on the B200 engine.

``RacingSimulator`` mirrors vehicle_control/simulation/racing.py:23-459 + simulator.py:13-101 without the
matplotlib/Qt animation that drives the reference's loop (no graphics stack on the GPU box): same construction
from (simconfig, carconfig, trackconfig), same start state (Ux=4, s=1, :49), same per-tick order (command ->
drive -> log, :230-241), same termination (s > L - 0.1, :219), same containers and the same on-disk run format
({name}_state_traj.npy, _action_traj.npy, _preds.npy, _elapsed.npy, _config.yaml; :441-459).

``BatchedRacingSimulator`` advances B vehicles of one controller type in lock step, everything on the device
(BASELINE config C4): horizon parameters -> batched NMPC solve -> batched plant step -> termination mask.
"""
from __future__ import annotations

import os
import time

import numpy as np
import torch
import yaml

from ..controllers.mpc.cascaded_mpc import BatchedCascadedMPC, CascadedMPC
from ..engine import Engine
from ..environment.track import Track
from ..models.dynamic_car import DynamicCar, DynamicCarAction, DynamicCarState
from ..models.dynamic_point_mass import DynamicPointMass
from ..utils.common_utils import as_config, config_dir, load_config


class RacingSimulator:
    def __init__(self, simconfig, carconfig, trackconfig, controller_configs=None, device=None):
        self.config = as_config(simconfig)
        self.names = list(self.config.controller_names)
        self.track = Track(trackconfig)
        carconfig = as_config(carconfig)
        self.cars, self.controllers = [], []
        for i, name in enumerate(self.names):
            # reference: config/controllers/{name}.yaml relative to the CWD (:53-56); here the packaged copies,
            # or explicit configs (e.g. the *_config.yaml saved next to a reference run)
            cc = controller_configs[i] if controller_configs else load_config(
                os.path.join(config_dir(), "controllers", f"{name}.yaml"))
            cc = as_config(cc)
            cc.obstacles = bool(self.config.obstacles)            # :57-58
            eng = Engine(carconfig, cc, self.track, device=device)
            car = DynamicCar(carconfig, self.track, engine=eng)
            car.state = DynamicCarState(Ux=4, s=1)                # :48-49
            pm = DynamicPointMass(carconfig, self.track, engine=eng)
            self.cars.append(car)
            self.controllers.append(CascadedMPC(car=car, point_mass=pm, config=cc, engine=eng))
        self.colors = [c.config.get("color") for c in self.controllers]
        self.loaded = False
        self.init_containers()
        self.data_path = os.path.join(self.config.get("data_dir", "experiments/data"), self.name)
        if self.config.get("load"):
            self.load()

    @property
    def name(self):
        return f"{self.config.name}_{self.config.track_name}"

    @property
    def state_len(self):
        return len(self.state_traj[self.names[0]])

    def init_containers(self):
        self.state_traj = {n: [c.state] for n, c in zip(self.names, self.cars)}
        self.action_traj = {n: [c.create_action()] for n, c in zip(self.names, self.cars)}   # zero placeholder :82-84
        self.elapsed = {n: [] for n in self.names}
        self.preds = {n: [] for n in self.names}
        self.iters = {n: [] for n in self.names}

    # ------------------------------------------------------------------ loop
    def finished(self) -> bool:
        return any(car.state.s > self.track.length - 0.1 for car in self.cars)

    def step(self, controller, car):
        try:
            action = controller.command(car.state)
            state = car.drive(action)
        except Exception as e:   # racing.py:416-423 prints and returns None
            print(e)
            return None
        return action, state

    def update(self, n) -> bool:
        """One tick of every controller; returns False when the lap is over (racing.py:217-242)."""
        if self.finished():
            if self.config.get("save_data"):
                self.save()
            return False
        for name, car, controller in zip(self.names, self.cars, self.controllers):
            start = time.time()
            res = self.step(controller, car)
            if res is None:
                return False
            action, state = res
            self.elapsed[name].append(time.time() - start)
            self.state_traj[name].append(state)
            self.action_traj[name].append(action)
            self.preds[name].append(controller.get_state_prediction())
            self.iters[name].append(int(controller.last_info["iters"]))
            if self.config.get("logging"):
                self.debug_print(n, car, self.elapsed[name], action, state)
        return True

    def run(self, max_ticks=100000):
        if self.loaded:
            return
        n = 0
        while n < max_ticks and self.update(n):
            n += 1
        return n

    # ------------------------------------------------------------------ logging / persistence
    def debug_print(self, n, car, elapsed, action, state):
        print("-" * 83)
        print(f"N: {n}")
        print(f"STATE: {state}")
        print(f"ACTION: {action}")
        print(f"AVERAGE ELAPSED TIME: {np.mean(elapsed):.3f}")
        print(f"MEDIAN ELAPSED TIME: {np.median(elapsed):.3f}")
        car.print(state, action)
        print("-" * 83 + "\n")

    def _arrays(self, name):
        st = np.array([np.asarray(getattr(s, "values", s), dtype=float) for s in self.state_traj[name]])
        ac = np.array([np.asarray(getattr(a, "values", a), dtype=float) for a in self.action_traj[name]])
        return st, ac, np.array(self.preds[name]), np.array(self.elapsed[name])

    def save(self):
        os.makedirs(self.data_path, exist_ok=True)
        for name, controller in zip(self.names, self.controllers):
            st, ac, pr, el = self._arrays(name)
            np.save(f"{self.data_path}/{name}_state_traj.npy", st)
            np.save(f"{self.data_path}/{name}_action_traj.npy", ac)
            np.save(f"{self.data_path}/{name}_preds.npy", pr)
            np.save(f"{self.data_path}/{name}_elapsed.npy", el)
            cfg = controller.config
            with open(f"{self.data_path}/{name}_config.yaml", "w") as f:
                yaml.safe_dump(cfg.to_dict() if hasattr(cfg, "to_dict") else dict(cfg), f)

    def load(self):
        for name, controller in zip(self.names, self.controllers):
            self.state_traj[name] = np.load(f"{self.data_path}/{name}_state_traj.npy")
            self.action_traj[name] = np.load(f"{self.data_path}/{name}_action_traj.npy")
            self.preds[name] = np.load(f"{self.data_path}/{name}_preds.npy")
            self.elapsed[name] = np.load(f"{self.data_path}/{name}_elapsed.npy")
            controller.config = load_config(f"{self.data_path}/{name}_config.yaml")
        self.loaded = True

    def summary(self):
        """The quantities racing.py:92-106 prints, as a dict per controller."""
        out = {}
        for name in self.names:
            st, ac, _, el = self._arrays(name)
            out[name] = {"laptime": float(st[-1, -1]), "average_time": float(np.mean(el)) if len(el) else 0.0,
                         "median_time": float(np.median(el)) if len(el) else 0.0,
                         "average_speed": float(np.mean(st[:, 0])), "mean_abs_Fx": float(np.mean(np.abs(ac[:, 0]))),
                         "std_Fx": float(np.std(ac[:, 0])), "mean_squared_ey": float(np.mean(np.square(st[:, 5])))}
        return out

    def summarize(self):
        print(self.name)
        for name, s in self.summary().items():
            print("-------------------------")
            print(name)
            print(f"Laptime: {s['laptime']}")
            print(f"Average time:{s['average_time']}")
            print(f"Median time:{s['median_time']}")
            print(f"Average speed: {s['average_speed']}")
            print(f"Mean Fx: {s['mean_abs_Fx']}")
            print(f"Std Fx: {s['std_Fx']}")
            print(f"Mean squared error: {s['mean_squared_ey']}")
            print("-------------------------")


class BatchedRacingSimulator:
    """B vehicles, one controller configuration, lock-step ticks, device resident (BASELINE config C4)."""

    def __init__(self, carconfig, trackconfig, controller_config, x0, obstacles=None, device=None, record=True,
                 tracks=None, track_id=None, engine=None):
        self.track = trackconfig if isinstance(trackconfig, Track) else Track(trackconfig)
        self.config = as_config(controller_config)
        # engine: an existing handle (e.g. a lane of pipeline.InFlight on its own stream) instead of a new one
        self.engine = engine if engine is not None else Engine(as_config(carconfig), self.config, self.track, device=device)
        e = self.engine
        x0 = np.ascontiguousarray(x0, dtype=float)
        self.B = x0.shape[0]
        car = DynamicCar(carconfig, self.track, engine=e)
        self.controller = BatchedCascadedMPC(car, None, self.config, self.B, engine=e, obstacles=obstacles)
        self.x = e.tensor(x0)
        # per-vehicle tracks: `tracks` (list of Track) and `track_id` [B]; the finish line is each vehicle's own
        self.track_id = None
        self.lengths = torch.full((self.B,), float(self.track.length), dtype=torch.float64, device=e.device)
        if tracks is not None:
            self.controller.set_tracks(tracks, track_id)
            self.track_id = self.controller.track_id
            self.lengths = e.tensor(np.array([tracks[int(i)].length for i in np.asarray(track_id)], dtype=float))
        self.active = torch.ones(self.B, dtype=torch.bool, device=e.device)
        self.ticks = torch.zeros(self.B, dtype=torch.int32, device=e.device)
        self.failed = torch.zeros(self.B, dtype=torch.bool, device=e.device)
        self.record = record
        self.state_traj, self.action_traj = [self.x.clone()], []
        self.total_solves = 0

    def update(self) -> int:
        """One lock-step tick.  Vehicles past the finish line (or whose solve failed) keep their state."""
        e = self.engine
        self.active &= ~(self.x[:, 4] > self.lengths - 0.1) & ~self.failed
        n_active = int(self.active.sum())
        if n_active == 0:
            return 0
        u = self.controller.command(self.x, active=self.active)
        self.failed |= self.active & (self.controller.info["status"] > 1)
        xn = self.x.clone()
        e.plant_step(xn, u, track_id=self.track_id)
        keep = (self.active & ~self.failed).unsqueeze(1)
        self.x = torch.where(keep, xn, self.x).contiguous()
        self.ticks += keep.squeeze(1).to(torch.int32)
        self.total_solves += n_active
        if self.record:
            self.state_traj.append(self.x.clone()); self.action_traj.append(u.clone())
        return n_active

    def run(self, max_ticks=3000):
        n = 0
        while n < max_ticks and self.update():
            n += 1
        return n

    def laptimes(self):
        return self.x[:, 7].cpu().numpy()

    # ------------------------------------------------------------------ persistence / summary (racing.py:441-459, 92-106)
    def _trajectories(self):
        """state_traj [T+1,B,8], action_traj [T+1,B,2] with the reference's zero placeholder in row 0 (racing.py:79-84)."""
        if not self.record:
            raise RuntimeError("BatchedRacingSimulator(record=False) keeps no trajectories")
        st = torch.stack(self.state_traj).cpu().numpy()
        ac = torch.stack([torch.zeros_like(self.action_traj[0])] + self.action_traj).cpu().numpy() if self.action_traj \
            else np.zeros((1, self.B, 2))
        return st, ac

    def save(self, data_dir, name="batched"):
        """One compressed `{name}_batched.npz` for all vehicles (+ `{name}_config.yaml`): state_traj [T+1,B,8],
        action_traj [T+1,B,2], ticks [B] (rows 0..ticks[b] of vehicle b are its lap), failed [B]."""
        os.makedirs(data_dir, exist_ok=True)
        st, ac = self._trajectories()
        np.savez_compressed(os.path.join(data_dir, f"{name}_batched.npz"), state_traj=st, action_traj=ac,
                            ticks=self.ticks.cpu().numpy(), failed=self.failed.cpu().numpy(),
                            track_id=np.zeros(self.B, np.int32) if self.track_id is None else self.track_id.cpu().numpy())
        cfg = self.config
        with open(os.path.join(data_dir, f"{name}_config.yaml"), "w") as f:
            yaml.safe_dump(cfg.to_dict() if hasattr(cfg, "to_dict") else dict(cfg), f)

    @staticmethod
    def load(data_dir, name="batched"):
        z = np.load(os.path.join(data_dir, f"{name}_batched.npz"))
        return {k: z[k] for k in z.files}, load_config(os.path.join(data_dir, f"{name}_config.yaml"))

    def export_vehicle(self, b, data_dir, name="cascaded"):
        """Vehicle b in the reference's own run format (racing.py:441-450): `{name}_state_traj.npy`, `_action_traj.npy`,
        `_elapsed.npy`, `_config.yaml` (and `_preds.npy` = rel2glob of the states, one row per tick), so that the
        reference's replay / summarize tooling reads it."""
        os.makedirs(data_dir, exist_ok=True)
        st, ac = self._trajectories()
        T = int(self.ticks[b])
        s, a = st[:T + 1, b], ac[:T + 1, b]
        np.save(os.path.join(data_dir, f"{name}_state_traj.npy"), s)
        np.save(os.path.join(data_dir, f"{name}_action_traj.npy"), a)
        np.save(os.path.join(data_dir, f"{name}_elapsed.npy"), np.zeros(T))
        e = self.engine
        xy = e.rel2glob(e.tensor(s[:-1, 4]), e.tensor(s[:-1, 5]), e.tensor(s[:-1, 6])).cpu().numpy()
        np.save(os.path.join(data_dir, f"{name}_preds.npy"), xy[:, None, :])
        cfg = self.config
        with open(os.path.join(data_dir, f"{name}_config.yaml"), "w") as f:
            yaml.safe_dump(cfg.to_dict() if hasattr(cfg, "to_dict") else dict(cfg), f)

    def summary(self):
        """racing.py:92-106 for every vehicle at once: arrays [B] of lap time, average speed, mean |Fx|, std Fx, mean
        squared lateral deviation (over each vehicle's own ticks), plus the fleet aggregates."""
        st, ac = self._trajectories()
        T = self.ticks.cpu().numpy()
        k = np.arange(st.shape[0])[:, None]
        ms = (k <= T[None]).astype(float)                       # states 0..T[b]
        ma = ((k >= 1) & (k <= T[None])).astype(float)          # actions 1..T[b] (row 0 is the placeholder)
        ns, na = ms.sum(0), np.maximum(ma.sum(0), 1.0)
        fx = ac[:, :, 0]
        # the reference averages |Fx| and std over the whole action array including the zero placeholder row
        nfull = na + 1.0
        mean_fx = (ma * np.abs(fx)).sum(0) / nfull
        mfx = (ma * fx).sum(0) / nfull
        out = {"laptime": st[np.minimum(T, st.shape[0] - 1), np.arange(self.B), 7], "ticks": T,
               "finished": ~self.failed.cpu().numpy() & (self.x[:, 4] > self.lengths - 0.1).cpu().numpy(),
               "average_speed": (ms * st[:, :, 0]).sum(0) / ns, "mean_abs_Fx": mean_fx,
               "std_Fx": np.sqrt(np.maximum(((ma * fx ** 2).sum(0) + 0.0) / nfull - mfx ** 2, 0.0)),
               "mean_squared_ey": (ms * st[:, :, 5] ** 2).sum(0) / ns}
        fin = out["finished"]
        out["fleet"] = {"vehicles": int(self.B), "finished": int(fin.sum()), "failed": int(self.failed.sum()),
                        "median_laptime": float(np.median(out["laptime"][fin])) if fin.any() else float("nan"),
                        "total_solves": int(self.total_solves), "retries": dict(self.controller.retries)}
        return out

    def summarize(self):
        s = self.summary()
        print(f"vehicles {s['fleet']['vehicles']}: finished {s['fleet']['finished']}, failed {s['fleet']['failed']}")
        print(f"Laptime: median {s['fleet']['median_laptime']}  min {np.min(s['laptime'])}  max {np.max(s['laptime'])}")
        print(f"Average speed: {np.mean(s['average_speed'])}")
        print(f"Mean Fx: {np.mean(s['mean_abs_Fx'])}")
        print(f"Mean squared error: {np.mean(s['mean_squared_ey'])}")
