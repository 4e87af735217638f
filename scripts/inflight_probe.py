"""How much throughput do several batches IN FLIGHT give on one GPU?  Each of n engines (own handle, own non-blocking
stream, own workspace) solves the C3 bench batch from its own host thread; while one is in its latency-bound tail the
bulk passes of another fill the GPU.
usage: inflight_probe.py [n_engines] [steps_per_engine] [batch] [stagger_s]"""
import os
import sys
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import bench_common as bc
from bench_nmpc import SEED, controller_config, make_batch, replay_lap
from vehicle_control_b200.engine import Engine
from vehicle_control_b200.environment.track import Track

n_eng = int(sys.argv[1]) if len(sys.argv) > 1 else 2
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
B = int(sys.argv[3]) if len(sys.argv) > 3 else 65536
stagger = float(sys.argv[4]) if len(sys.argv) > 4 else 0.0
cfg = controller_config()
tr = Track(bc.load_yaml("environment/ippodromo.yaml"))
car = bc.load_yaml("models/dynamic_car.yaml")
cache = os.path.join(bc.ROOT, "tmp_data", "lap_cache.npz")
eng0 = Engine(car, cfg, tr, device=0)
lap = dict(np.load(cache)) if os.path.exists(cache) else replay_lap(eng0, cfg, tr)
x0h, Xwh, Uwh, obh, _ = make_batch(lap, tr, B, SEED, cfg)
H = cfg["horizon"] + cfg["horizon_pm"]


class Lane:
    def __init__(self, i):
        self.stream = torch.cuda.Stream(device=0)
        self.eng = Engine(car, cfg, tr, device=0, stream=self.stream.cuda_stream)
        e = self.eng
        with torch.cuda.stream(self.stream):
            self.x0, self.Xw, self.Uw, self.ob = e.tensor(x0h), e.tensor(Xwh), e.tensor(Uwh), e.tensor(obh)
            self.X, self.U = torch.empty_like(self.Xw), torch.empty_like(self.Uw)
            self.ds, self.kap = e.empty(B, H), e.empty(B, H)
            self.out = {"obj": e.empty(B), "status": e.empty(B, dtype=torch.int32), "iters": e.empty(B, dtype=torch.int32),
                        "kkt": e.empty(B)}
        self.stream.synchronize()

    def step(self):
        with torch.cuda.stream(self.stream):
            self.X.copy_(self.Xw); self.U.copy_(self.Uw)
            self.eng.horizon_params(self.x0, self.X, self.ds, self.kap)
            self.eng.solve(self.x0, self.kap, self.ds, self.X, self.U, obstacles=self.ob, out=self.out)

    def run(self, n, delay):
        time.sleep(delay)
        for _ in range(n):
            self.step()
        self.stream.synchronize()


lanes = [Lane(i) for i in range(n_eng)]
for ln in lanes:       # warm-up (allocates the workspaces)
    ln.step(); ln.step()
torch.cuda.synchronize()
ref = {k: v.clone() for k, v in lanes[0].out.items()}
refU = lanes[0].U.clone()
t0 = time.perf_counter()
th = [threading.Thread(target=ln.run, args=(steps, i * stagger)) for i, ln in enumerate(lanes)]
for t in th:
    t.start()
for t in th:
    t.join()
torch.cuda.synchronize()
dt = time.perf_counter() - t0
tot = n_eng * steps * B
same = all(torch.equal(ln.U, refU) and torch.equal(ln.out["status"], ref["status"]) for ln in lanes)
print(f"engines {n_eng} steps/engine {steps} batch {B} stagger {stagger}: {dt:.3f} s -> {tot / dt:.0f} solves/s "
      f"({dt / (n_eng * steps) * 1e3:.1f} ms per batch); results identical across engines: {same}", flush=True)
