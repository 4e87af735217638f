import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tests.test_gpu_tracks import _setup, NAMES
from tests.conftest import load_yaml
from vehicle_control_b200.simulation.racing import BatchedRacingSimulator
B = 32
eng, cfg, tracks, tid, x0h, X0, U0 = _setup(B)
x0 = np.zeros((B, 8)); x0[:, 0] = 4.0; x0[:, 4] = 1.0
np.random.seed(31)
sim = BatchedRacingSimulator(load_yaml("models/dynamic_car.yaml"), tracks[0], cfg, x0, record=False, tracks=tracks, track_id=tid)
n = 0
prev = 0
while n < 1500 and sim.update() > 0:
    n += 1
    f = int(sim.failed.sum())
    if f != prev:
        st = sim.controller.info["status"].cpu().numpy(); it = sim.controller.info["iters"].cpu().numpy()
        bad = np.flatnonzero(sim.failed.cpu().numpy())
        print("tick", n, "failed", bad, "tracks", [NAMES[tid[b]] for b in bad], "status", st[bad], "iters", it[bad], "s", sim.x[bad, 4].cpu().numpy(), "retries", sim.controller.retries)
        prev = f
print("ticks", n, "retries", sim.controller.retries)
