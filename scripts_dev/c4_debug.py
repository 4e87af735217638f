import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from bench_common import load_yaml, ROOT
from vehicle_control_b200.simulation.racing import BatchedRacingSimulator
meta = json.load(open(os.path.join(ROOT, "tests", "golden", "runs.json")))
cfg = dict(meta["race2_shoe/cascaded"]["config"]); cfg["horizon_pm"] = int(sys.argv[1]) if len(sys.argv) > 1 else 35
B, T = 512, 6
rng = np.random.default_rng(1004)
x0 = np.zeros((B, 8)); x0[:, 0] = 4.0; x0[:, 4] = 1.0 + rng.uniform(0, 5, B); x0[:, 5] = rng.uniform(-1, 1, B)
x0[0] = [4, 0, 0, 0, 1, 0, 0, 0]
np.random.seed(31)
sim = BatchedRacingSimulator(load_yaml("models/dynamic_car.yaml"), load_yaml("environment/shoe.yaml"), cfg, x0)
for t in range(T):
    n = sim.update()
    st = sim.controller.info["status"].cpu().numpy(); it = sim.controller.info["iters"].cpu().numpy()
    print(f"tick {t}: active {n} status {dict(zip(*np.unique(st, return_counts=True)))} iters p50 {np.median(it)} max {it.max()}", flush=True)
    if t == 0:
        bad = np.flatnonzero(st > 1)[:8]
        print("   failed x0:", x0[bad][:, [4, 5]], "kkt", sim.controller.info["kkt"].cpu().numpy()[bad])
