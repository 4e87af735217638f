"""Which instances of the C3 bench batch do not end with status 0, after how many iterations and at which final NLP
error (kkt)?  usage: fail_stats.py [batch]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import bench_common as bc
from bench_nmpc import SEED, controller_config, make_batch, replay_lap
from vehicle_control_b200.engine import Engine
from vehicle_control_b200.environment.track import Track

B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
cfg = controller_config()
tr = Track(bc.load_yaml("environment/ippodromo.yaml"))
eng = Engine(bc.load_yaml("models/dynamic_car.yaml"), cfg, tr, device=0)
cache = os.path.join(bc.ROOT, "tmp_data", "lap_cache.npz")
lap = dict(np.load(cache)) if os.path.exists(cache) else replay_lap(eng, cfg, tr)
x0h, Xwh, Uwh, obh, idx = make_batch(lap, tr, B, SEED, cfg)
x0, X, U, ob = eng.tensor(x0h), eng.tensor(Xwh), eng.tensor(Uwh), eng.tensor(obh)
ds, kap = eng.horizon_params(x0, X)
out = eng.solve(x0, kap, ds, X, U, obstacles=ob)
st, it, kkt = out["status"].cpu().numpy(), out["iters"].cpu().numpy(), out["kkt"].cpu().numpy()
print("status counts", {int(k): int(v) for k, v in zip(*np.unique(st, return_counts=True))})
order = np.argsort(-it)
print("the 80 longest-running instances: (index, tick, status, iters, kkt)")
for b in order[:80]:
    print(int(b), int(idx[b]), int(st[b]), int(it[b]), f"{kkt[b]:.2e}")
bad = st > 1
print("iterations spent by instances that do not converge:", int(it[bad].sum()), "of", int(it.sum()))
print("passes > 60 are held by", int((it > 60).sum()), "instances;", int(((it > 60) & bad).sum()), "of them do not converge")
