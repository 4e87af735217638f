"""GPU side: evaluate the derivative blocks at the warm start of a lap tick (max_iter=0) and dump the workspace."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench_common as bc
from bench_nmpc import controller_config
from vehicle_control_b200.engine import Engine
from vehicle_control_b200.environment.track import Track
lap = np.load(sys.argv[1]); t = int(sys.argv[2]); obst = int(sys.argv[3])
cfg = controller_config(horizon_pm=35, obstacles=bool(obst))
tr = Track(bc.load_yaml("environment/ippodromo.yaml"))
eng = Engine(bc.load_yaml("models/dynamic_car.yaml"), cfg, tr)
eng.set_config(cfg, max_iter=0)
x = eng.tensor(lap["x0"][t][None]); X = eng.tensor(lap["Xw"][t][None]); U = eng.tensor(lap["Uw"][t][None])
ds, kap = eng.horizon_params(x, X)
ob = eng.tensor(tr.obstacle_array()[None]) if obst else None
out = eng.solve(x, kap, ds, X, U, obstacles=ob)
H = 55
names = {"Y": (0, 10), "LAM": (1, 8), "S": (2, 10), "Z": (3, 10), "HC": (4, 36), "AB": (5, 80), "JC": (6, 80), "RD": (7, 10),
         "GF": (8, 10), "V1": (9, 10), "V2": (10, 10), "GRES": (11, 8), "CRES": (12, 10), "FST": (13, 1)}
d = {k: eng.debug_read(0, a, H * w).reshape(H, w) for k, (a, w) in names.items()}
d["scal"] = eng.debug_read(0, -1, 64)
d["x0"] = lap["x0"][t]; d["Xw"] = lap["Xw"][t]; d["Uw"] = lap["Uw"][t]; d["ds"] = ds.cpu().numpy()[0]; d["kappa"] = kap.cpu().numpy()[0]
np.savez("gpurun_out/dump_eval.npz", **d)
print("status", out["status"].cpu().numpy(), "sf", d["scal"][1], "mu", d["scal"][0])
