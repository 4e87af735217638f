import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench_common as bc
from bench_nmpc import controller_config
from vehicle_control_b200.engine import Engine
from vehicle_control_b200.environment.track import Track
lap = np.load(sys.argv[1]); t = int(sys.argv[2]); obst = int(sys.argv[3])
cfg = controller_config(horizon_pm=35, obstacles=bool(obst))
tr = Track(bc.load_yaml("environment/ippodromo.yaml"))
eng = Engine(bc.load_yaml("models/dynamic_car.yaml"), cfg, tr)
x = eng.tensor(lap["x0"][t][None]); X = eng.tensor(lap["Xw"][t][None]); U = eng.tensor(lap["Uw"][t][None])
ds, kap = eng.horizon_params(x, X)
ob = eng.tensor(tr.obstacle_array()[None]) if obst else None
out = eng.solve(x, kap, ds, X, U, obstacles=ob)
print({k: v.cpu().numpy() for k, v in out.items()})
H = 55
for name, arr, w in (("GRES", 11, 8), ("DY", 15, 10), ("Y", 0, 10), ("S", 2, 10), ("Z", 3, 10), ("DSL", 17, 10), ("CRES", 12, 10)):
    a = eng.debug_read(0, arr, H * w).reshape(H, w)
    np.set_printoptions(linewidth=250, precision=4, suppress=False)
    print(name); print(a)
