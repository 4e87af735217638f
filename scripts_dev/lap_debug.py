import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, time
import bench_common as bc
from bench_nmpc import controller_config, replay_lap
from vehicle_control_b200.engine import Engine
from vehicle_control_b200.environment.track import Track
obst = int(sys.argv[1]) if len(sys.argv) > 1 else 1
pm = int(sys.argv[2]) if len(sys.argv) > 2 else 35
cfg = controller_config(horizon_pm=pm, obstacles=bool(obst))
tr = Track(bc.load_yaml("environment/ippodromo.yaml"))
eng = Engine(bc.load_yaml("models/dynamic_car.yaml"), cfg, tr)
t0 = time.time()
lap = replay_lap(eng, cfg, tr)
print("ticks", len(lap["x0"]), "time", time.time() - t0, "status tail", lap["status"][-5:], "iters", lap["iters"])
print("last x0", lap["x0"][-1])
np.savez(f"gpurun_out/lap_debug_{obst}.npz", **lap)
if lap["status"][-1] > 1:
    os.environ["VC_DEBUG"] = "1"
    t = len(lap["x0"]) - 1
    x = eng.tensor(lap["x0"][t][None]); X = eng.tensor(lap["Xw"][t][None]); U = eng.tensor(lap["Uw"][t][None])
    ds, kap = eng.horizon_params(x, X)
    ob = eng.tensor(tr.obstacle_array()[None]) if obst else None
    out = eng.solve(x, kap, ds, X, U, obstacles=ob)
    print({k: v.cpu().numpy() for k, v in out.items()})
