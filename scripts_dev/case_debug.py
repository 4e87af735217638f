"""Solve saved single-instance cases (scripts_dev/cases/*.npz) on the engine; VC_DEBUG=1 traces the iterations."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from bench_common import load_yaml, ROOT
from vehicle_control_b200.engine import Engine
from vehicle_control_b200.environment.track import Track
meta = json.load(open(os.path.join(ROOT, "tests", "golden", "runs.json")))
z = np.load(os.path.join(ROOT, "tests", "golden", "runs.npz"))
for name in sys.argv[1:]:
    c = np.load(os.path.join(ROOT, "scripts_dev", "cases", name + ".npz"))
    key, tick = str(c["key"]), int(c["tick"])
    m = meta[key]
    tr = Track(load_yaml(f"environment/{m['track']}.yaml"))
    eng = Engine(load_yaml("models/dynamic_car.yaml"), m["config"], tr)
    x = eng.tensor(c["x0"][None]); X = eng.tensor(c["Xw"][None]); U = eng.tensor(c["Uw"][None])
    ds, kap = eng.horizon_params(x, X)
    ob = eng.tensor(tr.obstacle_array()[None]) if m["config"].get("obstacles") else None
    print(f"==== {name}: {key} tick {tick}", flush=True)
    out = eng.solve(x, kap, ds, X, U, obstacles=ob)
    o = {k: v.cpu().numpy()[0] for k, v in out.items()}
    print(f"==== {name}: status {o['status']} iters {o['iters']} obj {o['obj']:.9f} kkt {o['kkt']:.2e} u0 {U[0, :, 0].cpu().numpy()} ref {z[key + '/action'][tick + 1]}", flush=True)
