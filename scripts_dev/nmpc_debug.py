import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from bench_common import load_yaml
from vehicle_control_b200.engine import Engine
from vehicle_control_b200.environment.track import Track
z = np.load("tests/golden/nmpc_oracle.npz"); meta = json.load(open("tests/golden/runs.json"))
tags = sorted({k.rsplit("/", 1)[0] for k in z.files})
pat = sys.argv[1] if len(sys.argv) > 1 else "cascaded1"
for tag in tags:
    if pat not in tag: continue
    rk = tag.rsplit("/", 1)[0]; m = meta[rk]
    eng = Engine(load_yaml("models/dynamic_car.yaml"), m["config"], Track(load_yaml(f"environment/{m['track']}.yaml")))
    f = {k: z[f"{tag}/{k}"] for k in ("x0", "kappa", "ds", "obst", "X0", "U0", "X", "U", "obj", "ref_u", "iters")}
    t = lambda a: eng.tensor(np.ascontiguousarray(a[None]))
    X, U = t(f["X0"]), t(f["U0"])
    out = eng.solve(t(f["x0"]), t(f["kappa"]), t(f["ds"]), X, U, obstacles=t(f["obst"]) if f["obst"].size else None)
    torch.cuda.synchronize()
    Un = U.cpu().numpy()[0]
    print(tag, {k: v.cpu().numpy()[0] for k, v in out.items()}, "u0", Un[:, 0], "oracle", f["U"][:, 0], "ref", f["ref_u"], "oracle obj", f["obj"], "iters", f["iters"], flush=True)
    print("max |dU|", np.abs(Un - f["U"]).max())
