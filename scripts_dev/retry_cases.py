"""Which fallback rescues the failing closed-loop ticks (gpurun_out/cases/*.npz)?"""
import sys, os, json, glob
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from bench_common import load_yaml, ROOT
from vehicle_control_b200.engine import Engine
from vehicle_control_b200.environment.track import Track
from vehicle_control_b200.controllers.mpc.cascaded_mpc import cold_guess
meta = json.load(open(os.path.join(ROOT, "tests", "golden", "runs.json")))
for f in sorted(glob.glob(os.path.join(ROOT, "scripts_dev", "cases", "*_[0-9][0-9][0-9].npz"))):
    c = np.load(f); key = str(c["key"]); m = meta[key]; cfg = dict(m["config"])
    tr = Track(load_yaml(f"environment/{m['track']}.yaml"))
    eng = Engine(load_yaml("models/dynamic_car.yaml"), cfg, tr)
    N, H = int(cfg["horizon"]), int(cfg["horizon"]) + int(cfg["horizon_pm"])
    ob = eng.tensor(tr.obstacle_array()[None]) if cfg.get("obstacles") else None
    def attempt(name, Xg, Ug):
        x = eng.tensor(c["x0"][None]); X = eng.tensor(Xg[None]); U = eng.tensor(Ug[None])
        ds, kap = eng.horizon_params(x, X)
        out = eng.solve(x, kap, ds, X, U, obstacles=ob)
        print(f"{os.path.basename(f)} {name}: status {int(out['status'][0])} iters {int(out['iters'][0])} obj {float(out['obj'][0]):.6f} u0 {U[0,:,0].cpu().numpy()}", flush=True)
        return int(out['status'][0]), X.cpu().numpy()[0], U.cpu().numpy()[0]
    attempt("warm", c["Xw"], c["Uw"])
    np.random.seed(31)
    Xc, Uc = cold_guess(N, H)
    attempt("cold", Xc, Uc)
    Xs, Us = c["Xw"].copy(), c["Uw"].copy()
    Xs[:, :N - 1] = Xs[:, 1:N]; Us[:, :N - 1] = Us[:, 1:N]
    attempt("shifted", Xs, Us)
    # cold guess placed at the vehicle: s, ey, epsi, Ux from x0
    Xp = Xc.copy(); Xp[0, :N] = max(c["x0"][0], 4.0); Xp[4, :N] = c["x0"][4] + np.arange(N) * 0.03 * max(c["x0"][0], 4.0)
    Xp[5, :N] = c["x0"][5]; Xp[6, :N] = 0; Xp[1:4, :N] = 0; Xp[7, :N] = c["x0"][7]
    if H > N:
        Xp[0, N:] = Xp[0, N - 1]; Xp[1, N:] = Xp[4, N - 1] + (1 + np.arange(H - N)) * float(cfg.get("ds_pm", 3)); Xp[2, N:] = c["x0"][5]; Xp[3, N:] = 0; Xp[4, N:] = c["x0"][7]
    Up = np.zeros_like(Uc)
    attempt("placed", Xp, Up)
