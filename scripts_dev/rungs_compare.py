"""Same 8192-instance C3 batch solved with the barrier ladder entered 0 / 1 / 2 / 3 rungs down: agreement of the answers."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench_common as bc
from bench_nmpc import controller_config, replay_lap, make_batch, SEED
from vehicle_control_b200.engine import Engine
from vehicle_control_b200.environment.track import Track
B = 8192
cfg = controller_config()
tr = Track(bc.load_yaml("environment/ippodromo.yaml"))
eng = Engine(bc.load_yaml("models/dynamic_car.yaml"), cfg, tr)
lap = dict(np.load("/tmp/lap.npz")) if os.path.exists("/tmp/lap.npz") else replay_lap(eng, cfg, tr)
x0h, Xwh, Uwh, obh, idx = make_batch(lap, tr, B, SEED, cfg)
res = {}
for r in (0, 1, 2, 3):
    eng.cfg.warm_rungs = r; eng.set_config(eng.cfg)
    x0, X, U, ob = eng.tensor(x0h), eng.tensor(Xwh), eng.tensor(Uwh), eng.tensor(obh)
    ds, kap = eng.horizon_params(x0, X)
    out = eng.solve(x0, kap, ds, X, U, obstacles=ob)
    res[r] = (U.cpu().numpy(), out["status"].cpu().numpy(), out["iters"].cpu().numpy(), out["obj"].cpu().numpy())
    st, it = res[r][1], res[r][2]
    print(f"rungs {r}: converged {np.mean(st <= 1):.4f} mean iters {it.mean():.2f} p50 {np.median(it)} p99 {np.percentile(it, 99)}")
U0, s0 = res[0][0], res[0][1]
for r in (1, 2, 3):
    U1, s1 = res[r][0], res[r][1]
    ok = (s0 <= 1) & (s1 <= 1)
    d = (np.abs(U1 - U0) / np.maximum(1.0, np.abs(U0))).reshape(B, -1).max(1)
    do = np.abs(res[r][3] - res[0][3]) / np.maximum(1.0, np.abs(res[0][3]))
    print(f"rungs {r} vs 0: both converged {ok.mean():.4f}; control diff p50 {np.median(d[ok]):.2e} p99 {np.percentile(d[ok], 99):.2e} "
          f"frac>1e-6 {np.mean(d[ok] > 1e-6):.4f}; first-action diff frac>1e-6 {np.mean((np.abs(U1[:, :, 0] - U0[:, :, 0]) / np.maximum(1, np.abs(U0[:, :, 0]))).max(1)[ok] > 1e-6):.4f}; obj rel diff p99 {np.percentile(do[ok], 99):.2e}")
