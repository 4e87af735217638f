"""Find the slowest / failing instances of the bench batch and trace them one at a time (VC_DEBUG)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench_common as bc
from bench_nmpc import controller_config, replay_lap, make_batch, SEED
from vehicle_control_b200.engine import Engine
from vehicle_control_b200.environment.track import Track
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
cfg = controller_config()
tr = Track(bc.load_yaml("environment/ippodromo.yaml"))
eng = Engine(bc.load_yaml("models/dynamic_car.yaml"), cfg, tr)
cache = "/tmp/lap.npz"
lap = dict(np.load(cache)) if os.path.exists(cache) else replay_lap(eng, cfg, tr)
x0h, Xwh, Uwh, obh, idx = make_batch(lap, tr, B, SEED, cfg)
x0, X, U, ob = eng.tensor(x0h), eng.tensor(Xwh), eng.tensor(Uwh), eng.tensor(obh)
ds, kap = eng.horizon_params(x0, X)
out = eng.solve(x0, kap, ds, X, U, obstacles=ob)
st, it = out["status"].cpu().numpy(), out["iters"].cpu().numpy()
print("status counts", np.unique(st, return_counts=True), "iters pct", np.percentile(it, [50, 90, 99, 99.9]))
for lo in (60, 80, 100, 150, 200, 250, 300):
    m = it >= lo
    print(f"iters >= {lo}: {int(m.sum())} instances, of which converged {int((m & (st <= 1)).sum())}")
order = np.argsort(-it)
slow_ok = [b for b in order if st[b] == 0][:4]
fails = [b for b in order if st[b] > 1][:5]
np.savez("gpurun_out/slow_cases.npz", b=np.array(slow_ok + fails), x0=x0h[slow_ok + fails], Xw=Xwh[slow_ok + fails],
         Uw=Uwh[slow_ok + fails], ob=obh[slow_ok + fails], tick=idx[slow_ok + fails])
os.environ["VC_DEBUG"] = "1"
for b in slow_ok + fails:
    print(f"==== instance {b} tick {idx[b]} status {st[b]} iters {it[b]} x0 {x0h[b]}", flush=True)
    sys.stderr.write(f"==== instance {b}\n"); sys.stderr.flush()
    x1, X1, U1, o1 = eng.tensor(x0h[b:b+1]), eng.tensor(Xwh[b:b+1]), eng.tensor(Uwh[b:b+1]), eng.tensor(obh[b:b+1])
    d1, k1 = eng.horizon_params(x1, X1)
    o = eng.solve(x1, k1, d1, X1, U1, obstacles=o1)
    print({k: v.cpu().numpy() for k, v in o.items()}, flush=True)
