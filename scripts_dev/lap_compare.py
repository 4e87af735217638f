"""Closed-loop laps on the engine vs the reference's saved runs, tick by tick (diagnostic; writes gpurun_out/lapcmp_*.npz)."""
import sys, os, json, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from bench_common import load_yaml, ROOT
from vehicle_control_b200.simulation.racing import RacingSimulator

z = np.load(os.path.join(ROOT, "tests", "golden", "runs.npz"))
meta = json.load(open(os.path.join(ROOT, "tests", "golden", "runs.json")))
keys = sys.argv[1:] or ["cascaded3_ippodromo/cascaded"]
os.makedirs("gpurun_out", exist_ok=True)
for key in keys:
    m = meta[key]
    name = key.split("/")[1]
    np.random.seed(31)
    sc = {"track_name": m["track"], "controller_names": [name], "obstacles": bool(m["config"].get("obstacles")),
          "slip_angle_plot": False, "name": "cmp", "save_gif": False, "save_images": False, "save_data": False,
          "load": False, "logging": False}
    sim = RacingSimulator(sc, load_yaml("models/dynamic_car.yaml"), load_yaml(f"environment/{m['track']}.yaml"),
                          controller_configs=[m["config"]])
    ctl = sim.controllers[0]
    S, A = z[key + "/state"], z[key + "/action"]
    t0 = time.time()
    warm = []
    n = 0
    while n < len(S) + 50:
        warm.append((ctl.state_prediction.copy(), ctl.action_prediction.copy()))
        if not sim.update(n):
            break
        n += 1
    st, ac, pr, el = sim._arrays(name)
    T = min(len(ac), len(A)) - 1
    eu = (np.abs(ac[1:T + 1] - A[1:T + 1]) / np.maximum(1.0, np.abs(A[1:T + 1]))).max(1)
    ex = (np.abs(st[:T + 1] - S[:T + 1]) / np.maximum(1.0, np.abs(S[:T + 1]))).max(1)
    bad = np.flatnonzero(eu > 1e-6)
    it = np.array(sim.iters[name])
    print(f"{key}: ticks {n} (ref {len(S) - 1}) laptime {st[-1, 7]:.2f} (ref {S[-1, 7]:.2f}) wall {time.time() - t0:.1f}s "
          f"mean iters {it.mean():.1f} max {it.max()} | max action err {eu.max():.2e} max state err {ex.max():.2e} | "
          f"ticks with action err>1e-6: {len(bad)} first {bad[:8]}", flush=True)
    if len(bad):
        k = bad[0]
        print(f"   tick {k}: ours {ac[k + 1]} ref {A[k + 1]} iters {it[k]} state err before {ex[k]:.2e}")
    np.savez(f"gpurun_out/lapcmp_{key.replace('/', '_')}.npz", st=st, ac=ac, eu=eu, ex=ex, iters=it,
             Xw=np.array([w[0] for w in warm]), Uw=np.array([w[1] for w in warm]))
