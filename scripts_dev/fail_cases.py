"""Replay the simulations whose lap stopped on a solver failure, dump the failing tick as a case and trace its re-solve."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from bench_common import load_yaml, ROOT
from vehicle_control_b200.simulation.racing import RacingSimulator

meta = json.load(open(os.path.join(ROOT, "tests", "golden", "runs.json")))
os.makedirs("gpurun_out/cases", exist_ok=True)
for spec in sys.argv[1:]:
    sim_name, names = spec.split(":")
    names = names.split(",")
    m0 = meta[f"{sim_name}/{names[0]}"]
    np.random.seed(31)
    sc = {"track_name": m0["track"], "controller_names": names, "obstacles": bool(m0["config"].get("obstacles")),
          "slip_angle_plot": False, "name": "cmp", "save_gif": False, "save_images": False, "save_data": False,
          "load": False, "logging": False}
    sim = RacingSimulator(sc, load_yaml("models/dynamic_car.yaml"), load_yaml(f"environment/{m0['track']}.yaml"),
                          controller_configs=[meta[f"{sim_name}/{n}"]["config"] for n in names])
    n = 0
    while n < 1200:
        warm = [(c.state_prediction.copy(), c.action_prediction.copy(), np.asarray(car.state.values, dtype=float).copy())
                for c, car in zip(sim.controllers, sim.cars)]
        if not sim.update(n):
            break
        n += 1
    print(f"{spec}: stopped at tick {n}, finished {sim.finished()}", flush=True)
    if sim.finished():
        continue
    for name, ctl, car, w in zip(names, sim.controllers, sim.cars, warm):
        if len(sim.action_traj[name]) - 1 > n:       # this controller did tick n: not the failing one
            continue
        tag = f"{sim_name}_{name}_{n}"
        np.savez(f"gpurun_out/cases/{tag}.npz", key=f"{sim_name}/{name}", tick=n, x0=w[2], Xw=w[0], Uw=w[1])
        print(f"---- failing controller {name}: case {tag}; x0 {w[2]}", flush=True)
        e = ctl._batched.engine
        x = e.tensor(w[2][None]); X = e.tensor(w[0][None]); U = e.tensor(w[1][None])
        ds, kap = e.horizon_params(x, X)
        ob = e.tensor(sim.track.obstacle_array()[None]) if m0["config"].get("obstacles") else None
        os.environ["VC_DEBUG"] = "1"
        out = e.solve(x, kap, ds, X, U, obstacles=ob)
        del os.environ["VC_DEBUG"]
        print(f"---- re-solve status {int(out['status'][0])} iters {int(out['iters'][0])}", flush=True)
