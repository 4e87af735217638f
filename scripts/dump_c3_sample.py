"""Dumps instances of the C3 bench batch (inputs and the engine's answers) for CPU analysis with the oracle.

Runs on the GPU box: builds the bench_nmpc batch (same seed as bench.py), solves it once and writes
  gpurun_out/c3_lap.npz     the untimed warm-start lap (so that any instance can be rebuilt on the CPU)
  gpurun_out/c3_sample.npz  stragglers (status > 1 or iterations > --slow) and an evenly spaced sample of the rest:
                            idx, x0, obst, Xw, Uw, kappa, ds (inputs) and X, U, obj, status, iters, kkt (outputs)

    python scripts/dump_c3_sample.py [--batch 65536] [--sample 96] [--slow 60]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=65536)
    ap.add_argument("--sample", type=int, default=96)
    ap.add_argument("--slow", type=int, default=60)
    ap.add_argument("--out", default="gpurun_out")
    ap.add_argument("--workload", default="c3", choices=["c3", "c2"])
    a = ap.parse_args()
    import torch
    import bench_common as bc
    import bench_nmpc as bn
    from vehicle_control_b200.engine import Engine
    from vehicle_control_b200.environment.track import Track
    os.makedirs(a.out, exist_ok=True)
    if a.workload == "c3":
        cfg, seed = bn.controller_config(), bn.SEED
    else:
        cfg, seed = bn.controller_config(horizon_pm=15, obstacles=False), 1002
    tr = Track(bc.load_yaml("environment/ippodromo.yaml"))
    eng = Engine(bc.load_yaml("models/dynamic_car.yaml"), cfg, tr, device=0)
    lap = bn.replay_lap(eng, cfg, tr)
    np.savez(os.path.join(a.out, f"{a.workload}_lap.npz"), **lap)
    B = a.batch
    x0h, Xwh, Uwh, obh, idx = bn.make_batch(lap, tr, B, seed, cfg)
    x0, X, U = eng.tensor(x0h), eng.tensor(Xwh), eng.tensor(Uwh)
    ob = eng.tensor(obh) if obh is not None else None
    ds, kap = eng.horizon_params(x0, X)
    out = eng.solve(x0, kap, ds, X, U, obstacles=ob)
    torch.cuda.synchronize()
    status, iters = out["status"].cpu().numpy(), out["iters"].cpu().numpy()
    hard = np.flatnonzero((status > 1) | (iters > a.slow))[:384]
    rest = np.setdiff1d(np.linspace(0, B - 1, a.sample).astype(int), hard)
    sel = np.concatenate([hard, rest])
    g = lambda t: t.cpu().numpy()[sel]
    np.savez(os.path.join(a.out, f"{a.workload}_sample.npz"), sel=sel, tick=idx[sel], n_hard=len(hard), x0=x0h[sel],
             obst=obh[sel] if obh is not None else np.zeros((len(sel), 0, 3)), Xw=Xwh[sel], Uw=Uwh[sel],
             kappa=g(kap), ds=g(ds), X=g(X), U=g(U), obj=g(out["obj"]), status=status[sel], iters=iters[sel], kkt=g(out["kkt"]))
    print(f"{a.workload}: B={B} status counts {dict(zip(*np.unique(status, return_counts=True)))} mean iters {iters.mean():.2f} "
          f"max {iters.max()} hard {len(hard)} sampled {len(sel)}")
    print("iters histogram (>=40):", {int(k): int(v) for k, v in zip(*np.unique(iters[iters >= 40], return_counts=True))})


if __name__ == "__main__":
    main()
