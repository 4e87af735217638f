"""bench leg C2: cascaded 20+15 NMPC on ippodromo, 4,096 perturbed instances per GPU, warm-started solve
timed, cold-start (`ones`) solve timed and reported separately.  Not the headline (that is C3, bench_nmpc.py); run with
`python bench.py --workload c2`."""
from __future__ import annotations

import json
import os
import time

import numpy as np

import bench_common as bc
from bench_nmpc import make_batch, replay_lap

SEED = 1002


def run_c2(a, rank, world, local):
    import torch
    from vehicle_control_b200.controllers.mpc.cascaded_mpc import cold_guess
    from vehicle_control_b200.engine import Engine
    from vehicle_control_b200.environment.track import Track
    B = a.batch or 4096
    with open(os.path.join(bc.ROOT, "tests", "golden", "runs.json")) as f:
        cfg = dict(json.load(f)["cascaded1_ippodromo/cascaded"]["config"])   # = race1_ippodromo/cascaded_config.yaml
    tr = Track(bc.load_yaml("environment/ippodromo.yaml"))
    eng = Engine(bc.load_yaml("models/dynamic_car.yaml"), cfg, tr, device=local)
    N, H = int(cfg["horizon"]), int(cfg["horizon"]) + int(cfg["horizon_pm"])
    np.random.seed(31)
    lap = replay_lap(eng, cfg, tr)
    x0h, Xwh, Uwh, _, _ = make_batch(lap, tr, B, SEED + rank, cfg)
    x0, Xw, Uw = eng.tensor(x0h), eng.tensor(Xwh), eng.tensor(Uwh)
    Xc0, Uc0 = cold_guess(N, H, rng_draw=np.random.RandomState(31).random_sample((2, H)))
    Xc, Uc = eng.tensor(np.repeat(Xc0[None], B, 0)), eng.tensor(np.repeat(Uc0[None], B, 0))
    X, U = torch.empty_like(Xw), torch.empty_like(Uw)
    ds, kap = eng.empty(B, H), eng.empty(B, H)
    out = {"obj": eng.empty(B), "status": eng.empty(B, dtype=torch.int32), "iters": eng.empty(B, dtype=torch.int32),
           "kkt": eng.empty(B)}

    def step_from(Xs, Us):
        def step():
            X.copy_(Xs); U.copy_(Us)
            eng.horizon_params(x0, X, ds, kap)
            eng.solve(x0, kap, ds, X, U, out=out)
        return step

    res = {}
    l0 = eng.launch_count()
    with bc.ClockSampler(local) as clk:
        for name, Xs, Us in (("warm", Xw, Uw), ("cold", Xc, Uc)):
            total_ms, _ = bc.time_steps(step_from(Xs, Us), a.steps, a.warmup)
            st, it = out["status"].cpu().numpy(), out["iters"].cpu().numpy()
            res[name] = {"solves_per_s": world * B * a.steps / (total_ms * 1e-3), "ms_per_step": total_ms / a.steps,
                         "converged_frac": float((st <= 1).mean()), "mean_iters": float(it.mean()), "max_iters": int(it.max())}
    launches = eng.launch_count() - l0
    # e2e: host buffers through vc_nmpc_solve_host (warm start)
    pin = lambda arr: torch.from_numpy(np.ascontiguousarray(arr)).pin_memory()
    keep = [pin(x0h), pin(kap.cpu().numpy()), pin(ds.cpu().numpy()), pin(Xwh), pin(Uwh)]
    hx0, hk, hd, hX, hU = (t.numpy() for t in keep)
    eng.solve_host(hx0, hk, hd, hX, hU)
    t = 0.0
    for _ in range(min(3, a.steps)):
        hX[:], hU[:] = Xwh, Uwh
        t1 = time.perf_counter(); eng.solve_host(hx0, hk, hd, hX, hU); t += time.perf_counter() - t1
    e2e_s = bc.max_over_ranks(t / min(3, a.steps))
    return {"metric": "NMPC solves/sec", "value": res["warm"]["solves_per_s"], "unit": "solves/s", "n_gpus": world,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": res["warm"]["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "C2: cascaded NMPC 20+15, ippodromo, 4,096 perturbed instances; warm start timed, "
                                   "cold start reported separately", "batch_per_gpu": B, "N": N, "M": H - N,
                       "warm_start": res["warm"], "cold_start": res["cold"],
                       "l2": "solver workspace %.2f GB per step > 126 MB L2" % (eng.workspace_bytes() / 1e9)},
            "clocks": clk.summary(), "gpu_launches": int(round(launches / (2 * (a.steps + a.warmup)) * a.steps)),
            "e2e": {"value": world * B / e2e_s, "unit": "solves/s",
                    "h2d_bytes_per_step": int(sum(v.nbytes for v in (hx0, hk, hd, hX, hU))),
                    "d2h_bytes_per_step": int(hX.nbytes + hU.nbytes + B * 24)}}
