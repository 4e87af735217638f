"""bench.py --impl reference: the reference's CPU implementation of the hot path on the box's host cores.

CasADi / IPOPT are not installable offline, so the reference arm times the CPU restatement under
oracle/ (numpy + torch-fp64 autodiff NLP and an IPOPT-style interior point: "kind": "port"), one single-threaded
worker process per host core, on a BOUNDED sample of the same workload the B200 arm runs: each step solves one
warm-started closed-loop tick of the C3 lap per worker (or, for the rollout workload, a block of trajectories per
worker).  The warm-start material comes from the committed fixture tests/golden/nmpc_oracle.npz (ticks of the
reference's own cascaded 20+35 run) so that no GPU is needed on this arm.
"""
from __future__ import annotations

import json
import os
import time

import numpy as np

import bench_common as bc


def _solve_tick(args):
    import torch
    torch.set_num_threads(1)
    from oracle import ipm, models as om, nlp as on
    cfg, x0, kap, ds, obst, X0, U0 = args
    c = on.MpcCfg.from_config(cfg)
    p = on.NLP(c, om.DEFAULT, x0, kap, ds, obst if c.obstacles else None)
    # bounded work per step: the interior-point loop is capped at 30 iterations (the B200 arm's median is 17, its
    # p90 37); a capped solve still counts as one solve, which can only favour this CPU arm
    r = ipm.solve(p, p.pack(X0, U0), ipm.IpmOptions(max_iter=30, max_restorations=0))
    return 1


def _rollout_block(args):
    import torch
    torch.set_num_threads(1)
    from oracle import models as om
    x0, U, kap, ds = args
    X = om.st_rollout_spatial(om.DEFAULT, x0, U, kap, ds)
    om.st_step_jac(om.DEFAULT, X[:, :-1].reshape(-1, 8), U.reshape(-1, 2), kap.reshape(-1), ds.reshape(-1))
    return x0.shape[0] * U.shape[1]


def _nmpc_tasks(n, c2=False):
    """n warm-started instances of the C3 workload: the reference run's warm ticks (fixture) with the bench's
    perturbation recipe applied (seed 1003), obstacles on with the 7 ippodromo obstacles jittered.
    c2: the C2 workload instead (cascaded 20+15 of the run cascaded1_ippodromo, no obstacles, seed 1002)."""
    from bench_nmpc import controller_config
    from oracle.track import OracleTrack
    tr = OracleTrack(bc.load_yaml("environment/ippodromo.yaml"))
    z = np.load(os.path.join(bc.ROOT, "tests", "golden", "nmpc_oracle.npz"))
    if c2:
        with open(os.path.join(bc.ROOT, "tests", "golden", "runs.json")) as f:
            cfg = dict(json.load(f)["cascaded1_ippodromo/cascaded"]["config"])
        ticks = ["cascaded1_ippodromo/cascaded/tick1", "cascaded1_ippodromo/cascaded/tick2"]
    else:
        cfg = controller_config()
        ticks = ["cascaded3_ippodromo/cascaded/tick1", "cascaded3_ippodromo/cascaded/tick2"]
    base = np.array([o[:3] for o in tr.obstacles], dtype=np.float64)
    rng = np.random.default_rng(1002 if c2 else 1003)
    from oracle import nlp as on
    c = on.MpcCfg.from_config(cfg)
    tasks = []
    for i in range(n):
        t = ticks[i % len(ticks)]
        x0 = z[t + "/x0"].copy()
        x0[0] = max(x0[0] + rng.uniform(-.5, .5), 3.2)
        x0[1] += rng.normal(0, .05); x0[2] += rng.normal(0, .02)
        x0[3] = float(np.clip(x0[3] + rng.uniform(-.02, .02), -.4, .4))
        x0[5] += rng.uniform(-.5, .5); x0[6] += rng.normal(0, .01); x0[7] = 0.0
        ob = base.copy()
        ob[:, 0] += rng.uniform(-1, 1, len(base)); ob[:, 1] += rng.uniform(-.3, .3, len(base))
        X0, U0 = z[t + "/X"].copy(), z[t + "/U"].copy()   # previous solution, un-shifted (cascaded_mpc.py:320-321)
        ds, kap = on.horizon_params(c, tr, x0, X0[0])
        tasks.append((cfg, x0, kap, ds, ob, X0, U0))
    return cfg, tasks


def run_reference(a, workload):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import multiprocessing as mp
    cores = max(1, len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1))
    procs = min(cores, 64)
    steps, warm = max(1, a.steps), max(0, a.warmup)
    # one single-threaded worker per core: keep BLAS / OpenMP from oversubscribing the cores (inherited by the workers)
    for v in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS", "NUMEXPR_NUM_THREADS"):
        os.environ[v] = "1"
    ctx = mp.get_context("spawn")
    with ctx.Pool(procs) as pool:
        if workload == "rollout":
            from oracle.track import OracleTrack
            tr = OracleTrack(bc.load_yaml("environment/ippodromo.yaml"))
            blk = 256
            mk = lambda i: bc.rollout_inputs(blk, 50, 1005 + i, tr, bc.golden_states())
            tasks = [mk(i) for i in range(procs)]
            fn, unit, metric = _rollout_block, "state-steps/s", "rollout+Jacobian state-steps/sec"
            wl = "C5: SingleTrack trajectories x N=50 spatial RK4 + [A|B] (ippodromo)"
            sample = f"{procs} x {blk} trajectories x N=50 per step (oracle numpy RK4 + torch.func.jacfwd)"
        else:
            cfg, tasks = _nmpc_tasks(procs, c2=workload == "c2")
            fn, unit, metric = _solve_tick, "solves/s", "NMPC solves/sec"
            wl = ("C2: cascaded NMPC 20+15, ippodromo, 4,096 perturbed instances; warm start timed, cold start reported "
                  "separately" if workload == "c2" else
                  "C3: cascaded NMPC 20+35 with obstacles, ippodromo, warm-started perturbed instances")
            sample = (f"{procs} warm-started perturbed instances per step (one per worker, interior-point loop capped at 30 "
                      f"iterations and counted as solved), oracle/nlp.py + oracle/ipm.py; "
                      f"the reference's own logs report 14-25 solves/s for IPOPT+MA27 on 1 thread (BASELINE.md)")
        for _ in range(min(warm, 1)):       # one warm-up pass is enough to page the workers in (each pass is ~30 s)
            pool.map(fn, tasks)
        done, t0 = 0, time.perf_counter()
        for _ in range(steps):
            done += sum(pool.map(fn, tasks))
        el = time.perf_counter() - t0
    value = done / el
    line = {"impl": "reference", "metric": metric, "value": value, "unit": unit, "n_gpus": a.gpus, "steps": steps,
            "warmup": warm, "ms_per_step": 1e3 * el / steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": {"workload": wl},
            "gpu_launches": 0,
            "cpu_baseline": {"value": value, "unit": unit, "cores": procs, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)
    return 0
