"""bench.py --impl reference, and the cpu_baseline leg: the reference's CPU path of the hot path on the box's host cores.

CasADi / IPOPT are not installable offline and the reference is pure Python on top of them, so it cannot be
compiled or run here.  The CPU arm is therefore the C++ fp64 restatement under oracle/cpp (kind "port-cpp": exact
derivatives by second-order jets, IPOPT's published algorithm with second-order correction and inertia correction, a
block-tridiagonal Bunch-Kaufman LDL^T in the role of MA27; g++ -O3, OpenMP over instances), with NO iteration cap
(max_iter = IPOPT's default 3000), on all host threads, on a bounded sample of the very workload the B200 arm times:

  nmpc / c2   a closed-loop lap is replayed on the CPU oracle from the reference's cold start (untimed, the protocol of
              bench_nmpc.replay_lap), bench_nmpc.make_batch with the B200 arm's seed builds the 65,536 (4,096) perturbed
              instances, and each step solves the first `--cpu-instances` (default 1,024) of them
  rollout     each step rolls out the first 16,384 trajectories x N = 50 with first-order jets for the [A|B] blocks
The reference's own logs give 14-25 solves/s for IPOPT+MA27 on one thread -- the restatement runs at a
comparable per-thread rate, so the ratio against this arm is quotable.
"""
from __future__ import annotations

import json
import os
import time

import numpy as np

import bench_common as bc


class OracleTrackView:
    """what bench_nmpc.make_batch needs of a track, served by the oracle's track (no product code on this arm)"""
    def __init__(self, tr):
        self.tr, self.length = tr, tr.length

    def obstacle_array(self):
        return np.array([o[:3] for o in self.tr.obstacles], dtype=float)


def oracle_lap(cfg, tr, max_ticks=2000, threads=0):
    """Closed loop of ONE vehicle on the CPU oracle (racing.py:217-242 semantics; same protocol as bench_nmpc.replay_lap):
    cold guess with the legacy MT19937 seed 31, un-shifted warm start, retry from the cold guess on a failure."""
    from oracle import cpp_oracle as co, nlp as on
    c = on.MpcCfg.from_config(cfg)
    cc = co.mpc_config(c)
    so = solver_options(cfg)
    X0, U0 = on.cold_start(c, np.random.RandomState(31).random_sample((2, c.H)))
    X, U = X0.copy(), U0.copy()
    x = np.array([4.0, 0, 0, 0, 1.0, 0, 0, 0])
    ob = np.array([o[:3] for o in tr.obstacles], dtype=float) if c.obstacles else None
    rows = {"x0": [], "Xw": [], "Uw": [], "iters": [], "status": []}
    for _ in range(max_ticks):
        if x[4] > tr.length - 0.1:
            break
        rows["x0"].append(x.copy()); rows["Xw"].append(X.copy()); rows["Uw"].append(U.copy())
        ds, kap = on.horizon_params(c, tr, x, X[0])
        r = co.solve(cc, x, kap, ds, X, U, obstacles=ob, threads=1, **so)
        if int(r["status"]) > 1:
            ds, kap = on.horizon_params(c, tr, x, X0[0])
            r = co.solve(cc, x, kap, ds, X0, U0, obstacles=ob, threads=1, **so)
        rows["iters"].append(int(r["iters"])); rows["status"].append(int(r["status"]))
        if rows["status"][-1] > 1:
            break
        X, U = r["X"], r["U"]
        x = co.plant_step(x[None], U[None, :, 0], np.atleast_1d(tr.k(x[4])), 0.05)[0]
    return {k: np.array(v) for k, v in rows.items()}


def nmpc_sample(workload, n, lap=None):
    """First n instances of the B200 arm's batch (same recipe, same seed); horizon parameters by the oracle's track."""
    import bench_nmpc as bn
    from oracle import nlp as on
    from oracle.track import OracleTrack
    c2 = workload == "c2"
    cfg = bn.controller_config(horizon_pm=15, obstacles=False) if c2 else bn.controller_config()
    tr = OracleTrack(bc.load_yaml("environment/ippodromo.yaml"))
    t0 = time.perf_counter()
    if lap is None:
        lap = oracle_lap(cfg, tr)
    lap_s = time.perf_counter() - t0
    B = 4096 if c2 else 65536
    x0, Xw, Uw, ob, _ = bn.make_batch(lap, OracleTrackView(tr), B, 1002 if c2 else bn.SEED, cfg)
    n = min(n, B)
    c = on.MpcCfg.from_config(cfg)
    ds, kap = np.empty((n, c.H)), np.empty((n, c.H))
    for i in range(n):
        ds[i], kap[i] = on.horizon_params(c, tr, x0[i], Xw[i, 0])
    return cfg, dict(x0=x0[:n], Xw=Xw[:n], Uw=Uw[:n], obst=None if ob is None else ob[:n], ds=ds, kappa=kap), lap_s, len(lap["x0"])


def solver_options(cfg):
    """IPOPT's defaults as the reference leaves them (max_iter 3000, max_soc 4, watchdog trigger 10) overridden by the
    controller config's `solver` block -- the same block the B200 arm reads (engine.mpc_config_from)."""
    o = {"max_iter": 3000, "max_soc": 4, "watchdog": 10}
    for k, v in dict(cfg.get("solver") or {}).items():
        o[{"watchdog_trigger": "watchdog"}.get(k, k)] = int(v)
    return {k: o[k] for k in ("max_iter", "max_soc", "watchdog")}


def cpu_solve(cfg, s, threads=0):
    from oracle import cpp_oracle as co
    return co.solve(cfg, s["x0"], s["kappa"], s["ds"], s["Xw"], s["Uw"], obstacles=s["obst"], threads=threads,
                    **solver_options(cfg))


def host_threads():
    return max(1, len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1))


def cpu_baseline_nmpc(cfg, sample, gpu_U=None, gpu_obj=None):
    """cpu_baseline leg of the B200 arm: the same first instances of the batch the GPU just solved, on all host threads;
    also reports how many of them the two arms agree on (controls and objective within 1e-6)."""
    th = host_threads()
    n = len(sample["x0"])
    t0 = time.perf_counter()
    r = cpu_solve(cfg, sample, th)
    el = time.perf_counter() - t0
    out = {"value": n / el, "unit": "solves/s", "cores": th, "kind": "port-cpp", "per_core": n / el / th,
           "sample": f"first {n} instances of this batch solved by oracle/cpp (C++ restatement of CasADi AD + IPOPT, same solver options, OpenMP over instances, {th} threads) in {el:.1f} s; mean {r['iters'].mean():.1f} iterations; "
                     f"the reference's own logs: 14-25 solves/s per thread (IPOPT+MA27, BASELINE.md)",
           "mean_iters": float(r["iters"].mean()), "converged_frac": float((r["status"] <= 1).mean())}
    if gpu_U is not None:
        rel = lambda a, b: np.abs(a - b) / np.maximum(1.0, np.abs(b))
        eu = rel(gpu_U, r["U"]).reshape(n, -1).max(1)
        eo = rel(gpu_obj, r["obj"])
        same = (eu < 1e-6) & (eo < 1e-6)
        out["parity"] = {"instances": n, "agree_1e-6": int(same.sum()), "frac": float(same.mean()),
                         "max_rel_err_controls_of_agreeing": float(eu[same].max()) if same.any() else None,
                         "note": "instances that disagree end in a different local solution of the non-convex NLP "
                                 "(DESIGN.md section 2)"}
    return out


def run_reference(a, workload, emit=None):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import cpp_oracle as co
    th = host_threads()
    steps, warm = max(1, a.steps), max(0, a.warmup)
    extra = {}
    if workload == "rollout":
        from oracle.track import OracleTrack
        tr = OracleTrack(bc.load_yaml("environment/ippodromo.yaml"))
        nb = 16384
        x0, U, kap, ds = bc.rollout_inputs(nb, 50, 1005, tr, bc.golden_states())
        fn = lambda: co.rollout_st(x0, U, kap, ds, jac=True, threads=th)
        units, unit, metric = nb * 50, "state-steps/s", "rollout+Jacobian state-steps/sec"
        wl = "C5: 1M SingleTrack trajectories x N=50 spatial RK4 + [A|B] Jacobians (ippodromo)"
        sample = f"{nb} trajectories x N=50 per step, oracle/cpp (RK4 in first-order jets), {th} threads"
    else:
        n = getattr(a, "cpu_instances", 0) or 1024
        cfg, s, lap_s, ticks = nmpc_sample(workload, n)
        n = len(s["x0"])
        last = {}
        def fn():
            last["r"] = cpu_solve(cfg, s, th)
        units, unit, metric = n, "solves/s", "NMPC solves/sec"
        wl = ("C2: cascaded NMPC 20+15, ippodromo, 4,096 perturbed instances; warm start timed, cold start reported separately"
              if workload == "c2" else "C3: cascaded NMPC 20+35 with obstacles, ippodromo, warm-started perturbed instances")
        sample = (f"first {n} instances of the batch per step (same recipe and seed as the B200 arm; warm-start lap of {ticks} "
                  f"ticks replayed on the CPU oracle in {lap_s:.0f} s, untimed), oracle/cpp with max_iter {solver_options(cfg)['max_iter']}, {th} threads")
    for _ in range(min(warm, 1)):
        fn()
    t0 = time.perf_counter()
    for _ in range(steps):
        fn()
    el = time.perf_counter() - t0
    value = units * steps / el
    if workload != "rollout":
        r = last["r"]
        extra = {"mean_iters": float(r["iters"].mean()), "max_iters": int(r["iters"].max()),
                 "converged_frac": float((r["status"] <= 1).mean()), "per_core": value / th}
    line = {"impl": "reference", "metric": metric, "value": value, "unit": unit, "n_gpus": a.gpus, "steps": steps,
            "warmup": warm, "ms_per_step": 1e3 * el / steps, "higher_is_better": True, "scaling": getattr(a, "scaling", "strong"),
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": {"workload": wl}, "gpu_launches": 0,
            "cpu_baseline": dict({"value": value, "unit": unit, "cores": th, "kind": "port-cpp", "sample": sample}, **extra),
            "e2e": {"value": value, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    (emit or (lambda obj: print(json.dumps(obj), flush=True)))(line)
    return 0
