"""bench leg C3 (headline): cascaded 20+35 NMPC with obstacles on ippodromo, batch 65,536 per GPU.

Workload: the controller config of the reference run race3_ippodromo (20 SingleTrack +
35 PointMass stages, max_speed 18) with obstacles enabled (weight 10, the 7 ippodromo obstacles).  A closed-loop
lap is first replayed on the engine itself from the reference's cold start (untimed); every tick gives a base
state and the warm start the reference protocol would use there (previous solution, un-shifted).  Instance b
takes tick b mod T, perturbs the initial state and jitters the obstacles (seed 1003 + rank).  The timed unit is
one `command()`-equivalent: horizon parameters + one batched interior-point solve from that warm start.
"""
from __future__ import annotations

import os
import time

import numpy as np

import bench_common as bc

SEED = 1003
MAX_ITER = 300


def controller_config(horizon_pm=35, obstacles=True, max_speed=18):
    c = bc.load_yaml("controllers/cascaded.yaml")
    c["horizon"], c["horizon_pm"], c["obstacles"] = 20, int(horizon_pm), bool(obstacles)
    c["state_constraints"]["max_speed"] = max_speed
    # solver options of the benchmark, applied to BOTH arms (B200 engine and CPU reference arm): IPOPT's defaults as the
    # reference leaves them, except the iteration limit (IPOPT 3000): a handful of the 65,536 perturbed instances crawl
    # along an obstacle pole for ~1000 iterations under IPOPT's algorithm without its restoration phase
    c["solver"] = {"max_iter": MAX_ITER}
    if os.environ.get("VC_BENCH_WIDE_MAX"):     # tuning runs: instances left below which a whole CTA serves one instance
        c["solver"]["wide_max"] = int(os.environ["VC_BENCH_WIDE_MAX"])
    return c


def replay_lap(eng, cfg, track, max_ticks=2000, obstacles=None):
    """Closed loop of ONE vehicle on the engine (racing.py:217-242 semantics).  Returns per tick the state the
    solve started from, the warm start it was given and the solution."""
    import torch
    from vehicle_control_b200.controllers.mpc.cascaded_mpc import cold_guess
    N, H = int(cfg["horizon"]), int(cfg["horizon"]) + int(cfg["horizon_pm"])
    X0, U0 = cold_guess(N, H, rng_draw=np.random.RandomState(31).random_sample((2, H)))
    X, U = eng.tensor(X0[None]), eng.tensor(U0[None])
    x = eng.tensor(np.array([[4.0, 0, 0, 0, 1.0, 0, 0, 0]]))
    ob = None
    if cfg.get("obstacles"):
        oa = track.obstacle_array() if obstacles is None else obstacles
        ob = eng.tensor(oa[None]) if len(oa) else None
    rows = {"x0": [], "Xw": [], "Uw": [], "iters": [], "status": []}
    for _ in range(max_ticks):
        if float(x[0, 4]) > track.length - 0.1:
            break
        rows["x0"].append(x.cpu().numpy()[0]); rows["Xw"].append(X.cpu().numpy()[0]); rows["Uw"].append(U.cpu().numpy()[0])
        ds, kap = eng.horizon_params(x, X)
        out = eng.solve(x, kap, ds, X, U, obstacles=ob)
        if int(out["status"][0]) > 1:      # what a user of the reference would do on an IPOPT failure: retry from the cold guess
            X.copy_(eng.tensor(X0[None])); U.copy_(eng.tensor(U0[None]))
            ds, kap = eng.horizon_params(x, X)
            out = eng.solve(x, kap, ds, X, U, obstacles=ob)
        rows["iters"].append(int(out["iters"][0])); rows["status"].append(int(out["status"][0]))
        if rows["status"][-1] > 1:
            break
        eng.plant_step(x, U[:, :, 0].contiguous())
    return {k: np.array(v) for k, v in rows.items()}


def make_batch(lap, track, B, seed, cfg):
    """Perturbed initial states + per-instance obstacle jitter around the lap's ticks (SURVEY C2/C3 recipe)."""
    rng = np.random.default_rng(seed)
    ok = np.flatnonzero(lap["status"] <= 1)
    ok = ok[ok >= 1]   # tick 0 has the cold guess, not a warm start
    T = len(ok)
    idx = ok[np.arange(B) % T]
    x0 = lap["x0"][idx].copy()
    x0[:, 0] = np.maximum(x0[:, 0] + rng.uniform(-.5, .5, B), 3.2)
    x0[:, 1] += rng.normal(0, .05, B)
    x0[:, 2] += rng.normal(0, .02, B)
    x0[:, 3] = np.clip(x0[:, 3] + rng.uniform(-.02, .02, B), -.40, .40)
    x0[:, 5] += rng.uniform(-.5, .5, B)
    x0[:, 6] += rng.normal(0, .01, B)
    x0[:, 7] = 0.0
    obst = None
    if cfg.get("obstacles"):
        base = track.obstacle_array()
        obst = np.repeat(base[None], B, 0)
        obst[:, :, 0] += rng.uniform(-1, 1, (B, len(base)))
        obst[:, :, 1] += rng.uniform(-.3, .3, (B, len(base)))
        # keep the (perturbed) start outside every obstacle's radius + 1 m: fall back to the un-perturbed state
        d = np.sqrt((x0[:, None, 4] - obst[:, :, 0]) ** 2 + (x0[:, None, 5] - obst[:, :, 1]) ** 2) - obst[:, :, 2]
        bad = (d < 1.0).any(1)
        x0[bad, 5] = lap["x0"][idx[bad], 5]
        d = np.sqrt((x0[:, None, 4] - obst[:, :, 0]) ** 2 + (x0[:, None, 5] - obst[:, :, 1]) ** 2) - obst[:, :, 2]
        bad = (d < 1.0).any(1)
        obst[bad] = base   # still too close: drop the jitter as well
    return x0, lap["Xw"][idx].copy(), lap["Uw"][idx].copy(), obst, idx


# algorithmic flops of one interior-point iteration (SURVEY 8d convention: one flop per add/mul/div/sqrt/
# compare-select/elementary function): N*(F_st_eval + F_ric) + M*(F_pm_eval + F_ric5)
def flops_per_iteration(N, M):
    return N * (15.0e3 + 4.0e3) + M * (0.5e3 + 1.5e3)   # SURVEY 8(d) convention (full-state Riccati: 4 k / 1.5 k per stage)


# algorithmic bytes of the two dominant kernels (doubles moved per unit)
EVAL_ST_BYTES = 8.0 * (10 + 8 + 10 + 10 + 8 + 8      # reads: y, lam(next), z, s, y(next) and the parameters
                       + 36 + 64 + 80 + 4 * 8 + 8 + 10 + 1)   # writes: Hessian block, dynamics / inequality Jacobians, gradients, residuals


def STEP_BYTES(H):
    """k_step per instance-iteration: per stage it reads the evaluation record (HC AB RD V1 V2 OBS GRES JC GF CRES:
    262 doubles) and the iterate (y lam s z: 38), writes + re-reads the Riccati factors (KM PV PM: 134, twice), writes the
    direction (38) and the new iterate (38), and one line-search trial re-reads iterate and direction (76)."""
    return 8.0 * H * (262 + 38 + 2 * 134 + 38 + 38 + 76)


def bytes_per_solve(H):
    """compulsory traffic: read x0, kappa, ds, warm start; write solution + (obj, status, iters, kkt)."""
    return 8.0 * ((8 + H + H + 10 * H) + (10 * H + 3))


def source_hash():
    """hash of the CUDA sources the ncu per-unit numbers (profiles/*_ncu_units.json) were captured for; comments and white
    space do not count (a comment-only edit leaves the SASS, and therefore the captured numbers, as they are)"""
    import hashlib
    import re
    h = hashlib.sha256()
    d = os.path.join(bc.ROOT, "vehicle_control_b200", "csrc")
    for f in sorted(os.listdir(d)):
        with open(os.path.join(d, f), "r", encoding="utf-8", errors="replace") as fh:
            txt = fh.read()
        txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
        txt = re.sub(r"//[^\n]*", "", txt)
        h.update(f.encode())
        h.update("".join(txt.split()).encode())
    return h.hexdigest()[:16]


def ncu_units():
    """per-unit DRAM bytes / executed flops of the kernels from the newest committed ncu capture, with a staleness flag"""
    import glob
    import json
    files = sorted(glob.glob(os.path.join(bc.ROOT, "profiles", "r*_ncu_units.json")))
    if not files:
        return None
    with open(files[-1]) as f:
        u = json.load(f)
    u["file"] = os.path.relpath(files[-1], bc.ROOT)
    u["stale"] = u.get("source_hash") != source_hash()
    return u


def time_in_flight(fl, step_fn, steps, warmup, stagger_s=0.0):
    """K steps over the lanes of `fl` (step i on lane i mod L, vehicle_control_b200/pipeline.py) between barrier +
    synchronize on both sides; device time from a start event to the last lane's end event, max over ranks."""
    import torch
    if warmup:
        fl.run(step_fn, warmup)
    torch.cuda.synchronize()
    bc.barrier()
    torch.cuda.synchronize()
    ends = {}

    def last(lane):
        e = torch.cuda.Event(enable_timing=True)
        e.record()                       # on the lane's stream (current inside the lane's context)
        ends[lane.index] = e

    start = torch.cuda.Event(enable_timing=True)
    start.record(fl.lanes[0].stream)     # nothing is pending anywhere: the time stamp is "now"
    fl.run(step_fn, steps, last=last, stagger_s=stagger_s)
    torch.cuda.synchronize()
    bc.barrier()
    torch.cuda.synchronize()
    return bc.max_over_ranks(max(start.elapsed_time(e) for e in ends.values()))


def run_nmpc(a, rank, world, local):
    import torch
    from vehicle_control_b200 import sharding
    from vehicle_control_b200.engine import Engine
    from vehicle_control_b200.environment.track import Track
    from vehicle_control_b200.pipeline import InFlight
    strong = a.scaling == "strong"
    B_total = (a.batch or 65536) * (1 if strong else world)
    cfg = controller_config()
    track_yaml = bc.load_yaml("environment/ippodromo.yaml")
    tr = Track(track_yaml)
    eng = Engine(bc.load_yaml("models/dynamic_car.yaml"), cfg, tr, device=local)
    N, M = cfg["horizon"], cfg["horizon_pm"]
    H = N + M
    t0 = time.perf_counter()
    cache = getattr(a, "lap_cache", None)
    if cache and os.path.exists(cache):      # profiling convenience: skip the (untimed) lap replay
        lap = dict(np.load(cache))
    else:
        lap = replay_lap(eng, cfg, tr)
        if cache and rank == 0:
            np.savez(cache, **lap)
    lap_s = time.perf_counter() - t0
    if strong:   # the configuration's batch, cut into contiguous shards (every rank builds the same batch)
        x0h, Xwh, Uwh, obh, _ = make_batch(lap, tr, B_total, SEED, cfg)
        lo, hi = sharding.shard_range(B_total, rank, world)
        x0h, Xwh, Uwh, obh = x0h[lo:hi], Xwh[lo:hi], Uwh[lo:hi], obh[lo:hi]
    else:
        x0h, Xwh, Uwh, obh, _ = make_batch(lap, tr, B_total // world, SEED + rank, cfg)
    B = len(x0h)
    x0, Xw, Uw, ob = eng.tensor(x0h), eng.tensor(Xwh), eng.tensor(Uwh), eng.tensor(obh)
    X, U = torch.empty_like(Xw), torch.empty_like(Uw)
    ds, kap = eng.empty(B, H), eng.empty(B, H)
    out = {"obj": eng.empty(B), "status": eng.empty(B, dtype=torch.int32), "iters": eng.empty(B, dtype=torch.int32),
           "kkt": eng.empty(B)}
    gathered = {}

    def step():
        X.copy_(Xw); U.copy_(Uw)                      # the warm start (solve is in place)
        eng.horizon_params(x0, X, ds, kap)            # _init_horizon
        eng.solve(x0, kap, ds, X, U, obstacles=ob, out=out)   # opti.solve()
        if strong and world > 1:                       # the one collective of the design: result gather (NCCL all_gather)
            gathered.update(sharding.gather_results({"u0": U[:, :, 0].contiguous(), "obj": out["obj"],
                                                     "status": out["status"], "iters": out["iters"]}, B_total))

    eng.set_profiling(True)
    l0 = eng.launch_count()
    with bc.ClockSampler(local) as clk:
        total_ms, per = bc.time_steps(step, a.steps, a.warmup)
    launches_total = eng.launch_count() - l0
    prof = eng.profile()
    eng.set_profiling(False)
    ms_single = total_ms / a.steps
    value_single = B_total / (ms_single * 1e-3)
    # ---- the same kernels with (nearly) every instance active in every launch: one profiled solve cut off after 6 iterations
    # (untimed).  The averages above include ~280 launches of the latency-bound tail with a few dozen instances each.
    eng.set_config(cfg, max_iter=6)
    eng.set_profiling(True)
    X.copy_(Xw); U.copy_(Uw)
    eng.horizon_params(x0, X, ds, kap)
    eng.solve(x0, kap, ds, X, U, obstacles=ob, out=out)
    torch.cuda.synchronize()
    prof_full = eng.profile()
    eng.set_profiling(False)
    eng.set_config(cfg)
    step()                                 # leaves X, U, out holding the converged solution again
    torch.cuda.synchronize()
    # ---- the headline: the same K steps with several batches in flight (lane = handle + stream + workspace + host thread;
    # the latency-bound tail of one batch runs under the full passes of the next; results are bit-identical)
    L = int(getattr(a, "in_flight", 0))
    if L <= 0:
        L = 6 if B >= 16384 else 8
    fl = None
    ms, value, launches_fl = ms_single, value_single, None
    if L > 1:
        # lanes spin while they wait for the GPU when every waiting thread of the box can have a core, else they sleep
        try:
            cores = len(os.sched_getaffinity(0))
        except AttributeError:
            cores = os.cpu_count() or 1
        ranks_here = int(os.environ.get("LOCAL_WORLD_SIZE", world))
        blocking = (L + 1) * ranks_here > cores
        if os.environ.get("VC_BENCH_SPIN") is not None:
            blocking = os.environ["VC_BENCH_SPIN"] == "0"
        fl = InFlight.create(bc.load_yaml("models/dynamic_car.yaml"), cfg, tr, local, L, blocking_sync=blocking)
        for ln in fl.lanes:
            with ln:
                ln.buf = {"X": torch.empty_like(Xw), "U": torch.empty_like(Uw), "ds": eng.empty(B, H), "kap": eng.empty(B, H),
                          "out": {"obj": eng.empty(B), "status": eng.empty(B, dtype=torch.int32),
                                  "iters": eng.empty(B, dtype=torch.int32), "kkt": eng.empty(B)}}
        torch.cuda.synchronize()

        def fstep(lane, i):
            b = lane.buf
            b["X"].copy_(Xw); b["U"].copy_(Uw)
            lane.eng.horizon_params(x0, b["X"], b["ds"], b["kap"])
            lane.eng.solve(x0, b["kap"], b["ds"], b["X"], b["U"], obstacles=ob, out=b["out"])
            if strong and world > 1:       # every rank issues the gathers in step order
                with fl.ordered(i):
                    gathered.update(sharding.gather_results({"u0": b["U"][:, :, 0].contiguous(), "obj": b["out"]["obj"],
                                                             "status": b["out"]["status"], "iters": b["out"]["iters"]}, B_total))

        fl.run(fstep, L)                   # allocates every lane's workspace
        lf0 = sum(ln.eng.launch_count() for ln in fl.lanes)
        with bc.ClockSampler(local) as clk:
            # offsetting the lanes' starts (pipeline.py: stagger_s) was measured at 0 / 100 / 150 / 200 / 300 ms: no effect
            # (663 ms per batch each time) -- the lanes are bound by the GPU's throughput, not by their phase
            stagger = float(os.environ.get("VC_BENCH_STAGGER", 0.0))
            total_fl = time_in_flight(fl, fstep, a.steps, max(a.warmup, L), stagger_s=stagger)
        launches_fl = sum(ln.eng.launch_count() for ln in fl.lanes) - lf0
        ms = total_fl / a.steps
        value = B_total / (ms * 1e-3)
        same = all(torch.equal(ln.buf["U"], U) and torch.equal(ln.buf["out"]["status"], out["status"]) for ln in fl.lanes)
        if not same:
            raise RuntimeError("in-flight lanes returned results that differ from the single-batch solve")
    status, iters = out["status"].cpu().numpy(), out["iters"].cpu().numpy()
    conv = float((status <= 1).mean())
    status_hist = {str(int(k)): int(v) for k, v in zip(*np.unique(status, return_counts=True))}
    it_pct = {f"p{q}": float(np.percentile(iters, q)) for q in (50, 90, 99, 99.9)}
    mean_it = float(iters.mean())
    runs = a.steps + a.warmup
    Ug, objg = U.cpu().numpy(), out["obj"].cpu().numpy()
    kap_h, ds_h = kap.cpu().numpy(), ds.cpu().numpy()
    # ---- e2e: `command()` from pinned host buffers through vc_nmpc_command_host: H2D of state, previous prediction and
    # obstacles, _init_horizon + solve on the device, D2H of the new prediction and the per-instance results, all timed
    Be = B if a.e2e_batch == 0 else min(B, a.e2e_batch)
    keep = []

    def hp(arr):   # page-locked host copy (numpy view of a pinned torch tensor)
        t = torch.from_numpy(np.ascontiguousarray(arr)).pin_memory()
        keep.append(t)
        return t.numpy()

    hx0, hob, hX, hU = hp(x0h[:Be]), hp(obh[:Be]), hp(Xwh[:Be]), hp(Uwh[:Be])
    eng.command_host(hx0, hX, hU, obstacles=hob)   # warm-up (allocates the staging arena)
    bc.barrier()
    e2e_t = 0.0
    e2e_n = max(1, min(a.steps, 3))
    for _ in range(e2e_n):
        hX[:], hU[:] = Xwh[:Be], Uwh[:Be]
        t1 = time.perf_counter()
        eng.command_host(hx0, hX, hU, obstacles=hob)
        e2e_t += time.perf_counter() - t1
    e2e_single_s = bc.max_over_ranks(e2e_t / e2e_n)
    e2e_s, e2e_steps = e2e_single_s, e2e_n
    if fl is not None:
        # the same call from the lanes: every step has its own pinned in/out buffers (filled before the clock starts), the
        # copies of one lane overlap the solve of another
        e2e_steps = max(L, min(a.steps, 6))
        hXs = [hp(Xwh[:Be]) for _ in range(e2e_steps)]
        hUs = [hp(Uwh[:Be]) for _ in range(e2e_steps)]
        res = [None] * e2e_steps

        def estep(lane, i):
            res[i] = lane.eng.command_host(hx0, hXs[i], hUs[i], obstacles=hob)

        wX, wU = [hp(Xwh[:Be]) for _ in range(L)], [hp(Uwh[:Be]) for _ in range(L)]
        fl.run(lambda lane, i: lane.eng.command_host(hx0, wX[i], wU[i], obstacles=hob), L)   # staging arenas
        del wX, wU
        torch.cuda.synchronize()
        bc.barrier()
        t1 = time.perf_counter()
        fl.run(estep, e2e_steps, stagger_s=stagger)   # vc_nmpc_command_host returns after its D2H copies have completed
        e2e_s = bc.max_over_ranks((time.perf_counter() - t1) / e2e_steps)
        if Be == B and not all(np.array_equal(r["status"], status) and np.array_equal(u, Ug) for r, u in zip(res, hUs)):
            raise RuntimeError("in-flight host-buffer calls returned results that differ from the device-resident solve")
    Be_total = int(bc.sum_over_ranks(Be))
    h2d = hx0.nbytes + hob.nbytes + hX.nbytes + hU.nbytes
    d2h = hX.nbytes + hU.nbytes + Be * (8 + 8 + 4 + 4)
    # ---- roofline of the dominant kernel: algorithmic work / measured launch time (CUDA events on the solve stream)
    peak_hbm, how = bc.measured_peaks()
    ms_st, ms_pm, ms_step = prof["ms"]
    n_st, n_step = max(1, prof["launches"][0]), max(1, prof["launches"][2])
    share = [m / max(1e-9, sum(prof["ms"])) for m in prof["ms"]]
    inst_passes = max(1, prof["instance_passes"])          # sum over timed+warm-up passes of the active instances
    fp64_peak = bc.FP64_PEAK_TFLOPS
    nu = ncu_units() or {}
    # units per launch: one launch of k_step (k_step_a -> k_step_b -> k_update, timed together with CUDA events on the
    # solve stream) advances every unfinished instance by one interior-point iteration; one launch of k_eval_st
    # evaluates N stages of every unfinished instance
    kern = {
        "k_step": {"units_per_launch": inst_passes / n_step, "unit": "instance-iteration", "bytes_per_unit": STEP_BYTES(H),
                   "flops_per_unit": N * 2.0e3 + M * 0.8e3 + (N * 0.73e3 + M * 0.1e3), "launch_ms": ms_step / n_step,
                   "ncu_dram_bytes_per_unit": (nu.get("k_step") or {}).get("dram_bytes_per_unit"),
                   "executed_flops_per_unit": (nu.get("k_step") or {}).get("executed_flops_per_unit")},
        "k_eval_st": {"units_per_launch": inst_passes * N / n_st, "unit": "stage evaluation", "bytes_per_unit": EVAL_ST_BYTES,
                      "flops_per_unit": 15.0e3, "launch_ms": ms_st / n_st,
                      "ncu_dram_bytes_per_unit": (nu.get("k_eval_st") or {}).get("dram_bytes_per_unit"),
                      "executed_flops_per_unit": (nu.get("k_eval_st") or {}).get("executed_flops_per_unit")},
    }
    dom = "k_step" if ms_step >= ms_st else "k_eval_st"
    kd = kern[dom]
    per_s = lambda v, per_unit: None if per_unit is None else v["units_per_launch"] * per_unit / (v["launch_ms"] * 1e-3)
    ach = per_s(kd, kd["bytes_per_unit"]) / 1e9
    solve_flops = mean_it * flops_per_iteration(N, M)
    roof = {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak_hbm, "unit": "GB/s", "frac": ach / peak_hbm,
            "traffic": None if kd["ncu_dram_bytes_per_unit"] is None else kd["units_per_launch"] * kd["ncu_dram_bytes_per_unit"],
            "peak_source": how,
            "traffic_source": ("dram__bytes_read+write per unit from the committed ncu --set full capture (%s%s) times this "
                               "run's average units per launch" % (nu.get("file"), ", STALE: kernel sources changed since" if nu.get("stale") else "")
                               if nu else "no committed ncu capture"),
            "units_per_launch": kd["units_per_launch"], "unit_of_work": kd["unit"],
            "algorithmic_bytes_per_unit": kd["bytes_per_unit"], "launch_ms": kd["launch_ms"],
            "share_of_step": share[2] if dom == "k_step" else share[0],
            "note": "the step kernels are bound by neither roof: they are latency / instruction-issue bound (sequential "
                    "Riccati and line-search chains, DESIGN.md section 4); the HBM fraction is reported as the contract asks",
            "fp64": {"achieved_tflops": per_s(kd, kd["flops_per_unit"]) / 1e12,
                     "executed_tflops": None if kd["executed_flops_per_unit"] is None else per_s(kd, kd["executed_flops_per_unit"]) / 1e12,
                     "peak_tflops": fp64_peak, "peak_source": "measured DFMA micro-benchmark (profiles/r01_fp64_peak.txt)"},
            "kernels": {k: {"launch_ms": v["launch_ms"], "units_per_launch": v["units_per_launch"],
                            "hbm_gbs": per_s(v, v["bytes_per_unit"]) / 1e9,
                            "fp64_tflops": per_s(v, v["flops_per_unit"]) / 1e12,
                            "fp64_executed_tflops": None if v["executed_flops_per_unit"] is None else per_s(v, v["executed_flops_per_unit"]) / 1e12}
                        for k, v in kern.items()},
            "whole_solve_fp64": {"achieved_tflops": value / world * solve_flops / 1e12,
                                 "frac": value / world * solve_flops / 1e12 / fp64_peak, "flops_per_solve": solve_flops},
            "kernel_share": {"k_eval_st": share[0], "k_eval_pm": share[1], "k_step": share[2]}}
    roof["fp64"]["frac"] = roof["fp64"]["achieved_tflops"] / fp64_peak
    fu = max(1, prof_full["instance_passes"]) / max(1, prof_full["launches"][2])      # instances per launch
    f_ms = prof_full["ms"][2] / max(1, prof_full["launches"][2])
    f_st_ms = prof_full["ms"][0] / max(1, prof_full["launches"][0])
    roof["full_launches"] = {
        "note": "the same kernels in a solve cut off after 6 iterations: (nearly) all instances active in every launch",
        "k_step": {"units_per_launch": fu, "launch_ms": f_ms, "hbm_gbs": fu * STEP_BYTES(H) / (f_ms * 1e-3) / 1e9,
                   "frac": fu * STEP_BYTES(H) / (f_ms * 1e-3) / 1e9 / peak_hbm},
        "k_eval_st": {"units_per_launch": fu * N, "launch_ms": f_st_ms, "hbm_gbs": fu * N * EVAL_ST_BYTES / (f_st_ms * 1e-3) / 1e9,
                      "fp64_tflops": fu * N * 15.0e3 / (f_st_ms * 1e-3) / 1e12,
                      "fp64_frac": fu * N * 15.0e3 / (f_st_ms * 1e-3) / 1e12 / fp64_peak}}
    line = {
        "metric": "NMPC solves/sec", "value": value, "unit": "solves/s", "n_gpus": world, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": a.scaling, "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": "C3: cascaded NMPC 20+35 with obstacles, ippodromo, batch 65,536 warm-started perturbed instances",
                   "batch_total": B_total, "batch_per_gpu": B, "in_flight_batches": L,
                   "in_flight": ("step i runs on lane i mod %d (handle + stream + workspace + host thread each): the latency-bound tail of "
                                 "one batch overlaps the full passes of the next; results bit-identical to single-batch solves" % L) if fl is not None else "one batch at a time",
                   "lane_stagger_ms": None if fl is None else 1e3 * stagger,
                   "host_wait": None if fl is None else ("sleep on a blocking-sync event" if blocking else "spin"),
                   "N": N, "M": M, "n_obstacles": int(obh.shape[1]), "tol": 1e-8,
                   "solver": {"max_iter": MAX_ITER, "max_soc": 4, "watchdog_trigger": 10, "note": "same on the CPU arm"},
                   "result_gather": "all_gather of (u0, obj, status, iters) inside the timed step" if strong and world > 1 else "none (one rank)" if world == 1 else "none (weak scaling)",
                   "gathered_rows": int(gathered["obj"].shape[0]) if gathered else None,
                   "converged_frac": conv, "status_counts": status_hist, "mean_iters": mean_it, "iters_percentiles": it_pct,
                   "max_iters": int(iters.max()), "passes": int(prof["last_passes"]),
                   "lap_ticks": int(len(lap["x0"])), "lap_replay_s": lap_s,
                   "l2": "per-step working set %.1f GB (solver workspace) >> 126 MB L2" % (eng.workspace_bytes() / 1e9)},
        "clocks": clk.summary(),
        "gpu_launches": int(round(launches_total * a.steps / runs if launches_fl is None else launches_fl * a.steps / (a.steps + max(a.warmup, L)))),
        "e2e": {"value": Be_total / e2e_s, "unit": "solves/s", "h2d_bytes_per_step": int(h2d),
                "d2h_bytes_per_step": int(d2h), "batch_per_gpu": Be, "steps": e2e_steps,
                "call": "vc_nmpc_command_host: H2D (state, previous prediction, obstacles), _init_horizon + solve on the device, D2H (prediction, results)"
                        + ("; %d calls in flight (one per lane, own pinned buffers per step)" % L if fl is not None else "")},
        "single_batch": {"value": value_single, "ms_per_step": ms_single, "e2e_value": Be_total / e2e_single_s,
                         "note": "one batch at a time on one handle (latency of a batch); kernel shares and the roofline block are "
                                 "measured in this mode"},
        "roofline": roof,
    }
    if rank == 0 and not a.no_cpu_baseline:
        import bench_reference as br
        n_cpu = min(B, a.cpu_instances or 512)
        sample = dict(x0=x0h[:n_cpu], Xw=Xwh[:n_cpu], Uw=Uwh[:n_cpu], obst=obh[:n_cpu], ds=ds_h[:n_cpu], kappa=kap_h[:n_cpu])
        line["cpu_baseline"] = br.cpu_baseline_nmpc(cfg, sample, Ug[:n_cpu], objg[:n_cpu])
    del X, U, Xw, Uw, fl
    return line
