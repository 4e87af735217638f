"""bench leg C5: SingleTrack spatial RK4 rollout + [A|B] Jacobians (BASELINE.json configs[4])."""
from __future__ import annotations

import os
import time

import numpy as np

import bench_common as bc

# algorithmic bytes / flops of one state-step with Jacobian
BYTES_PER_STEP = 736.0
FLOPS_PER_STEP = 5.0e3


def cpu_baseline_rollout(track, states, N, seconds=12.0):
    """Oracle port (numpy values + torch fp64 autodiff Jacobians) on a bounded sample, 1 thread."""
    import torch
    from oracle import models as om
    torch.set_num_threads(1)
    B = 512
    x0, U, kap, ds = bc.rollout_inputs(B, N, 1005, track, states)
    done, t0 = 0, time.perf_counter()
    while True:
        X = om.st_rollout_spatial(om.DEFAULT, x0, U, kap, ds)
        om.st_step_jac(om.DEFAULT, X[:, :-1].reshape(-1, 8), U.reshape(-1, 2), kap.reshape(-1), ds.reshape(-1))
        done += B * N
        el = time.perf_counter() - t0
        if el > seconds:
            break
    return {"value": done / el, "unit": "state-steps/s", "cores": 1, "kind": "port",
            "sample": f"{done // N} trajectories x N={N} (oracle numpy RK4 + torch.func.jacfwd), {el:.1f} s"}


def run_rollout(a, rank, world, local):
    import torch
    from vehicle_control_b200.engine import Engine
    from vehicle_control_b200.environment.track import Track
    N = 50
    B = a.batch or (1 << 20)
    tr = Track(bc.load_yaml("environment/ippodromo.yaml"))
    eng = Engine(bc.load_yaml("models/dynamic_car.yaml"), bc.load_yaml("controllers/cascaded.yaml"), tr, device=local)
    states = bc.golden_states()
    # inputs generated in chunks on the host (synthetic, seed 1005 + rank), resident in HBM before timing
    chunk = min(B, 1 << 17)
    parts = [bc.rollout_inputs(min(chunk, B - o), N, 1005 + 7919 * rank + o, tr, states) for o in range(0, B, chunk)]
    x0, U, kap, ds = (eng.tensor(np.concatenate([p[i] for p in parts], 0)) for i in range(4))
    X, AB = eng.empty(B, N + 1, 8), eng.empty(B, N, 8, 10)
    l0 = eng.launch_count()

    def step():
        eng.rollout_st_into(x0, U, kap, ds, X, AB)

    with bc.ClockSampler(local) as clk:
        total_ms, per = bc.time_steps(step, a.steps, a.warmup)
    launches = eng.launch_count() - l0 - a.warmup
    ms = total_ms / a.steps
    value = world * B * N / (ms * 1e-3)
    finite = bool(torch.isfinite(X[:, -1]).all())
    # ---- e2e: host buffers through the same C-ABI call, copies inside the timed region
    Be = min(B, 1 << 16)
    hx = [t[:Be].cpu().pin_memory() for t in (x0, U, kap, ds)]
    hX, hAB = torch.empty(Be, N + 1, 8, dtype=torch.float64).pin_memory(), torch.empty(Be, N, 8, 10, dtype=torch.float64).pin_memory()
    dx = [eng.empty(*t.shape) for t in hx]

    def e2e_step():
        for d, h in zip(dx, hx):
            d.copy_(h, non_blocking=True)
        eng.rollout_st_into(dx[0], dx[1], dx[2], dx[3], X[:Be], AB[:Be])
        hX.copy_(X[:Be], non_blocking=True)
        hAB.copy_(AB[:Be], non_blocking=True)
        torch.cuda.synchronize()

    for _ in range(2):
        e2e_step()
    bc.barrier()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        e2e_step()
    e2e_s = bc.max_over_ranks(time.perf_counter() - t0)
    h2d = sum(t.numel() * 8 for t in hx)
    d2h = hX.numel() * 8 + hAB.numel() * 8
    peak, how = bc.measured_peaks()
    kern_ms = float(np.mean(per))
    achieved = B * N * BYTES_PER_STEP / (kern_ms * 1e-3) / 1e9
    line = {
        "metric": "rollout+Jacobian state-steps/sec", "value": value, "unit": "state-steps/s", "n_gpus": world,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "C5: 1M SingleTrack trajectories x N=50 spatial RK4 + [A|B] (ippodromo)",
                   "trajectories_per_gpu": B, "N": N, "l2": "outputs 32 GB per step, far larger than L2",
                   "finite": finite},
        "clocks": clk.summary(), "gpu_launches": int(launches),
        "e2e": {"value": world * Be * N * a.steps / e2e_s, "unit": "state-steps/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "trajectories_per_step": Be},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     # ncu --set full of k_rollout_st_jac_k<2> (profiles/r01_ncu_rollout_k2.md): 744.2 B of DRAM traffic
                     # and 12.83 kflop executed (DFMA x2 + DMUL + DADD) per state-step
                     "traffic": B * N * 744.2, "peak_source": how, "kernel": "k_rollout_st_jac_k<2>",
                     "traffic_source": "dram__bytes_read+write per state-step from the committed ncu capture x this run's units",
                     "fp64_gflops_algorithmic": B * N * FLOPS_PER_STEP / (kern_ms * 1e-3) / 1e9,
                     "fp64_gflops_executed": B * N * 12.83e3 / (kern_ms * 1e-3) / 1e9,
                     "fp64_peak_gflops": bc.FP64_PEAK_TFLOPS * 1e3},
    }
    if rank == 0 and not a.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline_rollout(tr, states, N)
    return line
