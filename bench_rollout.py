"""bench leg C5: SingleTrack spatial RK4 rollout + [A|B] Jacobians (BASELINE.json configs[4]: 1M trajectories x N=50).

`value`  device-resident throughput (inputs in HBM, CUDA events around vc_rollout_st).
`e2e`    the whole batch through the same C-ABI call from pinned host buffers, in chunks of 65,536 trajectories: H2D of
         the chunk's inputs, kernel, D2H of the chunk's X and [A|B] into a pinned staging buffer on a second stream (the
         copy of chunk i overlaps the kernel of chunk i+1).  32 GB of Jacobians cross PCIe per step: e2e is PCIe bound.
"""
from __future__ import annotations

import time

import numpy as np

import bench_common as bc

# algorithmic bytes / flops of one state-step with Jacobian
BYTES_PER_STEP = 736.0
FLOPS_PER_STEP = 5.0e3
CHUNK = 1 << 16


def cpu_baseline_rollout(track, states, N):
    """oracle/cpp (RK4 in first-order jets, OpenMP over trajectories) on a bounded sample, all host threads."""
    import bench_reference as br
    from oracle import cpp_oracle as co
    th = br.host_threads()
    nb = 16384
    x0, U, kap, ds = bc.rollout_inputs(nb, N, 1005, track, states)
    co.rollout_st(x0[:256], U[:256], kap[:256], ds[:256], jac=True, threads=th)
    done, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < 4.0:
        co.rollout_st(x0, U, kap, ds, jac=True, threads=th)
        done += nb * N
    el = time.perf_counter() - t0
    return {"value": done / el, "unit": "state-steps/s", "cores": th, "kind": "port-cpp",
            "sample": f"{nb} trajectories x N={N} per pass, oracle/cpp first-order jets, {th} threads, {el:.1f} s"}


def run_rollout(a, rank, world, local, steps=None, warmup=None):
    import torch
    from vehicle_control_b200 import sharding
    from vehicle_control_b200.engine import Engine
    from vehicle_control_b200.environment.track import Track
    steps, warmup = steps or a.steps, warmup or a.warmup
    N = 50
    strong = a.scaling == "strong"
    B_total = (a.rollout_batch or (1 << 20)) * (1 if strong else world)
    lo, hi = sharding.shard_range(B_total, rank, world)
    B = hi - lo
    tr = Track(bc.load_yaml("environment/ippodromo.yaml"))
    eng = Engine(bc.load_yaml("models/dynamic_car.yaml"), bc.load_yaml("controllers/cascaded.yaml"), tr, device=local)
    states = bc.golden_states()
    # inputs generated in chunks on the host (synthetic, seed 1005 + global chunk offset), resident in HBM before timing
    parts = [bc.rollout_inputs(min(CHUNK, hi - o), N, 1005 + o, tr, states) for o in range(lo, hi, CHUNK)]
    hx = [np.concatenate([p[i] for p in parts], 0) for i in range(4)]
    x0, U, kap, ds = (eng.tensor(h) for h in hx)
    X, AB = eng.empty(B, N + 1, 8), eng.empty(B, N, 8, 10)
    l0 = eng.launch_count()

    def step():
        eng.rollout_st_into(x0, U, kap, ds, X, AB)

    with bc.ClockSampler(local) as clk:
        total_ms, per = bc.time_steps(step, steps, warmup)
    launches = (eng.launch_count() - l0) * steps // (steps + warmup)
    ms = total_ms / steps
    value = B_total / world * world * N / (ms * 1e-3) if strong else B_total * N / (ms * 1e-3)
    finite = bool(torch.isfinite(X[:, -1]).all())
    # ---- e2e over the WHOLE batch from pinned host buffers, chunked
    pin = [torch.from_numpy(h).pin_memory() for h in hx]
    nb = min(B, CHUNK)
    stage = [(torch.empty(nb, N + 1, 8, dtype=torch.float64).pin_memory(), torch.empty(nb, N, 8, 10, dtype=torch.float64).pin_memory())
             for _ in range(2)]
    dx = [eng.empty(nb, *t.shape[1:]) for t in pin]
    side = torch.cuda.Stream(device=local)
    main = torch.cuda.current_stream(local)

    def e2e_step():
        for i, o in enumerate(range(0, B, nb)):
            n = min(nb, B - o)
            for d, h in zip(dx, pin):
                d[:n].copy_(h[o:o + n], non_blocking=True)
            eng.rollout_st_into(dx[0][:n], dx[1][:n], dx[2][:n], dx[3][:n], X[o:o + n], AB[o:o + n])
            ev = torch.cuda.Event(); ev.record(main)
            side.wait_event(ev)
            with torch.cuda.stream(side):
                stage[i % 2][0][:n].copy_(X[o:o + n], non_blocking=True)
                stage[i % 2][1][:n].copy_(AB[o:o + n], non_blocking=True)
        torch.cuda.synchronize()

    e2e_step()
    bc.barrier()
    e2e_n = max(1, min(steps, 2))
    t0 = time.perf_counter()
    for _ in range(e2e_n):
        e2e_step()
    e2e_s = bc.max_over_ranks((time.perf_counter() - t0) / e2e_n)
    h2d = sum(t.numel() * 8 for t in pin)
    d2h = B * (N + 1) * 8 * 8 + B * N * 80 * 8
    peak, how = bc.measured_peaks()
    kern_ms = float(np.mean(per))
    achieved = B * N * BYTES_PER_STEP / (kern_ms * 1e-3) / 1e9
    import bench_nmpc as bn
    nu = (bn.ncu_units() or {})
    ku = nu.get("k_rollout_st_jac") or {}
    line = {
        "metric": "rollout+Jacobian state-steps/sec", "value": B_total * N / (ms * 1e-3), "unit": "state-steps/s", "n_gpus": world,
        "steps": steps, "warmup": warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": a.scaling,
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "C5: 1M SingleTrack trajectories x N=50 spatial RK4 + [A|B] Jacobians (ippodromo)",
                   "trajectories_total": B_total, "trajectories_per_gpu": B, "N": N,
                   "l2": "outputs %.1f GB per step, far larger than L2" % (B * N * 736 / 1e9), "finite": finite},
        "clocks": clk.summary(), "gpu_launches": int(launches),
        "e2e": {"value": B_total * N / e2e_s, "unit": "state-steps/s", "h2d_bytes_per_step": int(h2d),
                "d2h_bytes_per_step": int(d2h), "trajectories_per_step_per_gpu": B, "chunk": nb,
                "note": "whole batch through vc_rollout_st from pinned host buffers in chunks; D2H of the Jacobians dominates (PCIe bound)"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": None if ku.get("dram_bytes_per_unit") is None else B * N * ku["dram_bytes_per_unit"],
                     "peak_source": how, "kernel": "k_rollout_st_jac_k<2>",
                     "traffic_source": "dram__bytes_read+write per state-step from the committed ncu capture (%s) x this run's units" % nu.get("file"),
                     "algorithmic_bytes_per_unit": BYTES_PER_STEP, "unit_of_work": "state-step",
                     "fp64": {"achieved_tflops": B * N * FLOPS_PER_STEP / (kern_ms * 1e-3) / 1e12,
                              "executed_tflops": None if ku.get("executed_flops_per_unit") is None else B * N * ku["executed_flops_per_unit"] / (kern_ms * 1e-3) / 1e12,
                              "peak_tflops": bc.FP64_PEAK_TFLOPS,
                              "frac": B * N * FLOPS_PER_STEP / (kern_ms * 1e-3) / 1e12 / bc.FP64_PEAK_TFLOPS}},
    }
    if rank == 0 and not a.no_cpu_baseline:
        from oracle.track import OracleTrack
        line["cpu_baseline"] = cpu_baseline_rollout(OracleTrack(bc.load_yaml("environment/ippodromo.yaml")), states, N)
    del X, AB
    torch.cuda.empty_cache()
    return line
