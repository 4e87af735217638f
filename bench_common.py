"""Shared helpers of bench.py: clocks sampling, distributed setup, synthetic workloads."""
from __future__ import annotations

import json
import os
import subprocess
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
FP64_PEAK_TFLOPS = 37.0   # measured DFMA peak of this pool's B200 (scripts/fp64_peak.cu, profiles/r01_fp64_peak.txt)


def load_yaml(rel):
    import yaml
    with open(os.path.join(ROOT, "vehicle_control_b200", "config", rel)) as f:
        return yaml.safe_load(f)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d.get("hbm_gbs", 6650.0)), "measured"
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0, period=0.1):
        self.index, self.period, self.rows, self._stop, self._t = index, period, [], threading.Event(), None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                f = [x.strip() for x in out.strip().split(",")]
                if len(f) >= 6:
                    self.rows.append(f)
            except Exception:
                pass
            self._stop.wait(self.period)

    def __enter__(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=10)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
                "samples": len(self.rows)}


def rollout_inputs(B, N, seed, track, states):
    """C5 inputs: golden-state rows + perturbation, piecewise-constant controls."""
    rng = np.random.default_rng(seed)
    x0 = states[rng.integers(0, len(states), B)].copy()
    x0[:, 0] = np.maximum(x0[:, 0] + rng.uniform(-.5, .5, B), 3.2)
    x0[:, 1] += rng.normal(0, .05, B)
    x0[:, 2] += rng.normal(0, .02, B)
    x0[:, 3] = np.clip(x0[:, 3] + rng.uniform(-.02, .02, B), -.4, .4)
    x0[:, 5] += rng.uniform(-.5, .5, B)
    x0[:, 6] += rng.normal(0, .01, B)
    x0[:, 7] = 0.0
    nseg = (N + 4) // 5
    U = np.empty((B, N, 2))
    U[..., 0] = np.repeat(rng.uniform(-3000, 3000, (B, nseg)), 5, 1)[:, :N]
    U[..., 1] = np.repeat(rng.uniform(-.15, .15, (B, nseg)), 5, 1)[:, :N]
    ds = np.repeat(0.03 * x0[:, :1], N, 1)
    kap = track.k(x0[:, 4:5] + np.cumsum(ds, 1) - ds)
    return x0, U, kap, ds


def golden_states(track_name="ippodromo"):
    z = np.load(os.path.join(ROOT, "tests", "golden", "runs.npz"))
    with open(os.path.join(ROOT, "tests", "golden", "runs.json")) as f:
        meta = json.load(f)
    rows = [z[k + "/state"][5:-1] for k, m in sorted(meta.items()) if m["track"] == track_name]
    return np.concatenate(rows, 0)


# ------------------------------------------------------------------------------------------------
# distributed plumbing (one process per GPU; instances shard with no data-path collective)
def default_workload():
    return "nmpc"


def dist_setup(gpus):
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        import torch
        # one JSON line on stdout is the contract: NCCL writes its version banner / debug lines to stdout unless told otherwise
        os.environ.setdefault("NCCL_DEBUG", "WARN")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        if torch.cuda.is_available():
            torch.cuda.set_device(local)
            dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
        else:
            dist.init_process_group("gloo", rank=rank, world_size=world)
    return rank, world, local


def dist_teardown():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()


def barrier():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        dist.barrier()


def max_over_ranks(x: float) -> float:
    import torch
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        t = torch.tensor([x], dtype=torch.float64, device="cuda" if torch.cuda.is_available() else "cpu")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    return x


def sum_over_ranks(x: float) -> float:
    import torch
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        t = torch.tensor([x], dtype=torch.float64, device="cuda" if torch.cuda.is_available() else "cpu")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())
    return x


class L2Flusher:
    """Writes a buffer larger than the 126 MB L2 between timed iterations."""

    def __init__(self, device, mib=256):
        import torch
        self.buf = torch.empty(mib << 20, dtype=torch.uint8, device=device)
        self.k = 0

    def __call__(self):
        self.k = (self.k + 1) & 0x7F
        self.buf.fill_(self.k)


def time_steps(step_fn, steps, warmup, flush=None, per_step_events=True):
    """W untimed warm-up steps, then exactly K timed steps between barrier+synchronize on both sides.
    Returns (total_ms over the K steps excluding flushes, max over ranks; list of per-step ms on this rank)."""
    import torch
    for _ in range(warmup):
        if flush:
            flush()
        step_fn()
    torch.cuda.synchronize()
    barrier()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for i in range(steps):
        if flush:
            flush()
        ev[i][0].record()
        step_fn()
        ev[i][1].record()
    torch.cuda.synchronize()
    barrier()
    torch.cuda.synchronize()
    per = [a.elapsed_time(b) for a, b in ev]
    return max_over_ranks(sum(per)), per
