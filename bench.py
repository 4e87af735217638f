"""bench.py -- throughput of the hot path on N B200s (one process per GPU).

    python bench.py --gpus 1 --steps K --warmup W [--workload nmpc|rollout] [--impl reference]

Workloads (BASELINE.json configs):
  nmpc     cascaded 20+35 NMPC with obstacles on ippodromo, one `command()`-equivalent batched solve per
           step from the warm start of the un-perturbed closed-loop tick (config C3; default once built)
  rollout  SingleTrack spatial RK4 rollout + [A|B] Jacobians, N=50 (config C5)
  c2       cascaded 20+15, 4,096 perturbed instances: warm-started and cold-started solves (config C2)
A "step" is one pass of the hot path over one batch.  `value` is device-resident throughput (CUDA events
on the launch stream), `e2e` goes through the host-buffer C-ABI call with H2D/D2H inside the timed region.
The instances shard across ranks with no data-path collective (weak scaling: per-GPU batch is fixed).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import bench_common as bc  # noqa: E402


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=None, choices=["nmpc", "rollout", "c2"])
    ap.add_argument("--batch", type=int, default=0, help="instances per GPU (0 = config default)")
    ap.add_argument("--e2e-batch", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-instances", type=int, default=0, help="instances per step of the CPU arm (0 = default sample)")
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"],
                    help="strong: the configuration's batch is split over the GPUs; weak: every GPU gets the whole batch")
    ap.add_argument("--lap-cache", default=None, help="npz file to save/load the untimed warm-start lap (profiling runs)")
    return ap.parse_args()


def main():
    a = parse()
    a.warmup = max(a.warmup, 3) if a.impl == "b200" else a.warmup
    workload = a.workload or bc.default_workload()
    if a.impl == "reference":
        from bench_reference import run_reference
        return run_reference(a, workload)
    import torch
    rank, world, local = bc.dist_setup(a.gpus)
    torch.cuda.set_device(local)
    if workload == "rollout":
        from bench_rollout import run_rollout
        line = run_rollout(a, rank, world, local)
    elif workload == "c2":
        from bench_c2 import run_c2
        line = run_c2(a, rank, world, local)
    else:
        from bench_nmpc import run_nmpc
        line = run_nmpc(a, rank, world, local)
    if rank == 0:
        print(json.dumps(line), flush=True)
    bc.dist_teardown()


if __name__ == "__main__":
    main()
