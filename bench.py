"""bench.py -- throughput of the hot path on N B200s (one process per GPU).

    python bench.py --gpus 1 --steps K --warmup W [--workload nmpc|rollout] [--impl reference]

Workloads (BASELINE.json configs):
  nmpc     cascaded 20+35 NMPC with obstacles on ippodromo, batch 65,536: one `command()`-equivalent batched solve per
           step from the warm start of the un-perturbed closed-loop tick (config C3; the default line, which also
           carries the C5 numbers under "secondary.rollout")
  rollout  SingleTrack spatial RK4 rollout + [A|B] Jacobians, 1M trajectories x N=50 (config C5)
  c2       cascaded 20+15, 4,096 perturbed instances: warm-started and cold-started solves (config C2)
A "step" is one pass of the hot path over one batch.  `value` is device-resident throughput (CUDA events
on the launch stream), `e2e` goes through the host-buffer C-ABI call with H2D/D2H inside the timed region.
--scaling strong (default): the configuration's batch is cut into contiguous shards, one per GPU, and the per-instance
results are all-gathered (NCCL) inside the timed step; --scaling weak: every GPU gets the whole batch, no collective.
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
    ap.add_argument("--steps", type=int, default=12)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=None, choices=["nmpc", "rollout", "c2"])
    ap.add_argument("--batch", type=int, default=0, help="instances per GPU (0 = config default)")
    ap.add_argument("--e2e-batch", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-instances", type=int, default=0, help="instances per step of the CPU arm (0 = default sample)")
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"],
                    help="strong: the configuration's batch is split over the GPUs; weak: every GPU gets the whole batch")
    ap.add_argument("--in-flight", type=int, default=0, dest="in_flight",
                    help="C3: batches in flight per GPU (lanes: handle + stream + workspace + host thread each); 1 = one batch "
                         "at a time; 0 = by the per-GPU batch (6 at >= 16,384 instances, else 8: the tail of a batch lasts as long "
                         "whatever its size, so smaller batches need more of them in flight to hide it)")
    ap.add_argument("--lap-cache", default=None, help="npz file to save/load the untimed warm-start lap (profiling runs)")
    ap.add_argument("--rollout-batch", type=int, default=0, help="trajectories of the C5 leg (0 = 1,048,576)")
    ap.add_argument("--no-secondary", action="store_true", help="skip the C5 block of the default line")
    return ap.parse_args()


def main():
    a = parse()
    # ONE JSON line on stdout is the contract: whatever libraries print to fd 1 (NCCL's version banner / debug lines, a
    # stray print) is sent to stderr for the whole run, and the line is written to the saved descriptor at the end
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    emit = lambda obj: os.write(real_stdout, (json.dumps(obj) + "\n").encode())
    a.warmup = max(a.warmup, 3) if a.impl == "b200" else a.warmup
    workload = a.workload or bc.default_workload()
    if a.impl == "reference":
        from bench_reference import run_reference
        return run_reference(a, workload, emit)
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
        if not a.no_secondary and a.workload is None:
            # the second half of BASELINE.json's metric ("rollout+Jacobian states/sec") in the same driver-run record
            import gc
            gc.collect(); torch.cuda.empty_cache()
            from bench_rollout import run_rollout
            r = run_rollout(a, rank, world, local, steps=min(a.steps, 3), warmup=3)
            line["secondary"] = {"rollout": {k: r[k] for k in ("metric", "value", "unit", "ms_per_step", "steps", "warmup", "scaling",
                                                                "config", "e2e", "roofline", "gpu_launches", "clocks", "cpu_baseline") if k in r}}
    if rank == 0:
        emit(line)
    bc.dist_teardown()


if __name__ == "__main__":
    main()
