"""CPU: the N>1 path (contiguous instance shards + result gather) on the gloo backend, world size 2."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from vehicle_control_b200 import sharding


def test_shard_range_partitions_exactly():
    for n in (0, 1, 7, 64, 65536, 65537):
        for w in (1, 2, 3, 8):
            spans = [sharding.shard_range(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(w - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.shard_range(10, 2, 2)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, n, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        full_u = torch.arange(n * 2, dtype=torch.float64).reshape(n, 2)
        full_s = (torch.arange(n) % 3).to(torch.int32)
        # each rank "solves" only its shard (stand-in: the shard of the known full answer, tagged by rank)
        local = {"u0": sharding.shard(full_u).clone(), "status": sharding.shard(full_s).clone(),
                 "owner": torch.full((sharding.shard(full_s).shape[0],), rank, dtype=torch.int32)}
        got = sharding.gather_results(local, n)
        lo, hi = sharding.shard_range(n, rank, world)
        ok = (torch.equal(got["u0"], full_u) and torch.equal(got["status"], full_s)
              and got["owner"].shape[0] == n and bool((got["owner"][lo:hi] == rank).all()))
        # the timing reduction bench.py uses (max over ranks)
        t = torch.tensor([float(rank + 1)], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        q.put((rank, ok and float(t) == world))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n", [8, 9, 1])
def test_gather_results_world2_gloo(n):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, n, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=120) for _ in ps)
    for p in ps:
        p.join(60)
    assert res == [(0, True), (1, True)]


def _lanes_worker(rank, world, port, n, steps, q):
    """the strong-scaling bench protocol on the CPU: every rank runs its steps on three lanes (host threads) that finish
    out of order, and issues the result gather of step i inside InFlight.ordered(i) -- the collectives must pair up"""
    import random
    import time
    from vehicle_control_b200.pipeline import InFlight, Lane
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        fl = InFlight([Lane(i, engine=None, stream=None) for i in range(3)])
        rnd = random.Random(100 + rank)              # different delays on every rank
        got = [None] * steps

        def step(lane, i):
            time.sleep(0.001 * rnd.randint(0, 20))
            full = torch.arange(n, dtype=torch.float64) * (i + 1)      # step i's "solution"
            with fl.ordered(i):
                got[i] = sharding.gather_results({"obj": sharding.shard(full).clone()}, n)["obj"]

        fl.run(step, steps)
        ok = all(torch.equal(got[i], torch.arange(n, dtype=torch.float64) * (i + 1)) for i in range(steps))
        q.put((rank, ok))
    finally:
        dist.destroy_process_group()


def test_in_flight_lanes_gather_in_step_order_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_lanes_worker, args=(r, 2, port, 11, 10, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=180) for _ in ps)
    for p in ps:
        p.join(60)
    assert res == [(0, True), (1, True)]
