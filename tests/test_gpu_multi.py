"""GPU: the N>1 path on NCCL (needs two GPUs; skipped on a one-GPU box) and the batched simulator's run format.

test_gather_results_world2_nccl  two ranks, one GPU each: every rank solves its contiguous shard of a small C3 batch
    through the C-ABI and the per-instance results are all-gathered over NCCL (sharding.gather_results); the gathered
    arrays equal the one-GPU solve of the whole batch bit for bit (an instance's result does not depend on its shard).
"""
import os
import socket

import numpy as np
import pytest
import torch

from tests.conftest import load_yaml

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, q):
    import torch.distributed as dist
    from tests.test_gpu_bench_parity import _engine, _fx
    from vehicle_control_b200 import sharding
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        f = _fx("c3")
        n = len(f["x0"])
        lo, hi = sharding.shard_range(n, rank, world)
        from vehicle_control_b200.engine import Engine
        from vehicle_control_b200.environment.track import Track
        import bench_nmpc as bn
        eng = Engine(load_yaml("models/dynamic_car.yaml"), bn.controller_config(), Track(load_yaml("environment/ippodromo.yaml")), device=rank)
        x0, X, U = eng.tensor(f["x0"][lo:hi]), eng.tensor(f["Xw"][lo:hi]), eng.tensor(f["Uw"][lo:hi])
        ds, kap = eng.horizon_params(x0, X)
        out = eng.solve(x0, kap, ds, X, U, obstacles=eng.tensor(f["obst"][lo:hi]))
        got = sharding.gather_results({"u0": U[:, :, 0].contiguous(), "obj": out["obj"], "status": out["status"],
                                       "iters": out["iters"]}, n)
        ok = got["u0"].shape == (n, 2) and bool((got["status"] == 0).all())
        eu = (np.abs(got["u0"].cpu().numpy() - f["U"][:, :, 0]) / np.maximum(1, np.abs(f["U"][:, :, 0]))).max()
        q.put((rank, bool(ok and eu < 1e-6), got["u0"].cpu().numpy().tobytes() if rank == 0 else None))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (gpurun --gpus 2)")
def test_gather_results_world2_nccl():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted((q.get(timeout=300) for _ in ps), key=lambda r: r[0])
    for p in ps:
        p.join(60)
    assert [r[:2] for r in res] == [(0, True), (1, True)]
    # one GPU, whole batch: identical bits
    from tests.test_gpu_bench_parity import _engine, _fx
    f = _fx("c3")
    eng, _ = _engine("c3")
    x0, X, U = eng.tensor(f["x0"]), eng.tensor(f["Xw"]), eng.tensor(f["Uw"])
    ds, kap = eng.horizon_params(x0, X)
    eng.solve(x0, kap, ds, X, U, obstacles=eng.tensor(f["obst"]))
    assert U[:, :, 0].contiguous().cpu().numpy().tobytes() == res[0][2]


def test_batched_simulator_run_format_and_summary(golden, tmp_path):
    """BatchedRacingSimulator: finished vehicles are not solved again (compaction), save() / load() round-trip, the
    batched summary equals the single-vehicle summary (racing.py:92-106) for vehicle 0, export_vehicle() writes the
    reference's five files per vehicle (racing.py:441-450)."""
    from vehicle_control_b200.simulation.racing import BatchedRacingSimulator, RacingSimulator
    _, meta = golden
    cfg = dict(meta["cascaded1_ippodromo/cascaded"]["config"])
    B = 6
    x0 = np.zeros((B, 8)); x0[:, 0] = 4.0; x0[:, 4] = 1.0
    x0[1:, 4] = 295.0 + 3.0 * np.arange(1, B)      # vehicles 1.. start near the finish line and stop early
    np.random.seed(31)
    sim = BatchedRacingSimulator(load_yaml("models/dynamic_car.yaml"), load_yaml("environment/ippodromo.yaml"), cfg, x0)
    solves = []
    for _ in range(60):
        solves.append(sim.update())
    assert solves[0] == B and solves[-1] == 1 and sim.total_solves == sum(solves)
    T = sim.ticks.cpu().numpy()
    assert T[0] == 60 and (T[1:] < 60).all() and (np.diff(T[1:]) <= 0).all()
    s = sim.summary()
    assert s["fleet"]["finished"] == B - 1 and s["fleet"]["failed"] == 0
    # vehicle 0 against the single-vehicle simulator's summary of the same 60 ticks
    np.random.seed(31)
    one = RacingSimulator({"track_name": "ippodromo", "controller_names": ["cascaded"], "obstacles": False, "slip_angle_plot": False,
                           "name": "t", "save_gif": False, "save_images": False, "save_data": False, "load": False, "logging": False},
                          load_yaml("models/dynamic_car.yaml"), load_yaml("environment/ippodromo.yaml"), controller_configs=[cfg])
    for n in range(60):
        one.update(n)
    so = one.summary()["cascaded"]
    for k in ("laptime", "average_speed", "mean_abs_Fx", "std_Fx", "mean_squared_ey"):
        assert abs(s[k][0] - so[k]) <= 1e-9 * max(1.0, abs(so[k])), (k, s[k][0], so[k])
    sim.save(str(tmp_path), "fleet")
    z, c2 = BatchedRacingSimulator.load(str(tmp_path), "fleet")
    assert z["state_traj"].shape == (61, B, 8) and z["action_traj"].shape == (61, B, 2) and np.array_equal(z["ticks"], T)
    assert int(c2["horizon"]) == int(cfg["horizon"])
    sim.export_vehicle(2, str(tmp_path), "cascaded")
    st = np.load(tmp_path / "cascaded_state_traj.npy"); ac = np.load(tmp_path / "cascaded_action_traj.npy")
    pr = np.load(tmp_path / "cascaded_preds.npy")
    assert st.shape == (T[2] + 1, 8) and ac.shape == (T[2] + 1, 2) and pr.shape == (T[2], 1, 3) and np.all(ac[0] == 0)
    assert st[-1, 4] > sim.track.length - 0.1 and os.path.exists(tmp_path / "cascaded_config.yaml")
