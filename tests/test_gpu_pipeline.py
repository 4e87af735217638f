"""GPU: several batches in flight (vehicle_control_b200/pipeline.py).  Lanes (handle + non-blocking stream + workspace + host
thread each) solve sub-batches of the C3 bench fixtures concurrently, device-resident and through the host-buffer call
vc_nmpc_command_host; every step must return, bit for bit, what one handle returns solving the same sub-batch alone, and
the oracle's controls to 1e-6."""
import numpy as np
import pytest
import torch

from tests.conftest import load_yaml
from tests.test_gpu_bench_parity import _fx, _rel

pytestmark = pytest.mark.gpu


def test_lanes_return_the_single_handle_results():
    import bench_nmpc as bn
    from vehicle_control_b200.engine import Engine
    from vehicle_control_b200.environment.track import Track
    from vehicle_control_b200.pipeline import InFlight
    f = _fx("c3")
    cfg = bn.controller_config()
    tr = Track(load_yaml("environment/ippodromo.yaml"))
    car = load_yaml("models/dynamic_car.yaml")
    n = len(f["x0"])
    steps, L = 7, 3
    rng = np.random.default_rng(5)
    subs = [np.sort(rng.choice(n, size=int(rng.integers(5, n)), replace=False)) for _ in range(steps)]   # ragged sub-batches
    # ---- one handle, one batch at a time (default stream)
    eng = Engine(car, cfg, tr)
    ref = []
    for idx in subs:
        x0, X, U, ob = (eng.tensor(f[k][idx]) for k in ("x0", "Xw", "Uw", "obst"))
        ds, kap = eng.horizon_params(x0, X)
        out = eng.solve(x0, kap, ds, X, U, obstacles=ob)
        ref.append((U.cpu().numpy(), X.cpu().numpy(), out["status"].cpu().numpy(), out["iters"].cpu().numpy()))
        assert _rel(ref[-1][0], f["U"][idx]).max() < 1e-6
    # ---- three lanes, device-resident
    fl = InFlight.create(car, cfg, tr, torch.cuda.current_device(), L)
    got = [None] * steps

    def step(lane, i):
        idx = subs[i]
        e = lane.eng
        x0, X, U, ob = (e.tensor(f[k][idx]) for k in ("x0", "Xw", "Uw", "obst"))
        ds, kap = e.horizon_params(x0, X)
        out = e.solve(x0, kap, ds, X, U, obstacles=ob)
        with fl.ordered(i):
            got[i] = (U.cpu().numpy(), X.cpu().numpy(), out["status"].cpu().numpy(), out["iters"].cpu().numpy())

    fl.run(step, steps)
    fl.synchronize()
    for a, b in zip(ref, got):
        assert all(np.array_equal(x, y) for x, y in zip(a, b))
    # ---- three lanes, host buffers in and out (the e2e call)
    hx = [(np.ascontiguousarray(f["x0"][idx]), np.ascontiguousarray(f["Xw"][idx]), np.ascontiguousarray(f["Uw"][idx]),
           np.ascontiguousarray(f["obst"][idx])) for idx in subs]
    res = [None] * steps

    def hstep(lane, i):
        x0, X, U, ob = hx[i]
        res[i] = lane.eng.command_host(x0, X, U, obstacles=ob)

    fl.run(hstep, steps)
    for (U, X, st, it), (_, Xh, Uh, _), r in zip(ref, hx, res):
        assert np.array_equal(Uh, U) and np.array_equal(Xh, X) and np.array_equal(r["status"], st) and np.array_equal(r["iters"], it)


def test_fleet_in_lanes_equals_single_simulator(golden):
    """A closed-loop fleet split over lanes (pipeline.run_fleet_in_lanes) ends, vehicle for vehicle and bit for bit, where
    the single-handle BatchedRacingSimulator ends after the same number of ticks."""
    from vehicle_control_b200.pipeline import run_fleet_in_lanes
    from vehicle_control_b200.simulation.racing import BatchedRacingSimulator
    _, meta = golden
    cfg = meta["cascaded1_ippodromo/cascaded"]["config"]
    B, ticks = 24, 25
    rng = np.random.default_rng(1004)
    x0 = np.zeros((B, 8)); x0[:, 0] = 4.0; x0[:, 4] = 1.0 + rng.uniform(0, 5, B); x0[:, 5] = rng.uniform(-1, 1, B)
    car, trk = load_yaml("models/dynamic_car.yaml"), load_yaml("environment/ippodromo.yaml")
    np.random.seed(31)
    one = BatchedRacingSimulator(car, trk, cfg, x0, record=False)
    assert one.run(ticks) == ticks
    sims = run_fleet_in_lanes(car, trk, cfg, x0, 3, device=torch.cuda.current_device(), max_ticks=ticks)
    assert [s.B for s in sims] == [8, 8, 8]
    x = torch.cat([s.x for s in sims]).cpu().numpy()
    assert np.array_equal(x, one.x.cpu().numpy())
    assert sum(int(s.total_solves) for s in sims) == int(one.total_solves)
