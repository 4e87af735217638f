"""GPU: per-instance tracks ("track segments" as an instance dimension).

A batch whose instances drive on different tracks (ippodromo, shoe, bean, gokart) through the track atlas must give,
bit for bit, what each track gives on its own: horizon parameters (a8 + a10), the NMPC solve (a9) and the plant
step (a11); a mixed lock-step lap finishes every vehicle on its own track length.
"""
import numpy as np
import pytest
import torch

from tests.conftest import load_yaml

pytestmark = pytest.mark.gpu
NAMES = ["ippodromo", "shoe", "bean", "gokart"]


def _setup(B):
    from vehicle_control_b200.engine import Engine
    from vehicle_control_b200.environment.track import Track
    from vehicle_control_b200.controllers.mpc.cascaded_mpc import cold_guess
    cfg = dict(load_yaml("controllers/cascaded.yaml")); cfg["obstacles"] = False
    tracks = [Track(load_yaml(f"environment/{n}.yaml")) for n in NAMES]
    eng = Engine(load_yaml("models/dynamic_car.yaml"), cfg, tracks[0])
    eng.set_track_atlas(tracks)
    rng = np.random.default_rng(7)
    tid = rng.integers(0, len(tracks), B).astype(np.int32)
    x0 = np.zeros((B, 8)); x0[:, 0] = 4.0 + rng.uniform(0, 2, B)
    x0[:, 4] = np.array([rng.uniform(1.0, 0.8 * tracks[t].length) for t in tid]); x0[:, 5] = rng.uniform(-1, 1, B)
    N, H = int(cfg["horizon"]), int(cfg["horizon"]) + int(cfg["horizon_pm"])
    np.random.seed(31)
    X0, U0 = cold_guess(N, H)
    return eng, cfg, tracks, tid, x0, X0, U0


def test_mixed_track_batch_equals_per_track_results():
    from vehicle_control_b200.engine import Engine
    B = 96
    eng, cfg, tracks, tid, x0h, X0, U0 = _setup(B)
    x0, tidd = eng.tensor(x0h), eng.tensor(tid, dtype=torch.int32)
    X, U = eng.tensor(np.repeat(X0[None], B, 0)), eng.tensor(np.repeat(U0[None], B, 0))
    ds, kap = eng.horizon_params(x0, X, track_id=tidd)
    out = eng.solve(x0, kap, ds, X, U)
    assert float((out["status"] <= 1).float().mean()) > 0.95
    xn = x0.clone()
    eng.plant_step(xn, U[:, :, 0].contiguous(), track_id=tidd)
    for t, tr in enumerate(tracks):
        sel = np.flatnonzero(tid == t)
        assert len(sel) > 0
        e1 = Engine(load_yaml("models/dynamic_car.yaml"), cfg, tr)
        x1 = e1.tensor(x0h[sel])
        X1, U1 = e1.tensor(np.repeat(X0[None], len(sel), 0)), e1.tensor(np.repeat(U0[None], len(sel), 0))
        ds1, kap1 = e1.horizon_params(x1, X1)
        st = torch.as_tensor(sel, device=ds.device)
        assert torch.equal(ds1, ds[st]) and torch.equal(kap1, kap[st])
        e1.solve(x1, kap1, ds1, X1, U1)
        assert torch.equal(U1, U[st]) and torch.equal(X1, X[st])
        xs = x1.clone()
        e1.plant_step(xs, U1[:, :, 0].contiguous())
        assert torch.equal(xs, xn[st])


def test_mixed_track_lockstep_lap():
    from vehicle_control_b200.simulation.racing import BatchedRacingSimulator
    B = 32
    eng, cfg, tracks, tid, x0h, X0, U0 = _setup(B)
    x0 = np.zeros((B, 8)); x0[:, 0] = 4.0; x0[:, 4] = 1.0
    np.random.seed(31)
    sim = BatchedRacingSimulator(load_yaml("models/dynamic_car.yaml"), tracks[0], cfg, x0, record=False,
                                 tracks=tracks, track_id=tid)
    n = 0
    while n < 1500 and sim.update() > 0:
        n += 1
    x = sim.x.cpu().numpy()
    L = np.array([tracks[t].length for t in tid])
    assert not bool(sim.failed.any())
    assert (x[:, 4] > L - 0.1).all()
    # vehicles on the same track started from the same state: identical laps; different tracks: different lap times
    laps = {t: np.unique(x[tid == t, 7]) for t in range(len(tracks))}
    assert all(len(v) == 1 for v in laps.values())
    print({NAMES[t]: float(v[0]) for t, v in laps.items()}, "ticks", n, "retries", sim.controller.retries)
    assert len({float(v[0]) for v in laps.values()}) == len(tracks)
