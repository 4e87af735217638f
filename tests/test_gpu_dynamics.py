"""GPU parity: rollout / Jacobian / track / plant kernels (through the C-ABI) against the oracle."""
import numpy as np
import pytest
import torch

from oracle import models as om
from tests.conftest import load_yaml

pytestmark = pytest.mark.gpu
RTOL = 1e-9   # BASELINE.json north_star: rollouts and Jacobians within 1e-9 relative


@pytest.fixture(scope="module")
def eng():
    from vehicle_control_b200.engine import Engine
    from vehicle_control_b200.environment.track import Track
    tr = Track(load_yaml("environment/ippodromo.yaml"))
    cfg = load_yaml("controllers/cascaded.yaml")
    cfg["horizon_pm"] = 35
    return Engine(load_yaml("models/dynamic_car.yaml"), cfg, tr)


def _rel(a, b):
    return float(np.max(np.abs(a - b) / np.maximum(1.0, np.abs(b))))


def _inputs(golden, oracle_tracks, B, N, seed):
    z, _ = golden
    rng = np.random.default_rng(seed)
    S = z["cascaded3_ippodromo/cascaded/state"][5:-1]
    x0 = S[rng.integers(0, len(S), B)].copy()
    x0[:, 0] = np.maximum(x0[:, 0] + rng.uniform(-.5, .5, B), 3.2)
    x0[:, 1] += rng.normal(0, .05, B); x0[:, 2] += rng.normal(0, .02, B)
    x0[:, 5] += rng.uniform(-.5, .5, B); x0[:, 6] += rng.normal(0, .01, B)
    U = np.empty((B, N, 2))
    U[..., 0] = np.repeat(rng.uniform(-3000, 3000, (B, (N + 4) // 5)), 5, 1)[:, :N]
    U[..., 1] = np.repeat(rng.uniform(-.2, .2, (B, (N + 4) // 5)), 5, 1)[:, :N]
    ds = np.repeat(0.03 * x0[:, :1], N, 1)
    kap = oracle_tracks["ippodromo"].k(x0[:, 4:5] + np.cumsum(ds, 1) - ds)
    return x0, U, kap, ds


def test_rollout_st_states(eng, golden, oracle_tracks):
    x0, U, kap, ds = _inputs(golden, oracle_tracks, 777, 20, 1)
    X = eng.rollout_st(*(eng.tensor(a) for a in (x0, U, kap, ds))).cpu().numpy()
    Xo = om.st_rollout_spatial(om.DEFAULT, x0, U, kap, ds)
    assert np.isfinite(Xo).all()
    assert _rel(X, Xo) < RTOL


def test_rollout_st_jacobian(eng, golden, oracle_tracks):
    B, N = 130, 12   # ragged vs the 16-trajectory CTA tile
    x0, U, kap, ds = _inputs(golden, oracle_tracks, B, N, 2)
    X, AB = eng.rollout_st(*(eng.tensor(a) for a in (x0, U, kap, ds)), jac=True)
    X, AB = X.cpu().numpy(), AB.cpu().numpy()
    Xo = om.st_rollout_spatial(om.DEFAULT, x0, U, kap, ds)
    assert _rel(X, Xo) < RTOL
    Jo = om.st_step_jac(om.DEFAULT, Xo[:, :-1].reshape(-1, 8), U.reshape(-1, 2), kap.reshape(-1), ds.reshape(-1))
    Jo = Jo.reshape(B, N, 8, 10)
    scale = np.maximum(1.0, np.abs(Jo).max(axis=(2, 3), keepdims=True))
    assert float(np.max(np.abs(AB - Jo) / scale)) < RTOL


def test_rollout_pm(eng):
    rng = np.random.default_rng(3)
    B, M = 301, 35
    x0 = np.stack([rng.uniform(10, 25, B), rng.uniform(0, 200, B), rng.uniform(-2, 2, B), rng.normal(0, .05, B),
                   np.zeros(B)], 1)
    U = np.stack([rng.uniform(-1500, 3000, (B, M)), rng.uniform(-3000, 3000, (B, M))], -1)
    kap, ds = rng.uniform(0, 0.04, (B, M)), np.full((B, M), 3.0)
    X, AB = eng.rollout_pm(*(eng.tensor(a) for a in (x0, U, kap, ds)), jac=True)
    Xo = om.pm_rollout_spatial(om.DEFAULT, x0, U, kap, ds)
    ok = np.isfinite(Xo).all(axis=(1, 2)) & (Xo[:, :, 0].min(1) > 2.0)
    assert ok.sum() > B // 4
    assert _rel(X.cpu().numpy()[ok], Xo[ok]) < RTOL
    Jo = om.pm_step_jac(om.DEFAULT, Xo[:, :-1].reshape(-1, 5), U.reshape(-1, 2), kap.reshape(-1), ds.reshape(-1))
    Jo = Jo.reshape(B, M, 5, 7)[ok]
    scale = np.maximum(1.0, np.abs(Jo).max(axis=(2, 3), keepdims=True))
    assert float(np.max(np.abs(AB.cpu().numpy()[ok] - Jo) / scale)) < RTOL


def test_plant_step_golden_lap(eng, golden):
    """The device plant step replays every golden ippodromo run (racing_car.py:34-46)."""
    z, meta = golden
    worst = 0.0
    for key, m in meta.items():
        if m["track"] != "ippodromo":
            continue
        X, U = z[key + "/state"], z[key + "/action"]
        x = eng.tensor(X[:-1].copy())
        eng.plant_step(x, eng.tensor(U[1:].copy()))
        worst = max(worst, _rel(x.cpu().numpy(), X[1:]))
    assert worst < 2e-9, worst


def test_curvature_rel2glob_and_horizon(eng, oracle_tracks, golden):
    T = oracle_tracks["ippodromo"]
    rng = np.random.default_rng(5)
    s = np.concatenate([rng.uniform(-3, T.length + 60, 5000), [0.0, 0.05, 0.1, T.s_table[-1], T.s_table[-2]]])
    k = eng.curvature(eng.tensor(s)).cpu().numpy()
    assert np.abs(k - T.k(s)).max() < 1e-12
    ey, ep = rng.uniform(-3, 3, len(s)), rng.normal(0, .2, len(s))
    sg = np.abs(s)
    out = eng.rel2glob(eng.tensor(sg), eng.tensor(ey), eng.tensor(ep)).cpu().numpy()
    x, y, psi = T.rel2glob(sg, ey, ep)
    # inside the table: 1e-9; beyond the finish line both sides extrapolate the end cubic (x/y are evaluated
    # without fmod, track.py:102-107), which amplifies rounding by (distance in knots)^3 -> looser bar there
    inside = sg <= T.length
    for col, ref in ((0, x), (1, y)):
        assert np.abs(out[inside, col] - ref[inside]).max() < 1e-9
        assert (np.abs(out[~inside, col] - ref[~inside]) / np.maximum(1.0, np.abs(ref[~inside]))).max() < 1e-6
    d = np.abs(out[:, 2] - psi); d = np.minimum(d, 2 * np.pi - d)
    assert d.max() < 1e-9
    # horizon parameters (cascaded_mpc.py:316-338), both code paths (global table / smem-staged table)
    z, _ = golden
    for B in (37, 1024):
        H, N = eng.H, eng.cfg.N
        S = z["cascaded3_ippodromo/cascaded/state"]
        x0 = S[rng.integers(0, len(S) - 1, B)].copy()
        Xprev = rng.uniform(3, 20, (B, 8, H))
        ds, kap = eng.horizon_params(eng.tensor(x0), eng.tensor(Xprev))
        dso = np.concatenate([0.03 * Xprev[:, 0, :N], np.full((B, H - N), 3.0)], 1)
        st = np.cumsum(dso[:, :N], 1) - dso[:, :1] + x0[:, 4:5]
        sp = np.cumsum(dso[:, N:], 1) - dso[:, N - 1:N] + st[:, -1:]
        ko = T.k(np.concatenate([st, sp], 1))
        assert np.abs(ds.cpu().numpy() - dso).max() < 1e-15
        assert np.abs(kap.cpu().numpy() - ko).max() < 1e-12


def test_tyre_forces_log_kat(eng, golden):
    """The reference's own printed tyre forces (experiments/logs/*.log, written by racing.py:425-439 through
    dynamic_car.py:20-47): block k of a log = Fx_f, Fx_r, Fy_f, Fy_r at (state_traj[k+1], action_traj[k+1]) of the same
    run, 6 significant digits (tests/golden/log_kat.json, tests/golden/make_log_kat.py).  SURVEY App. B: block 200 of
    cascaded4_ippodromo = 4021.53, 1.23627e-05, 3480.71, 3584.5."""
    import json
    import os
    from tests.conftest import ROOT
    z, _ = golden
    kat = json.load(open(os.path.join(ROOT, "tests", "golden", "log_kat.json")))
    assert kat["cascaded4_ippodromo/cascaded"]["200"] == [4021.53, 1.23627e-05, 3480.71, 3584.5]
    n = 0
    for key, blocks in kat.items():
        ks = np.array(sorted(int(k) for k in blocks))
        ref = np.array([blocks[str(k)] for k in ks])
        S, A = z[key + "/state"], z[key + "/action"]
        f = eng.tyre_forces(eng.tensor(S[ks + 1]), eng.tensor(A[ks + 1])).cpu().numpy()[:, :4]
        assert (np.abs(f - ref) <= 6e-6 * np.abs(ref) + 1e-8).all(), key      # print precision of the log
        n += len(ks)
    assert n >= 80
    # symmetry of the lateral forces
    x = eng.tensor(np.array([[15.0, 0.2, 0.3, 0.05, 0, 0, 0, 0], [15.0, -0.2, -0.3, -0.05, 0, 0, 0, 0]]))
    u = eng.tensor(np.array([[4000.0, 0.0], [4000.0, 0.0]]))
    f = eng.tyre_forces(x, u).cpu().numpy()
    assert np.allclose(f[0, 2:4], -f[1, 2:4], rtol=1e-12)
