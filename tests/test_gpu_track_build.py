"""GPU: track CONSTRUCTION on the device (include/vc_api.h: vc_build_tracks).

The host `Track` mirror (environment/track.py) is pinned by the reference's golden runs and the oracle; here the device
pipeline -- corner polygon -> waypoints -> smoothing -> centre-line splines -> length -> curvature table -> Track.k
(track.py:254-296, 206-252, 109-119, 156-167) and the obstacle placement (:131-138) -- is compared with it for the four
shipped tracks and a generated polygon, built in ONE batched call.  Tolerances: waypoints are bit-identical (numpy's
pairwise summation is restated on the device); spline coefficients and the length agree to rounding; the curvature is a
second difference of coefficients 0.1 m apart, which amplifies rounding-level differences of the two banded solvers by
~1e5, hence 1e-8 relative to the largest curvature there.
"""
import numpy as np
import pytest
import torch

from tests.conftest import load_yaml

pytestmark = pytest.mark.gpu
NAMES = ["ippodromo", "shoe", "bean", "gokart"]


def _configs():
    cfgs = [dict(load_yaml(f"environment/{n}.yaml")) for n in NAMES]
    gen = {"name": "generated", "width": 8, "resolution": 0.125, "smoothing": 120, "obstacle_data": [[20, 1, 1], [90, -2, 2]],
           "corners": [[0, 0], [45, 0], [70, 20], [70, 55], [20, 75], [-30, 40], [-30, 0], [0, 0]]}
    return cfgs + [gen]


@pytest.fixture(scope="module")
def built():
    from vehicle_control_b200.engine import Engine
    from vehicle_control_b200.environment.track import Track
    cfgs = _configs()
    host = [Track(c) for c in cfgs]
    ccfg = dict(load_yaml("controllers/cascaded.yaml")); ccfg["obstacles"] = False
    eng = Engine(load_yaml("models/dynamic_car.yaml"), ccfg, host[0])
    info = eng.build_tracks(cfgs)
    return eng, cfgs, host, info, ccfg


def test_built_tracks_match_host_construction(built):
    eng, cfgs, host, info, _ = built
    for t, tr in enumerate(host):
        tab = tr.device_tables()
        assert info[t]["n_waypoints"] == tr.n_waypoints and info[t]["n_k"] == len(tr.s_table)
        wx, wy = eng.built_track_table(t, "waypoints_x"), eng.built_track_table(t, "waypoints_y")
        assert np.array_equal(wx, tr.waypoints_xy[:, 0]) and np.array_equal(wy, tr.waypoints_xy[:, 1])   # bit for bit
        for name in ("x_coefs", "y_coefs"):
            c = eng.built_track_table(t, name)
            assert np.max(np.abs(c - tab[name])) < 1e-11 * max(1.0, np.max(np.abs(tab[name])))
        assert abs(info[t]["length"] - tr.length) < 1e-12 * tr.length
        kt = eng.built_track_table(t, "curvature_table")
        kmax = float(np.max(np.abs(tr.curvature_table)))
        assert np.max(np.abs(kt - tr.curvature_table)) < 1e-8 * kmax
        kc = eng.built_track_table(t, "k_coefs")
        assert np.max(np.abs(kc - tab["k_coefs"])) < 1e-8 * kmax
        print(f"{cfgs[t]['name']}: n_wp {tr.n_waypoints}, length {tr.length:.6f} (dev {info[t]['length'] - tr.length:+.1e}), "
              f"kappa table dev {np.max(np.abs(kt - tr.curvature_table)) / kmax:.1e}, k coefs dev {np.max(np.abs(kc - tab['k_coefs'])) / kmax:.1e}")


def test_obstacle_placement_on_built_tracks(built):
    eng, cfgs, host, info, _ = built
    for t, tr in enumerate(host):
        if not tr.obstacles:
            continue
        s = eng.tensor([o.s for o in tr.obstacles]); ey = eng.tensor([o.ey for o in tr.obstacles])
        xy = eng.built_track_rel2glob(t, s, ey).cpu().numpy()
        ref = np.array([[o.cx, o.cy] for o in tr.obstacles])
        assert np.max(np.abs(xy - ref)) < 1e-9 * max(1.0, np.max(np.abs(ref)))


def test_solve_on_built_tracks_equals_host_built_atlas(built):
    """horizon parameters, NMPC solve and plant step on the device-built atlas against the host-built atlas of the same
    tracks (tests/test_gpu_tracks.py pins that one): curvature to 1e-8, controls to 1e-6."""
    from vehicle_control_b200.engine import Engine
    from vehicle_control_b200.controllers.mpc.cascaded_mpc import cold_guess
    eng, cfgs, host, info, ccfg = built
    ref = Engine(load_yaml("models/dynamic_car.yaml"), ccfg, host[0])
    ref.set_track_atlas(host)
    B = 40
    rng = np.random.default_rng(11)
    tid = (np.arange(B) % len(host)).astype(np.int32)
    x0 = np.zeros((B, 8)); x0[:, 0] = 4.0 + rng.uniform(0, 2, B)
    x0[:, 4] = np.array([rng.uniform(1.0, 0.8 * host[t].length) for t in tid]); x0[:, 5] = rng.uniform(-1, 1, B)
    N, H = int(ccfg["horizon"]), int(ccfg["horizon"]) + int(ccfg["horizon_pm"])
    np.random.seed(31)
    X0, U0 = cold_guess(N, H)
    res = []
    for e in (eng, ref):
        xd, td = e.tensor(x0), e.tensor(tid, dtype=torch.int32)
        X, U = e.tensor(np.repeat(X0[None], B, 0)), e.tensor(np.repeat(U0[None], B, 0))
        ds, kap = e.horizon_params(xd, X, track_id=td)
        out = e.solve(xd, kap, ds, X, U)
        xn = xd.clone()
        e.plant_step(xn, U[:, :, 0].contiguous(), track_id=td)
        res.append((kap.cpu().numpy(), U.cpu().numpy(), out["status"].cpu().numpy(), xn.cpu().numpy()))
    (k1, U1, s1, x1), (k2, U2, s2, x2) = res
    assert np.max(np.abs(k1 - k2)) < 1e-8 * max(1e-3, np.max(np.abs(k2)))
    ok = (s1 <= 1) & (s2 <= 1)
    assert ok.mean() > 0.9
    scale = np.maximum(1.0, np.abs(U2[ok]))
    assert np.max(np.abs(U1[ok] - U2[ok]) / scale) < 1e-6
    assert np.max(np.abs(x1[ok] - x2[ok]) / np.maximum(1.0, np.abs(x2[ok]))) < 1e-9


def test_build_rejects_bad_input(built):
    eng = built[0]
    bad = dict(_configs()[0]); bad["smoothing"] = 100000
    with pytest.raises(RuntimeError):
        eng.build_tracks([bad])
    eng.build_tracks(_configs())      # the handle still works afterwards
