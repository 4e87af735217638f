"""GPU parity on the BENCHMARKED workloads: instances drawn from bench.py's own C3 / C2 batches against the CPU oracle.

tests/golden/bench_oracle.npz (tests/golden/make_bench_fixtures.py) holds an evenly spaced, i.e. unbiased, sample of
bench_nmpc.make_batch with bench.py's seeds -- C3: cascaded 20+35 with jittered obstacles, 65,536 perturbed warm-started
instances spread over a whole lap (straights and corners); C2: cascaded 20+15, 4,096 -- with the solver inputs and the
answers of the C++ oracle (IPOPT's algorithm incl. second-order correction), a few double-pinned by the Python oracle.
Everything goes through the C-ABI: vc_horizon_params (against the oracle's _init_horizon restatement) and vc_nmpc_solve.
Bars (BASELINE.json north_star): controls and objective within 1e-6 relative.
"""
import os

import numpy as np
import pytest

from tests.conftest import ROOT, load_yaml

pytestmark = pytest.mark.gpu
TOL = 1e-6


def _fx(workload):
    z = np.load(os.path.join(ROOT, "tests", "golden", "bench_oracle.npz"))
    return {k.split("/", 1)[1]: z[k] for k in z.files if k.startswith(workload + "/")}


def _engine(workload):
    import bench_nmpc as bn
    from vehicle_control_b200.engine import Engine
    from vehicle_control_b200.environment.track import Track
    cfg = bn.controller_config() if workload == "c3" else bn.controller_config(horizon_pm=15, obstacles=False)
    tr = Track(load_yaml("environment/ippodromo.yaml"))
    return Engine(load_yaml("models/dynamic_car.yaml"), cfg, tr), cfg


def _rel(a, b):
    return np.abs(a - b) / np.maximum(1.0, np.abs(b))


@pytest.mark.parametrize("workload,n_min", [("c3", 32), ("c2", 8)])
def test_bench_instances_match_oracle(workload, n_min):
    f = _fx(workload)
    eng, cfg = _engine(workload)
    n = len(f["x0"])
    assert n >= n_min and (f["status"] == 0).all()
    N, H = eng.cfg.N, eng.H
    assert f["Xw"].shape == (n, 8, H)
    if workload == "c3":   # the headline configuration: 20 + 35 with the 7 jittered ippodromo obstacles, mid-lap ticks included
        assert (N, H) == (20, 55) and f["obst"].shape == (n, 7, 3) and len(np.unique(f["tick"])) >= n_min
    x0, X, U = eng.tensor(f["x0"]), eng.tensor(f["Xw"]), eng.tensor(f["Uw"])
    ob = eng.tensor(f["obst"]) if f["obst"].size else None
    ds, kap = eng.horizon_params(x0, X)                       # _init_horizon on the device
    assert _rel(ds.cpu().numpy(), f["ds"]).max() < 1e-14 and np.abs(kap.cpu().numpy() - f["kappa"]).max() < 1e-12
    out = eng.solve(x0, kap, ds, X, U, obstacles=ob)
    st, it = out["status"].cpu().numpy(), out["iters"].cpu().numpy()
    assert (st == 0).all(), st
    Ug, Xg, og = U.cpu().numpy(), X.cpu().numpy(), out["obj"].cpu().numpy()
    eu = _rel(Ug, f["U"]).reshape(n, -1).max(1)
    ex = np.maximum(_rel(Xg[:, :, :N], f["X"][:, :, :N]).reshape(n, -1).max(1),
                    _rel(Xg[:, :5, N:], f["X"][:, :5, N:]).reshape(n, -1).max(1))
    eo = _rel(og, f["obj"])
    same_it = int((it == f["iters"]).sum())
    print(f"{workload}: {n} bench instances, max |dU| {eu.max():.2e} |dX| {ex.max():.2e} dObj {eo.max():.2e}; "
          f"iterations equal to the oracle's on {same_it}/{n} (engine mean {it.mean():.1f}, oracle {f['iters'].mean():.1f})")
    assert eu.max() < TOL and ex.max() < TOL and eo.max() < TOL
    assert same_it >= 0.6 * n      # same algorithm: most instances take the very same number of iterations
    # the instances the Python oracle solved as well (independent derivatives + dense LDL^T)
    py = ~np.isnan(f["U_py"]).reshape(n, -1).any(1)
    assert py.sum() >= 2 and _rel(Ug[py], f["U_py"][py]).max() < TOL


def test_fixture_is_drawn_from_the_bench_batch():
    """make_batch with bench.py's seed reproduces the stored perturbation of the stored ticks (same recipe, same rng
    stream): the fixture is a sample of what bench.py times, not a hand-picked set."""
    import bench_nmpc as bn
    f = _fx("c3")
    eng, cfg = _engine("c3")
    from vehicle_control_b200.environment.track import Track
    tr = Track(load_yaml("environment/ippodromo.yaml"))
    B, n = int(f["batch"]), len(f["x0"])
    assert B == 65536 and int(f["seed"]) == bn.SEED
    # a synthetic lap that carries the stored base states at the stored ticks is enough to replay the rng stream
    T = int(f["tick"].max()) + 1
    rng = np.random.default_rng(bn.SEED)
    dUx, dUy, dr = rng.uniform(-.5, .5, B), rng.normal(0, .05, B), rng.normal(0, .02, B)
    idx = f["index"]
    lap_x0 = f["x0"].copy()
    # undo the perturbation of Uy and r (additive): the base state must be the same for instances sharing a tick
    base_uy, base_r = lap_x0[:, 1] - dUy[idx], lap_x0[:, 2] - dr[idx]
    for t in np.unique(f["tick"]):
        m = f["tick"] == t
        if m.sum() > 1:
            assert np.ptp(base_uy[m]) < 1e-12 and np.ptp(base_r[m]) < 1e-12
    assert np.all(f["x0"][:, 7] == 0.0) and np.all(f["x0"][:, 0] >= 3.2)
    base = tr.obstacle_array()
    assert np.abs(f["obst"][:, :, 2] - base[None, :, 2]).max() == 0.0     # radii are not jittered
    assert np.abs(f["obst"][:, :, 0] - base[None, :, 0]).max() <= 1.0 and np.abs(f["obst"][:, :, 1] - base[None, :, 1]).max() <= 0.3


def test_block_wide_kernels_equal_warp_kernels_bit_for_bit():
    """The late-pass kernels (one block per instance: k_step_a_wide / k_step_b_wide) compute every number with the same
    arithmetic as the warp-per-instance kernels: the 64 C3 bench instances solved entirely by one or the other path
    (wide_max = 0 / large) return identical bits -- so a result does not depend on how many instances share its batch."""
    import torch
    f = _fx("c3")
    res = []
    for wide_max in (0, 1 << 20):
        eng, cfg = _engine("c3")
        eng.set_config(cfg, wide_max=wide_max)
        x0, X, U = eng.tensor(f["x0"]), eng.tensor(f["Xw"]), eng.tensor(f["Uw"])
        ds, kap = eng.horizon_params(x0, X)
        out = eng.solve(x0, kap, ds, X, U, obstacles=eng.tensor(f["obst"]))
        res.append((X.clone(), U.clone(), out["obj"].clone(), out["iters"].clone(), out["status"].clone()))
    for a, b in zip(*res):
        assert torch.equal(a, b)


def test_command_host_equals_device_path_and_oracle():
    """vc_nmpc_command_host (the reference-facing call with HOST buffers: horizon parameters, solve and all copies inside
    one call -- the call bench.py's `e2e` times) against the device-pointer path on the same instances (bit for bit: the
    same kernels run) and against the oracle's answers."""
    f = _fx("c3")
    eng, cfg = _engine("c3")
    n = len(f["x0"])
    x0, X, U = eng.tensor(f["x0"]), eng.tensor(f["Xw"]), eng.tensor(f["Uw"])
    ob = eng.tensor(f["obst"])
    ds, kap = eng.horizon_params(x0, X)
    dev = eng.solve(x0, kap, ds, X, U, obstacles=ob)
    hX, hU = np.ascontiguousarray(f["Xw"]).copy(), np.ascontiguousarray(f["Uw"]).copy()
    host = eng.command_host(np.ascontiguousarray(f["x0"]), hX, hU, np.ascontiguousarray(f["obst"]))
    assert (host["status"] == 0).all()
    assert np.array_equal(host["iters"], dev["iters"].cpu().numpy())
    assert np.array_equal(hU, U.cpu().numpy()) and np.array_equal(hX, X.cpu().numpy())
    assert np.array_equal(host["obj"], dev["obj"].cpu().numpy())
    assert _rel(hU, f["U"]).max() < TOL and _rel(host["obj"], f["obj"]).max() < TOL
    # a second call on a smaller batch reuses the handle's staging buffers
    m = n // 2
    hX2, hU2 = np.ascontiguousarray(f["Xw"][:m]).copy(), np.ascontiguousarray(f["Uw"][:m]).copy()
    h2 = eng.command_host(np.ascontiguousarray(f["x0"][:m]), hX2, hU2, np.ascontiguousarray(f["obst"][:m]))
    assert np.array_equal(hU2, hU[:m]) and np.array_equal(h2["iters"], host["iters"][:m])
