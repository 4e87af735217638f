"""GPU parity: the batched NMPC solve (through the C-ABI) against the CPU oracle and the reference's saved answers.

Fixtures: tests/golden/nmpc_oracle.npz (tests/golden/make_nmpc_fixtures.py): closed-loop ticks of reference runs,
solver inputs + the oracle's solution + the reference's own first action (action_traj[k+1]).
Bars (BASELINE.json north_star): optimal controls within 1e-6 (relative to max(1,|u|)), objective within 1e-6 relative.
"""
import json
import os

import numpy as np
import pytest

from tests.conftest import ROOT, load_yaml

pytestmark = pytest.mark.gpu
TOL_U = 1e-6
TOL_OBJ = 1e-6


def _fixtures():
    z = np.load(os.path.join(ROOT, "tests", "golden", "nmpc_oracle.npz"))
    tags = sorted({k.rsplit("/", 1)[0] for k in z.files})
    return z, tags


def _engine(run_key, golden):
    from vehicle_control_b200.engine import Engine
    from vehicle_control_b200.environment.track import Track
    _, meta = golden
    m = meta[run_key]
    tr = Track(load_yaml(f"environment/{m['track']}.yaml"))
    return Engine(load_yaml("models/dynamic_car.yaml"), m["config"], tr)


def _rel(a, b):
    return np.abs(a - b) / np.maximum(1.0, np.abs(b))


def _solve(eng, f, B=1):
    H = eng.H
    rep = lambda a: np.ascontiguousarray(np.repeat(a[None], B, 0))
    x0, kap, ds = (eng.tensor(rep(f[k])) for k in ("x0", "kappa", "ds"))
    X, U = eng.tensor(rep(f["X0"])), eng.tensor(rep(f["U0"]))
    obst = eng.tensor(rep(f["obst"])) if f["obst"].size else None
    out = eng.solve(x0, kap, ds, X, U, obstacles=obst)
    return X.cpu().numpy(), U.cpu().numpy(), {k: v.cpu().numpy() for k, v in out.items()}


def test_solve_matches_oracle_and_reference(golden):
    z, tags = _fixtures()
    assert tags
    for tag in tags:
        run_key = tag.rsplit("/", 1)[0]
        f = {k: z[f"{tag}/{k}"] for k in ("x0", "kappa", "ds", "obst", "X0", "U0", "X", "U", "obj", "ref_u", "iters")}
        eng = _engine(run_key, golden)
        X, U, out = _solve(eng, f, B=3)
        assert (out["status"] <= 1).all(), (tag, out)
        # all copies of the instance agree bit for bit (no cross-instance interference)
        assert np.array_equal(U[0], U[1]) and np.array_equal(U[0], U[2])
        N = eng.cfg.N
        eu = _rel(U[0], f["U"]).max()
        ex = _rel(X[0, :, :N], f["X"][:, :N]).max()
        if eng.H > N:   # PointMass columns only use rows 0-4 (cascaded_mpc.py:70,182)
            ex = max(ex, _rel(X[0, :5, N:], f["X"][:5, N:]).max())
        eo = abs(out["obj"][0] - f["obj"]) / max(1.0, abs(f["obj"]))
        er = _rel(U[0, :, 0], f["ref_u"]).max()
        print(f"{tag}: iters {out['iters'][0]} (oracle {int(f['iters'])}) |dU| {eu:.2e} |dX| {ex:.2e} dObj {eo:.2e} vs reference u0 {er:.2e}")
        assert eu < TOL_U and ex < TOL_U and eo < TOL_OBJ, tag
        assert er < TOL_U, tag      # the reference's own saved first action (IPOPT)


def test_eval_matches_oracle_objective(golden):
    z, tags = _fixtures()
    tag = tags[0]
    run_key = tag.rsplit("/", 1)[0]
    eng = _engine(run_key, golden)
    f = {k: z[f"{tag}/{k}"] for k in ("x0", "kappa", "ds", "obst", "X", "U", "obj")}
    t = lambda a: eng.tensor(np.ascontiguousarray(a[None]))
    out = eng.evaluate(t(f["x0"]), t(f["kappa"]), t(f["ds"]), t(f["X"]), t(f["U"])).cpu().numpy()[0]
    assert abs(out[0] - f["obj"]) / max(1.0, abs(f["obj"])) < 1e-12
    assert out[1] < 1e-7 and out[2] > -1e-7


def _simconfig(name, track="ippodromo", obstacles=False):
    return {"track_name": track, "controller_names": [name], "obstacles": obstacles, "slip_angle_plot": False,
            "name": "test", "save_gif": False, "save_images": False, "save_data": False, "load": False, "logging": False}


@pytest.mark.parametrize("run_key,ticks", [("cascaded3_ippodromo/cascaded", 60), ("singletrack3_ippodromo/singletrack", 30)])
def test_closed_loop_follows_reference_run(golden, run_key, ticks):
    """RacingSimulator (command -> drive per tick, racing.py:217-242) against the reference's saved trajectories:
    the controller's first actions and the plant states of the first ticks."""
    from vehicle_control_b200.simulation.racing import RacingSimulator
    z, meta = golden
    m = meta[run_key]
    np.random.seed(31)   # cascaded_mpc.py:13 (the cold guess is drawn at construction)
    sim = RacingSimulator(_simconfig(run_key.split("/")[1]), load_yaml("models/dynamic_car.yaml"),
                          load_yaml(f"environment/{m['track']}.yaml"), controller_configs=[m["config"]])
    for n in range(ticks):
        assert sim.update(n)
    name = sim.names[0]
    st, ac, pr, el = sim._arrays(name)
    S, A, P = z[run_key + "/state"], z[run_key + "/action"], z[run_key + "/preds"]
    assert st.shape == (ticks + 1, 8) and ac.shape == (ticks + 1, 2) and pr.shape == (ticks,) + P.shape[1:]
    eu = (np.abs(ac[1:] - A[1:ticks + 1]) / np.maximum(1.0, np.abs(A[1:ticks + 1]))).max()
    ex = (np.abs(st - S[:ticks + 1]) / np.maximum(1.0, np.abs(S[:ticks + 1]))).max()
    ep = np.abs(pr[:, :, :2] - P[:ticks, :, :2]).max()
    print(f"{run_key}: {ticks} ticks, actions {eu:.2e}, states {ex:.2e}, predicted footprints {ep:.2e}, iters {sim.iters[name]}")
    assert eu < 1e-6 and ex < 1e-6 and ep < 1e-5


def test_full_lap_cascaded_matches_reference_lap(golden, tmp_path):
    """Whole lap 20+35 on ippodromo: same number of ticks and lap time as the reference run; the closed loop follows
    the reference to 1e-6 up to the first tick where IPOPT's iterate path and ours end in different local solutions
    of the non-convex NLP (tick 194); run files round-trip."""
    from vehicle_control_b200.simulation.racing import RacingSimulator
    z, meta = golden
    key = "cascaded3_ippodromo/cascaded"
    np.random.seed(31)
    cfg = _simconfig("cascaded"); cfg["save_data"] = True; cfg["data_dir"] = str(tmp_path)
    sim = RacingSimulator(cfg, load_yaml("models/dynamic_car.yaml"), load_yaml("environment/ippodromo.yaml"),
                          controller_configs=[meta[key]["config"]])
    n = sim.run()
    S, A = z[key + "/state"], z[key + "/action"]
    st, ac, pr, el = sim._arrays("cascaded")
    print(f"ticks {n} (reference {len(S) - 1}), laptime {st[-1, 7]:.2f} (reference {S[-1, 7]:.2f}), "
          f"mean iters {np.mean(sim.iters['cascaded']):.1f}, ms/tick {1e3 * np.median(el):.2f} (reference {1e3 * np.median(z[key + '/elapsed']):.1f})")
    # after the branch tick the two closed loops are different trajectories: the lap ends within one tick of the reference's
    assert abs(n - (len(S) - 1)) <= 1 and abs(st[-1, 7] - S[-1, 7]) < 0.05 + 1e-9
    T = min(len(ac), len(A))
    ac, A, stc, Sc = ac[:T], A[:T], st[:T], S[:T]
    eu = (np.abs(ac[1:] - A[1:]) / np.maximum(1.0, np.abs(A[1:]))).max(1)
    ex = (np.abs(stc - Sc) / np.maximum(1.0, np.abs(Sc))).max(1)
    first_bad = int(np.flatnonzero(eu > 1e-6)[0]) if (eu > 1e-6).any() else n
    print(f"first tick with action error > 1e-6: {first_bad}; max state error before it {ex[:first_bad + 1].max():.2e}")
    assert first_bad >= 190 and ex[:first_bad + 1].max() < 1e-6
    assert np.abs(stc[:, 5] - Sc[:, 5]).max() < 1.0      # never further than 1 m of lateral offset from the reference line
    # the run format of racing.py:441-459
    cfg2 = dict(cfg); cfg2["load"] = True; cfg2["save_data"] = False
    np.random.seed(31)
    sim2 = RacingSimulator(cfg2, load_yaml("models/dynamic_car.yaml"), load_yaml("environment/ippodromo.yaml"),
                           controller_configs=[meta[key]["config"]])
    assert sim2.loaded and np.array_equal(sim2.state_traj["cascaded"], st) and sim2.preds["cascaded"].shape == pr.shape
    assert abs(sim2.summary()["cascaded"]["laptime"] - st[-1, 7]) < 1e-9


@pytest.mark.parametrize("run_key,dt_lap", [("singletrack3_ippodromo/singletrack", 0.0), ("race2_shoe/cascaded", 0.05),
                                            ("cascaded_obstacles1_ippodromo/cascaded", 0.5),
                                            ("cascaded_giantObstacle3_ippodromo/cascaded", 2.0),
                                            ("singletrack_shoe/singletrack", 0.05)])
def test_laps_complete_like_reference(golden, run_key, dt_lap):
    """Every solve of a whole lap converges (the reference raises on IPOPT failure) and the lap time equals the
    reference's (within dt_lap seconds where the two solvers take different sides of a branch of the NLP)."""
    from vehicle_control_b200.simulation.racing import RacingSimulator
    z, meta = golden
    m = meta[run_key]
    np.random.seed(31)
    sim = RacingSimulator(_simconfig(run_key.split("/")[1], track=m["track"], obstacles=bool(m["config"].get("obstacles"))),
                          load_yaml("models/dynamic_car.yaml"), load_yaml(f"environment/{m['track']}.yaml"),
                          controller_configs=[m["config"]])
    n = sim.run()
    S = z[run_key + "/state"]
    st, ac, pr, el = sim._arrays(run_key.split("/")[1])
    print(f"{run_key}: ticks {n} (reference {len(S) - 1}), laptime {st[-1, 7]:.2f} (reference {S[-1, 7]:.2f}), max iters {max(sim.iters[run_key.split('/')[1]])}")
    assert st[-1, 4] > sim.track.length - 0.1            # finished, no solver failure on the way
    assert abs(st[-1, 7] - S[-1, 7]) <= dt_lap + 1e-9


def test_batched_simulator_lockstep(golden):
    """16 vehicles with perturbed starts advance in lock step; vehicle 0 (unperturbed) equals the single-vehicle loop."""
    from vehicle_control_b200.simulation.racing import BatchedRacingSimulator
    _, meta = golden
    cfg = meta["cascaded1_ippodromo/cascaded"]["config"]
    rng = np.random.default_rng(1004)
    B = 16
    x0 = np.zeros((B, 8)); x0[:, 0] = 4.0; x0[:, 4] = 1.0 + rng.uniform(0, 5, B); x0[:, 5] = rng.uniform(-1, 1, B)
    x0[0] = [4, 0, 0, 0, 1, 0, 0, 0]
    np.random.seed(31)
    sim = BatchedRacingSimulator(load_yaml("models/dynamic_car.yaml"), load_yaml("environment/ippodromo.yaml"), cfg, x0)
    for _ in range(12):
        assert sim.update() == B
    assert not bool(sim.failed.any())
    z, _ = golden
    S = z["cascaded1_ippodromo/cascaded/state"]
    x = sim.x.cpu().numpy()
    assert (np.abs(x[0] - S[12]) / np.maximum(1, np.abs(S[12]))).max() < 1e-6
    assert np.isfinite(x).all() and (x[:, 7] > 0.59).all()


@pytest.mark.parametrize("case", ["cascaded_obstacles_shoe_cascaded_155", "race1_shoe_singletrack_393"])
def test_failed_line_search_is_retried(golden, case):
    """Closed-loop ticks (found by scripts/lap_parity_all.py, saved in tests/golden/hard_ticks.npz) on which the filter
    line search stalls without IPOPT's restoration phase -- the CPU oracle stalls on them in the same way.  The
    controller re-solves such instances from the shifted previous solution, then from the cold guess
    (`retry_failed`, cascaded_mpc.py mirror); with the retries switched off the failure is reported like the
    reference reports an IPOPT failure (exception)."""
    from vehicle_control_b200.simulation.racing import RacingSimulator
    _, meta = golden
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "hard_ticks.npz"))
    key = str(z[case + "/key"])
    m = meta[key]
    name = key.split("/")[1]
    for retry in (False, True):
        cfg = dict(m["config"]); cfg["retry_failed"] = retry
        np.random.seed(31)
        sc = {"track_name": m["track"], "controller_names": [name], "obstacles": bool(m["config"].get("obstacles")),
              "slip_angle_plot": False, "name": "t", "save_gif": False, "save_images": False, "save_data": False,
              "load": False, "logging": False}
        sim = RacingSimulator(sc, load_yaml("models/dynamic_car.yaml"), load_yaml(f"environment/{m['track']}.yaml"),
                              controller_configs=[cfg])
        ctl, car = sim.controllers[0], sim.cars[0]
        ctl.state_prediction, ctl.action_prediction = z[case + "/Xw"], z[case + "/Uw"]
        car.state = type(car.state)(*z[case + "/x0"])
        if not retry:
            # without the retries the tick either solves or is reported like the reference reports an IPOPT failure
            try:
                ctl.command(car.state)
                assert int(ctl.last_info["status"]) <= 1
            except RuntimeError as e:
                assert "NMPC solve failed" in str(e)
        else:
            ctl.command(car.state)
            assert int(ctl.last_info["status"]) <= 1
            r = ctl._batched.retries
            assert r["unrecovered"] == 0 and r["shifted"] + r["cold"] <= 1


def test_iteration_limit_is_reported_for_every_instance(golden):
    """max_iter = 3: every instance comes back with status VC_MAX_ITER, exactly 3 iterations and a defined objective /
    trajectory (nothing is left uninitialised when the pass loop ends with unfinished instances)."""
    z, tags = _fixtures()
    tag = "cascaded3_ippodromo/cascaded/tick1"
    f = {k: z[f"{tag}/{k}"] for k in ("x0", "kappa", "ds", "obst", "X0", "U0")}
    eng = _engine(tag.rsplit("/", 1)[0], golden)
    _, meta = golden
    eng.set_config(meta[tag.rsplit("/", 1)[0]]["config"], max_iter=3)
    X, U, out = _solve(eng, f, B=5)
    assert (out["status"] == 2).all() and (out["iters"] == 3).all()
    assert np.isfinite(out["obj"]).all() and np.isfinite(X[:, :, :eng.cfg.N]).all() and np.isfinite(U).all()
    assert not np.array_equal(U[0], f["U0"])      # three steps were taken


@pytest.mark.parametrize("run_key,min_1e6,max_branch", [("singletrack2_ippodromo/singletrack", 1.0, 0),
                                                        ("cascaded4_ippodromo/cascaded", 0.93, 12)])
def test_follow_mode_per_tick_parity_with_ipopt(golden, run_key, min_1e6, max_branch):
    """Per-tick parity with IPOPT over a WHOLE saved lap, on the CUDA path (the GPU counterpart of
    profiles/r02_branch_ticks.md).  Follow mode: every tick starts from the reference's own saved state, the warm start is
    our previous solution (cascaded_mpc.py:316-338), and the first action is compared with IPOPT's saved one
    (action_traj[k+1]).  The singletrack lap must agree on EVERY tick; on the cascaded 20+35 lap at least 93 % of the ticks
    agree to 1e-6 and at most 12 ticks (two short branch clusters of the non-convex NLP, listed in the profile) differ
    by more than 1e-3."""
    from vehicle_control_b200.controllers.mpc.cascaded_mpc import cold_guess
    z, meta = golden
    m = meta[run_key]
    eng = _engine(run_key, golden)
    N, H = eng.cfg.N, eng.H
    S, A = z[run_key + "/state"], z[run_key + "/action"]
    X0, U0 = cold_guess(N, H, rng_draw=np.random.RandomState(31).random_sample((2, H)))
    X, U = eng.tensor(X0[None]), eng.tensor(U0[None])
    ob = None
    if m["config"].get("obstacles"):
        oa = eng.track.obstacle_array()
        ob = eng.tensor(oa[None]) if len(oa) else None
    err, iters = [], []
    for k in range(len(A) - 1):
        x = eng.tensor(S[k][None])
        Xk, Uk = X.clone(), U.clone()
        ds, kap = eng.horizon_params(x, Xk)
        out = eng.solve(x, kap, ds, Xk, Uk, obstacles=ob)
        if int(out["status"][0]) > 1:
            err.append(np.inf)           # keep the warm start, as the survey does
            continue
        u = Uk[0, :, 0].cpu().numpy()
        err.append(float((np.abs(u - A[k + 1]) / np.maximum(1.0, np.abs(A[k + 1]))).max()))
        iters.append(int(out["iters"][0]))
        X, U = Xk, Uk
    err = np.array(err)
    frac = float((err <= 1e-6).mean())
    big = int((err > 1e-3).sum())
    print(f"{run_key}: {len(err)} ticks, {100 * frac:.1f} % within 1e-6 of IPOPT, {int((err <= 1e-5).sum())} within 1e-5, {big} beyond 1e-3 "
          f"(ticks {np.flatnonzero(err > 1e-3).tolist()}), mean iterations {np.mean(iters):.1f}")
    assert frac >= min_1e6 and big <= max_branch
