"""GPU: the BASELINE.json configurations at (or near) their full sizes, through size-independent properties.

C2  cascaded 20+15, 4 096 perturbed warm-started instances: every returned point is a feasible KKT point of its own
    NLP (checked by the independent evaluation entry point), results do not depend on the position of an instance in
    the batch (bitwise), re-solving from a solution returns the same solution.
C4  shoe track, lock-step closed loop of many vehicles: vehicle 0 equals the single-vehicle simulator.
C5  SingleTrack rollout + Jacobians at 65 536 x 50: a rollout equals two chained half rollouts (bitwise); the [A|B]
    blocks predict a finite-difference perturbation of the rollout; edge cases (B = 1, N = 1, ragged B).
"""
import numpy as np
import pytest
import torch

from tests.conftest import load_yaml

pytestmark = pytest.mark.gpu


def _c2_batch(golden, B, seed=1002):
    """base states = rows of the reference's cascaded1 lap, warm start = the engine's own
    closed-loop prediction at that tick, perturbation recipe of the survey."""
    from bench_nmpc import make_batch, replay_lap
    from vehicle_control_b200.engine import Engine
    from vehicle_control_b200.environment.track import Track
    _, meta = golden
    cfg = dict(meta["cascaded1_ippodromo/cascaded"]["config"])
    tr = Track(load_yaml("environment/ippodromo.yaml"))
    eng = Engine(load_yaml("models/dynamic_car.yaml"), cfg, tr)
    np.random.seed(31)
    lap = replay_lap(eng, cfg, tr, max_ticks=120)
    x0, Xw, Uw, ob, idx = make_batch(lap, tr, B, seed, cfg)
    return eng, x0, Xw, Uw


def test_c2_batch_kkt_permutation_idempotence(golden):
    B = 4096
    eng, x0h, Xwh, Uwh = _c2_batch(golden, B)
    x0, X, U = eng.tensor(x0h), eng.tensor(Xwh), eng.tensor(Uwh)
    ds, kap = eng.horizon_params(x0, X)
    out = eng.solve(x0, kap, ds, X, U)
    st = out["status"].cpu().numpy()
    ok = st <= 1
    print(f"C2: converged {ok.mean():.4f}, mean iters {out['iters'].float().mean().item():.1f}, max {int(out['iters'].max())}")
    assert ok.mean() > 0.99
    # (1) feasibility + objective through the independent evaluation kernel
    ev = eng.evaluate(x0, kap, ds, X, U).cpu().numpy()
    assert np.abs(ev[ok, 0] - out["obj"].cpu().numpy()[ok]).max() < 1e-9 * max(1.0, np.abs(ev[ok, 0]).max())
    assert ev[ok, 1].max() < 1e-7 and ev[ok, 2].min() > -1e-7
    assert out["kkt"].cpu().numpy()[ok].max() <= 1e-6
    # (2) permutation invariance, bit for bit
    perm = np.random.default_rng(5).permutation(B)
    x0p, Xp, Up = eng.tensor(x0h[perm]), eng.tensor(Xwh[perm]), eng.tensor(Uwh[perm])
    dsp, kapp = eng.horizon_params(x0p, Xp)
    eng.solve(x0p, kapp, dsp, Xp, Up)
    assert torch.equal(Up, U[torch.as_tensor(perm, device=U.device)])
    assert torch.equal(Xp, X[torch.as_tensor(perm, device=U.device)])
    # (3) re-solving from the solution (same parameters) stays at the solution
    X2, U2 = X.clone(), U.clone()
    out2 = eng.solve(x0, kap, ds, X2, U2)
    ok2 = ok & (out2["status"].cpu().numpy() <= 1)
    assert ok2.sum() >= 0.99 * ok.sum()
    du = (U2 - U).abs().cpu().numpy()[ok2] / np.maximum(1.0, U.abs().cpu().numpy()[ok2])
    print(f"C2: re-solve max control change {du.max():.2e}")
    assert du.max() < 1e-6


def test_c2_empty_and_single_instance(golden):
    eng, x0h, Xwh, Uwh = _c2_batch(golden, 3)
    x0, X, U = eng.tensor(x0h[:1]), eng.tensor(Xwh[:1]), eng.tensor(Uwh[:1])
    ds, kap = eng.horizon_params(x0, X)
    out = eng.solve(x0, kap, ds, X, U)
    assert int(out["status"][0]) <= 1
    xa, Xa, Ua = eng.tensor(x0h), eng.tensor(Xwh), eng.tensor(Uwh)
    dsa, kapa = eng.horizon_params(xa, Xa)
    eng.solve(xa, kapa, dsa, Xa, Ua)
    assert torch.equal(Ua[0], U[0])          # B = 1 and B = 3 give the same bits for instance 0


def test_c4_shoe_lockstep_matches_single_vehicle(golden):
    from vehicle_control_b200.simulation.racing import BatchedRacingSimulator, RacingSimulator
    _, meta = golden
    cfg = dict(meta["race2_shoe/cascaded"]["config"])
    cfg["horizon_pm"] = 35                                     # BASELINE config C4: 20+35 on shoe
    B, T = 512, 25
    rng = np.random.default_rng(1004)
    x0 = np.zeros((B, 8)); x0[:, 0] = 4.0; x0[:, 4] = 1.0 + rng.uniform(0, 5, B); x0[:, 5] = rng.uniform(-1, 1, B)
    x0[0] = [4, 0, 0, 0, 1, 0, 0, 0]
    np.random.seed(31)
    sim = BatchedRacingSimulator(load_yaml("models/dynamic_car.yaml"), load_yaml("environment/shoe.yaml"), cfg, x0)
    for _ in range(T):
        assert sim.update() == B
    assert not bool(sim.failed.any())
    np.random.seed(31)
    sc = {"track_name": "shoe", "controller_names": ["cascaded"], "obstacles": False, "slip_angle_plot": False,
          "name": "t", "save_gif": False, "save_images": False, "save_data": False, "load": False, "logging": False}
    one = RacingSimulator(sc, load_yaml("models/dynamic_car.yaml"), load_yaml("environment/shoe.yaml"),
                          controller_configs=[cfg])
    for n in range(T):
        assert one.update(n)
    xs = np.asarray(one.cars[0].state.values, dtype=float)
    xb = sim.x.cpu().numpy()
    assert np.array_equal(xb[0], xs)                           # same kernels, same bits
    assert np.isfinite(xb).all() and (xb[:, 4] > x0[:, 4]).all()


def test_c5_rollout_composition_and_linearisation(golden):
    from bench_common import golden_states, rollout_inputs
    from vehicle_control_b200.engine import Engine
    from vehicle_control_b200.environment.track import Track
    tr = Track(load_yaml("environment/ippodromo.yaml"))
    eng = Engine(load_yaml("models/dynamic_car.yaml"), load_yaml("controllers/cascaded.yaml"), tr)
    B, N = 65536, 50
    x0, U, kap, ds = (eng.tensor(a) for a in rollout_inputs(B, N, 1005, tr, golden_states()))
    X, AB = eng.rollout_st(x0, U, kap, ds, jac=True)
    assert bool(torch.isfinite(X).all()) and bool(torch.isfinite(AB).all())
    Xs = eng.rollout_st(x0, U, kap, ds)
    # the Jacobian kernel carries the same states (dual-number and plain arithmetic contract FMAs differently)
    # (rounding-level differences grow along trajectories that random controls drive into a spin: bound the first
    # steps tightly and the whole horizon in the median)
    rel = (X - Xs).abs() / Xs.abs().clamp(min=1.0)
    assert rel[:, :6].max().item() < 1e-12 and rel.amax((1, 2)).median().item() < 1e-12
    h = N // 2                                                 # N steps = h steps followed by N-h steps, bit for bit
    Xa = eng.rollout_st(x0, U[:, :h].contiguous(), kap[:, :h].contiguous(), ds[:, :h].contiguous())
    Xb = eng.rollout_st(Xa[:, -1].contiguous(), U[:, h:].contiguous(), kap[:, h:].contiguous(), ds[:, h:].contiguous())
    assert torch.equal(Xa, Xs[:, :h + 1]) and torch.equal(Xb, Xs[:, h:])
    # first-order prediction of a perturbed first step by [A|B] (directional finite difference)
    rng = np.random.default_rng(3)
    dx = torch.as_tensor(rng.normal(0, 1, (B, 8)), device=X.device) * 1e-6
    dx[:, 4] = 0.0; dx[:, 7] = 0.0
    du = torch.as_tensor(rng.normal(0, 1, (B, 2)), device=X.device) * torch.tensor([1e-2, 1e-6], device=X.device)
    U1 = U[:, :1].clone(); U1[:, 0] += du
    Xp = eng.rollout_st((x0 + dx).contiguous(), U1, kap[:, :1].contiguous(), ds[:, :1].contiguous())
    Xm = eng.rollout_st((x0 - dx).contiguous(), (2 * U[:, :1] - U1).contiguous(), kap[:, :1].contiguous(), ds[:, :1].contiguous())
    fd = (Xp[:, 1] - Xm[:, 1]) / 2
    lin = torch.einsum("bij,bj->bi", AB[:, 0], torch.cat([dx, du], 1))
    err = ((fd - lin).abs().max() / lin.abs().max()).item()
    print(f"C5: central finite difference vs [A|B]: {err:.2e}")
    assert err < 1e-5


def test_c5_edge_shapes():
    from bench_common import golden_states, rollout_inputs
    from vehicle_control_b200.engine import Engine
    from vehicle_control_b200.environment.track import Track
    tr = Track(load_yaml("environment/ippodromo.yaml"))
    eng = Engine(load_yaml("models/dynamic_car.yaml"), load_yaml("controllers/cascaded.yaml"), tr)
    full = [eng.tensor(a) for a in rollout_inputs(37, 7, 11, tr, golden_states())]
    Xf, ABf = eng.rollout_st(*full, jac=True)
    for B, N in ((1, 1), (1, 7), (37, 1), (17, 3)):
        part = [full[0][:B].contiguous(), full[1][:B, :N].contiguous(), full[2][:B, :N].contiguous(), full[3][:B, :N].contiguous()]
        X, AB = eng.rollout_st(*part, jac=True)
        assert torch.equal(X, Xf[:B, :N + 1]) and torch.equal(AB, ABf[:B, :N])
