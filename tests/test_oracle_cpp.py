"""CPU: the C++ oracle (oracle/cpp -> oracle/_ref/libvc_oracle.so) against the reference's golden data and against the
independent Python oracle (oracle/nlp.py + oracle/ipm.py: torch autodiff + dense LDL^T).

The two oracles share no code: different derivative mechanism (second-order jets vs torch.func), different KKT
factorisation (block-tridiagonal Bunch-Kaufman LDL^T vs dense scipy LDL^T), different language.  With the second-order
correction switched off they run the same published algorithm, so they must agree iterate for iterate.
"""
import json
import os

import numpy as np
import pytest

from oracle import cpp_oracle as co, models as om, nlp as on
from tests.conftest import ROOT


@pytest.fixture(scope="module")
def fx():
    z = np.load(os.path.join(ROOT, "tests", "golden", "nmpc_oracle.npz"))
    with open(os.path.join(ROOT, "tests", "golden", "runs.json")) as f:
        meta = json.load(f)
    return z, sorted({k.rsplit("/", 1)[0] for k in z.files}), meta


def test_library_builds_and_loads():
    co.build()
    assert co.load().vco_version() >= 2 and co.num_threads() >= 1


def test_plant_step_reproduces_golden_runs(golden, oracle_tracks):
    """X[k+1] = RK4_t(X[k], U[k+1], k(s_k), 0.05) for all 44 saved reference runs (racing_car.py:34-46)."""
    z, meta = golden
    worst, steps = 0.0, 0
    for key, m in meta.items():
        X, U = z[key + "/state"], z[key + "/action"]
        Xn = co.plant_step(X[:-1], U[1:], oracle_tracks[m["track"]].k(X[:-1, 4]), 0.05)
        worst = max(worst, (np.abs(Xn - X[1:]) / np.maximum(1.0, np.abs(X[1:]))).max())
        steps += len(Xn)
    assert steps == 23997 and worst < 2e-9, worst


def test_rollout_and_jacobians_match_python_oracle(golden, oracle_tracks):
    """Spatial RK4 rollout + [A|B] (dynamic_car.py:188-191): jets vs torch forward-mode autodiff, <= 1e-10."""
    from bench_common import golden_states, rollout_inputs
    x0, U, kap, ds = rollout_inputs(48, 6, 11, oracle_tracks["ippodromo"], golden_states())
    X, AB = co.rollout_st(x0, U, kap, ds, jac=True)
    Xo = om.st_rollout_spatial(om.DEFAULT, x0, U, kap, ds)
    assert (np.abs(X - Xo) / np.maximum(1, np.abs(Xo))).max() < 1e-12
    for n in (0, 5):
        J = om.st_step_jac(om.DEFAULT, Xo[:, n], U[:, n], kap[:, n], ds[:, n])
        assert (np.abs(AB[:, n] - J) / np.maximum(1, np.abs(J))).max() < 1e-10


def test_nmpc_fixtures_iterate_for_iterate(fx):
    """Same published algorithm (no second-order correction) on the 12 fixture ticks: same iteration count, controls
    <= 1e-9, objective <= 1e-12 against the Python oracle; first action <= 1e-6 against IPOPT's saved one."""
    z, tags, meta = fx
    for t in tags:
        cfg = on.MpcCfg.from_config(meta[t.rsplit("/", 1)[0]]["config"])
        ob = z[t + "/obst"] if cfg.obstacles else None
        r = co.solve(cfg, z[t + "/x0"], z[t + "/kappa"], z[t + "/ds"], z[t + "/X0"], z[t + "/U0"], obstacles=ob, max_soc=0, watchdog=0)
        U, ref, obj = z[t + "/U"], z[t + "/ref_u"], float(z[t + "/obj"])
        assert r["status"] == 0 and int(r["iters"]) == int(z[t + "/iters"]), (t, r["iters"], z[t + "/iters"])
        assert (np.abs(r["U"] - U) / np.maximum(1, np.abs(U))).max() < 1e-9, t
        assert abs(r["obj"] - obj) < 1e-12 * max(1, abs(obj)), t
        assert (np.abs(r["U"][:, 0] - ref) / np.maximum(1, np.abs(ref))).max() < 1e-6, t


def test_nmpc_with_second_order_correction_reaches_ipopts_answers(fx):
    """IPOPT's default (max_soc = 4) changes the path, not the point: same fixtures, first action <= 1e-6 vs IPOPT."""
    z, tags, meta = fx
    B = len(tags)
    for t in tags:
        cfg = on.MpcCfg.from_config(meta[t.rsplit("/", 1)[0]]["config"])
        ob = z[t + "/obst"] if cfg.obstacles else None
        r = co.solve(cfg, z[t + "/x0"], z[t + "/kappa"], z[t + "/ds"], z[t + "/X0"], z[t + "/U0"], obstacles=ob, max_soc=4)
        ref = z[t + "/ref_u"]
        assert r["status"] == 0
        assert (np.abs(r["U"][:, 0] - ref) / np.maximum(1, np.abs(ref))).max() < 1e-6, t
        assert (np.abs(r["U"] - z[t + "/U"]) / np.maximum(1, np.abs(z[t + "/U"]))).max() < 1e-6, t
    assert B >= 12


def test_eval_matches_python_nlp(fx):
    z, tags, meta = fx
    for t in tags[:4] + tags[-2:]:
        cfg = on.MpcCfg.from_config(meta[t.rsplit("/", 1)[0]]["config"])
        ob = z[t + "/obst"] if cfg.obstacles else None
        p = on.NLP(cfg, om.DEFAULT, z[t + "/x0"], z[t + "/kappa"], z[t + "/ds"], ob)
        f, g, c = p.values(p.pack(z[t + "/X0"], z[t + "/U0"]))
        e = co.evaluate(cfg, z[t + "/x0"][None], z[t + "/kappa"][None], z[t + "/ds"][None], z[t + "/X0"][None],
                        z[t + "/U0"][None], obstacles=None if ob is None else ob[None])[0]
        assert abs(e[0] - f) < 1e-12 * max(1, abs(f))
        assert abs(e[1] - np.abs(g).max()) < 1e-12 * max(1, np.abs(g).max()) and abs(e[2] - c.min()) < 1e-9 * max(1, abs(c.min()))


def test_batch_is_thread_invariant(fx):
    """OpenMP over instances: results do not depend on the thread count."""
    z, tags, meta = fx
    ts = [t for t in tags if t.startswith("cascaded3")]
    cfg = on.MpcCfg.from_config(meta["cascaded3_ippodromo/cascaded"]["config"])
    st = lambda k: np.stack([z[t + "/" + k] for t in ts])
    a = co.solve(cfg, st("x0"), st("kappa"), st("ds"), st("X0"), st("U0"), threads=1)
    b = co.solve(cfg, st("x0"), st("kappa"), st("ds"), st("X0"), st("U0"), threads=3)
    assert np.array_equal(a["U"], b["U"]) and np.array_equal(a["iters"], b["iters"])
