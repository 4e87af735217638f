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
        ex = max(_rel(X[0, :, :N], f["X"][:, :N]).max(), _rel(X[0, :5, N:], f["X"][:5, N:]).max())
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
