"""CPU: the NMPC oracle (oracle/nlp.py + oracle/ipm.py) against the reference's saved closed-loop actions.

tests/golden/nmpc_oracle.npz holds oracle solutions of ticks of five reference runs (made by
tests/golden/make_nmpc_fixtures.py); `ref_u` is the reference's own action_traj[k+1] (CasADi + IPOPT + MA27).
Bar (BASELINE.json north_star): optimal controls within 1e-6 relative to max(1,|u|).
"""
import json
import os

import numpy as np
import pytest

from tests.conftest import ROOT, load_yaml

TOL_U = 1e-6


@pytest.fixture(scope="module")
def fx():
    z = np.load(os.path.join(ROOT, "tests", "golden", "nmpc_oracle.npz"))
    tags = sorted({k.rsplit("/", 1)[0] for k in z.files})
    return z, tags


def test_fixture_answers_match_reference_actions(fx):
    z, tags = fx
    assert len(tags) >= 12
    runs = {t.rsplit("/", 1)[0] for t in tags}
    assert {"singletrack3_ippodromo/singletrack", "cascaded3_ippodromo/cascaded", "race2_shoe/cascaded",
            "cascaded_obstacles1_ippodromo/cascaded", "cascaded1_ippodromo/cascaded"} <= runs
    for t in tags:
        u, ref = z[t + "/U"][:, 0], z[t + "/ref_u"]
        err = np.abs(u - ref) / np.maximum(1.0, np.abs(ref))
        assert err.max() < TOL_U, (t, u, ref)


def test_fixture_solutions_are_feasible_kkt_points(fx, golden, oracle_tracks):
    """Every stored solution satisfies the oracle's own NLP: equalities to 1e-7, inequalities >= -1e-7."""
    from oracle import models as om, nlp as on
    z, tags = fx
    _, meta = golden
    for t in tags:
        m = meta[t.rsplit("/", 1)[0]]
        cfg = on.MpcCfg.from_config(m["config"])
        ob = z[t + "/obst"] if cfg.obstacles else None
        p = on.NLP(cfg, om.DEFAULT, z[t + "/x0"], z[t + "/kappa"], z[t + "/ds"], ob)
        f, g, c = p.values(p.pack(z[t + "/X"], z[t + "/U"]))
        assert abs(f - float(z[t + "/obj"])) < 1e-9 * max(1, abs(f))
        assert np.abs(g).max() < 1e-7 and c.min() > -1e-7, t


def test_horizon_params_of_fixtures(fx, golden, oracle_tracks):
    """ds / kappa stored with the fixtures follow _init_horizon (cascaded_mpc.py:316-338)."""
    from oracle import nlp as on
    z, tags = fx
    _, meta = golden
    for t in tags:
        m = meta[t.rsplit("/", 1)[0]]
        cfg = on.MpcCfg.from_config(m["config"])
        ds, kap = on.horizon_params(cfg, oracle_tracks[m["track"]], z[t + "/x0"], z[t + "/X0"][0])
        assert np.array_equal(ds, z[t + "/ds"]) and np.abs(kap - z[t + "/kappa"]).max() < 1e-15


def test_oracle_solves_a_warm_tick_again(fx, golden):
    """Re-runs the oracle IPM on one warm-started tick (about 15 s) and reproduces fixture and reference."""
    import torch
    from oracle import ipm, models as om, nlp as on
    torch.set_num_threads(2)
    z, _ = fx
    _, meta = golden
    t = "singletrack3_ippodromo/singletrack/tick1"
    cfg = on.MpcCfg.from_config(meta["singletrack3_ippodromo/singletrack"]["config"])
    p = on.NLP(cfg, om.DEFAULT, z[t + "/x0"], z[t + "/kappa"], z[t + "/ds"], None)
    r = ipm.solve(p, p.pack(z[t + "/X0"], z[t + "/U0"]))
    assert r.status == "solved"
    _, U = p.unpack(r.w)
    assert np.abs(U - z[t + "/U"]).max() < 1e-6
    ref = z[t + "/ref_u"]
    assert (np.abs(U[:, 0] - ref) / np.maximum(1, np.abs(ref))).max() < TOL_U
