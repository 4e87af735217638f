"""CPU: the oracle restatement against the reference's saved closed-loop runs (tests/golden)."""
import numpy as np
import pytest

from oracle import models as om


def test_plant_step_reproduces_all_golden_runs(golden, oracle_tracks):
    """X[k+1] = RK4_t(X[k], U[k+1], k(s_k), 0.05) for all 44 runs (racing_car.py:34-46)."""
    z, meta = golden
    assert len(meta) == 44
    worst, steps = 0.0, 0
    for key, m in meta.items():
        X, U = z[key + "/state"], z[key + "/action"]
        T = oracle_tracks[m["track"]]
        Xn = om.st_step_temporal(np, om.DEFAULT, X[:-1], U[1:], T.k(X[:-1, 4]), 0.05)
        err = np.abs(Xn - X[1:]) / np.maximum(1.0, np.abs(X[1:]))
        worst = max(worst, err.max())
        steps += len(Xn)
    assert steps == 23997
    assert worst < 2e-9, worst


def test_straight_segment_is_bit_exact(golden, oracle_tracks):
    z, _ = golden
    X, U = z["cascaded3_ippodromo/cascaded/state"], z["cascaded3_ippodromo/cascaded/action"]
    Xn = om.st_step_temporal(np, om.DEFAULT, X[:40], U[1:41], np.zeros(40), 0.05)
    assert np.abs(Xn - X[1:41]).max() < 1e-12


def test_track_known_answers(oracle_tracks):
    ip, sh = oracle_tracks["ippodromo"], oracle_tracks["shoe"]
    assert ip.n_waypoints == 3600 and ip.length == 314.6118402640666
    assert sh.n_waypoints == 8202 and sh.length == 745.4576718474385
    assert len(ip.s_table) == 6291 and len(sh.s_table) == 14908
    assert abs(ip.kappa_table.max() - 0.04706) < 1e-4


@pytest.mark.parametrize("key", ["cascaded3_ippodromo/cascaded", "singletrack3_ippodromo/singletrack",
                                 "race2_shoe/cascaded"])
def test_rel2glob_matches_saved_predictions(golden, oracle_tracks, key):
    """preds[:,0,:] is rel2glob of the state the solve started from (cascaded_mpc.py:340-352)."""
    z, meta = golden
    P, X = z[key + "/preds"], z[key + "/state"]
    T = oracle_tracks[meta[key]["track"]]
    x, y, psi = T.rel2glob(X[:-1, 4], X[:-1, 5], X[:-1, 6])
    assert np.abs(P[:, 0, 0] - x).max() < 1e-9
    assert np.abs(P[:, 0, 1] - y).max() < 1e-9
    assert np.abs(P[:, 0, 2] - psi).max() < 1e-9


def test_jacobian_sanity_vector(golden, oracle_tracks):
    """SURVEY App. B Jacobian spot values (autodiff of the restatement)."""
    z, _ = golden
    x = z["cascaded3_ippodromo/cascaded/state"][200][None]
    u = z["cascaded3_ippodromo/cascaded/action"][201][None]
    kap = oracle_tracks["ippodromo"].k(x[:, 4])
    ds = 0.03 * x[:, 0]
    J = om.st_step_jac(om.DEFAULT, x, u, kap, ds)[0]
    assert abs(J[0, 0] - 0.9959445) < 1e-6 and abs(J[1, 1] - 0.7245994) < 1e-6
    assert abs(J[1, 3] - 1.189958) < 1e-5 and abs(J[7, 0] + 1.745812e-03) < 1e-8
    assert abs(J[0, 8] - 1.802946e-05) < 1e-10 and abs(J[3, 9] - 3.023466e-02) < 1e-7
    assert np.array_equal(J[:, 4], np.eye(8)[4]) and np.array_equal(J[:, 7], np.eye(8)[7])


def test_tyre_forces_match_the_reference_logs(golden):
    """experiments/logs/*.log known answers (tests/golden/log_kat.json): block k = forces at (state[k+1], action[k+1]),
    printed by the reference with 6 significant digits (dynamic_car.py:20-47)."""
    import json
    import os
    from tests.conftest import ROOT
    z, _ = golden
    kat = json.load(open(os.path.join(ROOT, "tests", "golden", "log_kat.json")))
    for key, blocks in kat.items():
        S, A = z[key + "/state"], z[key + "/action"]
        for k, ref in blocks.items():
            x, u = S[int(k) + 1], A[int(k) + 1]
            f = [om.Fx_f(np, om.DEFAULT, u[0]), om.Fx_r(np, om.DEFAULT, u[0]), om.Fy_f(np, om.DEFAULT, *x[:4], u[0]),
                 om.Fy_r(np, om.DEFAULT, *x[:4], u[0])]
            assert all(abs(a - b) <= 6e-6 * abs(b) + 1e-8 for a, b in zip(f, ref)), (key, k, f, ref)
