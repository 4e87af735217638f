"""Generates tests/golden/nmpc_oracle_variants.npz: NMPC instances on the tracks and the parameter set the reference
ships but never saved a run for (config/environment/bean.yaml, gokart.yaml,
config/models/paper_dynamic_car.yaml), solved by the CPU oracle alone.

There is no reference artefact for these configurations, so the fixtures pin the CUDA path to the oracle only
("parity unpinned by the reference" for these rows).  Protocol = the reference's closed loop (racing.py:49,217-242,
cascaded_mpc.py:13,72-76,316-338): start state (Ux=4, s=1), cold guess from np.random.seed(31), un-shifted warm
start, plant advanced by the temporal RK4 with the oracle's own first action.

    python tests/golden/make_variant_fixtures.py         (minutes of CPU; output committed)
"""
import os
import sys
import time

import numpy as np
import yaml

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ipm, models as om, nlp as on  # noqa: E402
from oracle.track import OracleTrack  # noqa: E402

CFG = os.path.join(ROOT, "vehicle_control_b200", "config")
CASES = [  # (tag, track, model yaml, controller yaml, overrides, ticks)
    ("bean_paper_cascaded", "bean", "paper_dynamic_car", "cascaded", {"horizon_pm": 15}, 2),
    ("gokart_default_cascaded", "gokart", "dynamic_car", "cascaded", {"horizon_pm": 15}, 2),
    ("bean_paper_obstacles", "bean", "paper_dynamic_car", "cascaded", {"horizon_pm": 15, "obstacles": True}, 1),
]


def load(kind, name):
    with open(os.path.join(CFG, kind, name + ".yaml")) as f:
        return yaml.safe_load(f)


def main():
    out = {}
    for tag, tn, mn, cn, over, ticks in CASES:
        ccfg = load("controllers", cn)
        ccfg.update(over)
        ccfg.setdefault("obstacles", False)
        cfg = on.MpcCfg.from_config(ccfg)
        par = om.CarParams.from_config(load("models", mn))
        tr = OracleTrack(load("environment", tn))
        np.random.seed(31)
        X, U = on.cold_start(cfg, np.random.random((2, cfg.H)))
        x = np.array([4.0, 0, 0, 0, 1.0, 0, 0, 0])
        obst = np.array([o[:3] for o in tr.obstacles]) if cfg.obstacles else np.zeros((0, 3))
        for k in range(ticks):
            ds, kap = on.horizon_params(cfg, tr, x, X[0])
            p = on.NLP(cfg, par, x, kap, ds, obst if cfg.obstacles else None)
            t0 = time.time()
            r = ipm.solve(p, p.pack(X, U))
            Xs, Us = p.unpack(r.w)
            print(f"{tag}/tick{k}: {r.status} it={r.iters} obj={r.obj:.10f} u0=({Us[0, 0]:.10f}, {Us[1, 0]:.6e}) "
                  f"{time.time() - t0:.0f}s", flush=True)
            xn = om.st_step_temporal(np, par, x[None], Us[:, 0][None], np.array([float(tr.k(x[4]))]), par.dt)[0]
            for name, arr in (("x0", x), ("kappa", kap), ("ds", ds), ("obst", obst), ("X0", X), ("U0", U),
                              ("X", Xs), ("U", Us), ("obj", np.array(r.obj)), ("iters", np.array(r.iters)),
                              ("x_next", xn)):
                out[f"{tag}/tick{k}/{name}"] = np.asarray(arr, dtype=float)
            assert r.status in ("solved", "acceptable"), r.status
            X, U, x = Xs, Us, xn
    path = os.path.join(HERE, "nmpc_oracle_variants.npz")
    np.savez_compressed(path, **out)
    print("wrote", path)


if __name__ == "__main__":
    main()
