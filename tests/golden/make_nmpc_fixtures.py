"""Generates tests/golden/nmpc_oracle.npz: NMPC instances solved by the CPU oracle (oracle/nlp.py + oracle/ipm.py).

For each listed reference run the closed loop is replayed from tick 0 with the reference's protocol
(cold guess with np.random.seed(31), un-shifted warm start, ds from the previous predicted Ux;
cascaded_mpc.py:13,72-76,316-338) for a few ticks, the plant being advanced with the reference's own
saved actions (so every tick's x0 is the reference's).  Each tick stores the solver inputs
(x0, kappa, ds, obstacles, X/U guess) and the oracle's answer, next to the reference's saved first action
(action_traj[k+1]) that pins it.  The GPU parity tests replay these instances through the C-ABI.

    python tests/golden/make_nmpc_fixtures.py            (minutes of CPU; output committed)
"""
import json
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

RUNS = [  # (run key, ticks, draws of random((2,H)) consumed before this controller's, see racing.py:60-63)
    ("cascaded1_ippodromo/cascaded", 3, []),
    ("singletrack3_ippodromo/singletrack", 2, []),
    ("cascaded3_ippodromo/cascaded", 3, []),
    ("cascaded_obstacles1_ippodromo/cascaded", 2, []),
    ("race2_shoe/cascaded", 2, ["singletrack"]),
]


def main(only=None):
    meta = json.load(open(os.path.join(HERE, "runs.json")))
    z = np.load(os.path.join(HERE, "runs.npz"))
    tracks = {}
    out = {}
    for key, ticks, before in RUNS:
        if only and only not in key:
            continue
        m = meta[key]
        cfg = on.MpcCfg.from_config(m["config"])
        tn = m["track"]
        if tn not in tracks:
            tracks[tn] = OracleTrack(yaml.safe_load(open(os.path.join(
                ROOT, "vehicle_control_b200", "config", "environment", tn + ".yaml"))))
        tr = tracks[tn]
        np.random.seed(31)
        for other in before:   # controllers constructed earlier draw first
            mo = meta[key.split("/")[0] + "/" + other]["config"]
            np.random.random((2, int(mo["horizon"]) + int(mo["horizon_pm"])))
        X, U = on.cold_start(cfg, np.random.random((2, cfg.H)))
        S, A = z[key + "/state"], z[key + "/action"]
        obst = np.array([o[:3] for o in tr.obstacles]) if cfg.obstacles else np.zeros((0, 3))
        for k in range(ticks):
            x0 = S[k]
            ds, kap = on.horizon_params(cfg, tr, x0, X[0])
            p = on.NLP(cfg, om.DEFAULT, x0, kap, ds, obst if cfg.obstacles else None)
            t0 = time.time()
            r = ipm.solve(p, p.pack(X, U))
            Xs, Us = p.unpack(r.w)
            tag = f"{key}/tick{k}"
            print(f"{tag}: {r.status} it={r.iters} obj={r.obj:.10f} sf={r.obj_scale:.6g} mu={r.mu:.3e} "
                  f"u0=({Us[0, 0]:.10f}, {Us[1, 0]:.6e}) ref=({A[k + 1, 0]:.10f}, {A[k + 1, 1]:.6e}) "
                  f"{time.time() - t0:.0f}s", flush=True)
            for name, arr in (("x0", x0), ("kappa", kap), ("ds", ds), ("obst", obst), ("X0", X), ("U0", U),
                              ("X", Xs), ("U", Us), ("obj", np.array(r.obj)), ("iters", np.array(r.iters)),
                              ("mu", np.array(r.mu)), ("sf", np.array(r.obj_scale)), ("ref_u", A[k + 1])):
                out[f"{tag}/{name}"] = np.asarray(arr, dtype=float)
            if r.status not in ("solved", "acceptable"):
                break
            X, U = Xs, Us
    path = os.path.join(HERE, "nmpc_oracle.npz" if not only else f"nmpc_oracle_{only.replace('/', '_')}.npz")
    np.savez_compressed(path, **out)
    print("wrote", path)


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else None)
