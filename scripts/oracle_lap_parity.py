"""Closed-loop parity of the CPU ORACLE (oracle/cpp) against the reference's saved runs, tick by tick.

Test infrastructure, not product: replays a saved run of the reference (tests/golden/runs.npz) with the C++ oracle as the
controller (cold guess with the legacy MT19937 seed 31, un-shifted warm start, cascaded_mpc.py:13,72-76,316-338) and the
oracle's plant step, and reports the first tick whose action differs from IPOPT's saved one by more than 1e-6 relative.
Used to study which of IPOPT's mechanisms (second-order correction, watchdog) decide the branch ticks.

    python scripts/oracle_lap_parity.py [--soc 4] [--watchdog 10] [--ticks 250] run_key ...
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from bench_common import ROOT, load_yaml
from oracle import cpp_oracle as co, nlp as on
from oracle.track import OracleTrack

_tracks = {}


def track(name):
    if name not in _tracks:
        _tracks[name] = OracleTrack(load_yaml(f"environment/{name}.yaml"))
    return _tracks[name]


def replay(key, meta, z, opts, max_ticks, skip_draws=0, follow=False, verbose=False):
    m = meta[key]
    c = on.MpcCfg.from_config(m["config"])
    tr = track(m["track"])
    rs = np.random.RandomState(31)
    for h in ([] if not skip_draws else skip_draws):
        rs.random_sample((2, h))
    X, U = on.cold_start(c, rs.random_sample((2, c.H)))
    S, A = z[key + "/state"], z[key + "/action"]
    x = S[0].copy()
    ob = np.array([o[:3] for o in tr.obstacles], dtype=float) if c.obstacles else None
    cc = co.mpc_config(c)
    first, iters, errs = None, [], []
    T = min(len(A) - 1, max_ticks)
    for k in range(T):
        ds, kap = on.horizon_params(c, tr, x, X[0])
        r = co.solve(cc, x, kap, ds, X, U, obstacles=ob, threads=1, **opts)
        if int(r["status"]) > 1:
            print(f"   tick {k}: oracle status {co.STATUS_NAMES[int(r['status'])]} after {int(r['iters'])} iterations")
            break
        X, U = r["X"], r["U"]
        iters.append(int(r["iters"]))
        e = (np.abs(U[:, 0] - A[k + 1]) / np.maximum(1.0, np.abs(A[k + 1]))).max()
        errs.append(e)
        if e > 1e-6 and first is None:
            first = k
            if verbose:
                print(f"   tick {k}: u0 {U[:, 0]} ref {A[k + 1]} err {e:.2e} obj {float(r['obj']):.8f} iters {int(r['iters'])}")
            if not follow:
                break
        u = A[k + 1] if follow else U[:, 0]
        x = co.plant_step(x[None], u[None], np.atleast_1d(tr.k(x[4])), 0.05)[0]
    return first, iters, errs


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("keys", nargs="*")
    ap.add_argument("--soc", type=int, default=4)
    ap.add_argument("--watchdog", type=int, default=-1)
    ap.add_argument("--ticks", type=int, default=10000)
    ap.add_argument("--follow", action="store_true", help="advance the plant with the reference's saved actions")
    a = ap.parse_args()
    meta = json.load(open(os.path.join(ROOT, "tests", "golden", "runs.json")))
    z = np.load(os.path.join(ROOT, "tests", "golden", "runs.npz"))
    opts = {"max_soc": a.soc}
    if a.watchdog >= 0:
        opts["watchdog"] = a.watchdog
    keys = a.keys or ["singletrack3_ippodromo/singletrack", "cascaded3_ippodromo/cascaded", "cascaded1_ippodromo/cascaded",
                      "cascaded4_ippodromo/cascaded", "cascaded5_ippodromo/cascaded", "cascaded6_ippodromo/cascaded"]
    for key in keys:
        t0 = time.time()
        first, iters, errs = replay(key, meta, z, opts, a.ticks, follow=a.follow, verbose=True)
        print(f"{key}: first mismatch tick {first}; {len(iters)} ticks replayed, mean iters {np.mean(iters):.1f} max {max(iters)}; "
              f"max err before {max(errs[:first] if first else errs):.2e}  ({time.time() - t0:.0f} s)", flush=True)


if __name__ == "__main__":
    main()
