"""Per-tick parity of the solver ALGORITHM against IPOPT over every saved lap of the reference, on the CPU oracle.

Test infrastructure (oracle/cpp = the algorithm the CUDA engine implements iteration for iteration, tests/
test_gpu_bench_parity.py).  Every saved run (tests/golden/runs.npz) is replayed in FOLLOW mode: the plant is advanced with
the reference's own saved actions, so each tick's initial state is the reference's, and the warm start is our previous
solution (cascaded_mpc.py:316-338 protocol).  Per tick the first action is compared with IPOPT's saved one.
A closed-loop replay hides what this shows: most "first mismatch" ticks of a closed loop are 1e-6..3e-6 deviations on a
weakly determined Fx (both solvers stop inside tol = 1e-8 of the same KKT point) that then accumulate in the state;
real branch events -- a different local solution of the non-convex NLP -- are few and come in short clusters, because
the next ticks are warm-started from our branch until the two solutions merge again.

For the first tick of every cluster the instance is solved again with the second-order correction switched off and
from IPOPT-like multiplier initialisation, and the objectives are reported: which KKT point is lower, which option
reproduces IPOPT's saved action.

    python scripts/branch_ticks.py [--out profiles/r02_branch_ticks.md] [run ...]
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


def follow(key, meta, z, draws):
    m = meta[key]
    c = on.MpcCfg.from_config(m["config"])
    tr = track(m["track"])
    rs = np.random.RandomState(31)
    for h in draws:
        rs.random_sample((2, h))
    X, U = on.cold_start(c, rs.random_sample((2, c.H)))
    S, A = z[key + "/state"], z[key + "/action"]
    ob = np.array([o[:3] for o in tr.obstacles], dtype=float) if c.obstacles else None
    cc = co.mpc_config(c)
    rows = []
    for k in range(len(A) - 1):
        x = S[k]
        ds, kap = on.horizon_params(c, tr, x, X[0])
        r = co.solve(cc, x, kap, ds, X, U, obstacles=ob, threads=1, max_soc=4, watchdog=10)
        if int(r["status"]) > 1:
            rows.append((k, np.nan, int(r["iters"]), int(r["status"]), np.nan, None))
            continue   # keep the warm start
        e = float((np.abs(r["U"][:, 0] - A[k + 1]) / np.maximum(1.0, np.abs(A[k + 1]))).max())
        inst = (x, kap, ds, X, U, ob) if e > 1e-3 else None
        rows.append((k, e, int(r["iters"]), 0, float(r["obj"]), inst))
        X, U = r["X"], r["U"]
        if k == 0 and e > 1e-3:
            break    # cold guess not reproduced (construction order / settings not in the saved files)
    return c, cc, rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("keys", nargs="*")
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "r02_branch_ticks.md"))
    a = ap.parse_args()
    meta = json.load(open(os.path.join(ROOT, "tests", "golden", "runs.json")))
    z = np.load(os.path.join(ROOT, "tests", "golden", "runs.npz"))
    keys = a.keys or sorted(meta)
    lines, clusters_all = [], []
    tot = {"ticks": 0, "le1e6": 0, "le1e5": 0, "big": 0, "fail": 0, "clusters": 0}
    for key in keys:
        sim, name = key.split("/")
        others = [k.split("/")[1] for k in meta if k.startswith(sim + "/") and k != key]
        cands = [[]] + [[int(meta[f"{sim}/{o}"]["config"]["horizon"]) + int(meta[f"{sim}/{o}"]["config"].get("horizon_pm", 0))] for o in others]
        t0 = time.time()
        best = None
        for d in cands:
            c, cc, rows = follow(key, meta, z, d)
            if best is None or (np.isfinite(rows[0][1]) and rows[0][1] < best[2][0][1]):
                best = (c, cc, rows, d)
            if np.isfinite(rows[0][1]) and rows[0][1] <= 1e-5:
                break
        c, cc, rows, d = best
        e = np.array([r[1] for r in rows])
        if len(rows) == 1 and e[0] > 1e-3:
            lines.append(f"| {key} | {c.N}+{c.M} | - | - | - | - | - | tick 0 differs: the run was recorded with settings the saved files do not contain |")
            continue
        fin = np.isfinite(e)
        big = np.flatnonzero(fin & (e > 1e-3))
        cl = [] if len(big) == 0 else [g for g in np.split(big, np.flatnonzero(np.diff(big) > 3) + 1)]
        tot["ticks"] += len(e); tot["le1e6"] += int((e[fin] <= 1e-6).sum()); tot["le1e5"] += int((e[fin] <= 1e-5).sum())
        tot["big"] += len(big); tot["fail"] += int((~fin).sum()); tot["clusters"] += len(cl)
        mid = e[fin & (e > 1e-6) & (e <= 1e-3)]
        lines.append(f"| {key} | {c.N}+{c.M}{' obst' if c.obstacles else ''} | {len(e)} | {int((e[fin] <= 1e-6).sum())} | {int((e[fin] <= 1e-5).sum())} | "
                     f"{len(big)} in {len(cl)} cluster(s): {', '.join(f'{g[0]}-{g[-1]}' for g in cl)} | {int((~fin).sum())} | "
                     f"largest non-branch deviation {mid.max() if len(mid) else 0:.1e}; mean {np.mean([r[2] for r in rows]):.1f} iterations ({time.time() - t0:.0f} s) |")
        print(lines[-1], flush=True)
        A = z[key + "/action"]
        for g in cl:     # first tick of each cluster: which option lands on IPOPT's answer?
            k = int(g[0])
            x, kap, ds, X, U, ob = rows[k][5]
            ours = rows[k][4]
            alt = []
            for label, kw in (("max_soc 0", dict(max_soc=0, watchdog=10)), ("watchdog off", dict(max_soc=4, watchdog=0)),
                              ("max_soc 0, watchdog off", dict(max_soc=0, watchdog=0))):
                r = co.solve(cc, x, kap, ds, X, U, obstacles=ob, threads=1, **kw)
                err = float((np.abs(r["U"][:, 0] - A[k + 1]) / np.maximum(1.0, np.abs(A[k + 1]))).max())
                alt.append((label, int(r["status"]), float(r["obj"]), err))
            hit = [al for al in alt if al[1] <= 1 and al[3] <= 1e-5]
            clusters_all.append(f"| {key} | {k} | {len(g)} | {ours:.8f} | " + " ; ".join(f"{l}: obj {o:.8f}, |du0| {er:.1e}" for l, s, o, er in alt) +
                                f" | {('IPOPT point = ' + hit[0][0] + f', objective {hit[0][2]:.8f} (' + ('lower' if hit[0][2] < ours else 'higher') + ' than ours)') if hit else 'none of the variants reproduces IPOPT here'} |")
            print(clusters_all[-1], flush=True)
    with open(a.out, "w") as f:
        f.write("# r02 -- per-tick parity with IPOPT over the reference's saved laps (follow mode, CPU oracle = the engine's algorithm)\n\n")
        f.write(__doc__.split("    python")[0].split("\n", 2)[2])
        f.write(f"\nTotals over {len(lines)} controller laps: {tot['ticks']} ticks, {tot['le1e6']} within 1e-6 of IPOPT's saved action "
                f"({100.0 * tot['le1e6'] / max(1, tot['ticks']):.1f} %), {tot['le1e5']} within 1e-5, {tot['big']} ticks in {tot['clusters']} branch clusters, "
                f"{tot['fail']} ticks where the oracle's solve failed (no restoration phase).\n\n")
        f.write("| run | horizon | ticks | <= 1e-6 | <= 1e-5 | > 1e-3 (branch clusters: first-last tick) | failed | notes |\n|---|---|---|---|---|---|---|---|\n")
        f.write("\n".join(lines) + "\n\n")
        f.write("## first tick of every branch cluster\n\n| run | tick | cluster length | our objective (max_soc 4, watchdog 10) | the same instance with other options | verdict |\n|---|---|---|---|---|---|\n")
        f.write("\n".join(clusters_all) + "\n")
    print("wrote", a.out)


if __name__ == "__main__":
    main()
