"""Closed-loop parity survey: every saved run of the reference (tests/golden/runs.npz, 44 controller laps in 34
simulations) replayed on the engine through the RacingSimulator mirror, tick by tick against the saved
*_state_traj / *_action_traj arrays.

For each controller lap it reports the tick count and lap time (ours / reference), the first tick whose action differs
from IPOPT's by more than 1e-6 (relative), the largest state error before that tick, and the iteration statistics.
Simulations with two controllers ("race*") tick both in one simulator like racing.py:217-242; the legacy MT19937 cold
guesses are drawn in construction order (cascaded_mpc.py:13,72-76), which the saved files do not record, so both orders
are tried and the one that reproduces tick 0 is kept.

    python scripts/lap_parity_all.py [--out gpurun_out/lap_parity_all.json] [sim ...]
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from bench_common import ROOT, load_yaml
from vehicle_control_b200.simulation.racing import RacingSimulator


def run_group(sim_name, names, meta, z):
    m0 = meta[f"{sim_name}/{names[0]}"]
    np.random.seed(31)
    sc = {"track_name": m0["track"], "controller_names": list(names), "obstacles": bool(m0["config"].get("obstacles")),
          "slip_angle_plot": False, "name": "cmp", "save_gif": False, "save_images": False, "save_data": False,
          "load": False, "logging": False}
    sim = RacingSimulator(sc, load_yaml("models/dynamic_car.yaml"), load_yaml(f"environment/{m0['track']}.yaml"),
                          controller_configs=[meta[f"{sim_name}/{n}"]["config"] for n in names])
    ref_len = max(len(z[f"{sim_name}/{n}/state"]) for n in names)
    t0, n = time.time(), 0
    while n < ref_len + 100 and sim.update(n):
        n += 1
    rows = {}
    for name in names:
        key = f"{sim_name}/{name}"
        S, A = z[key + "/state"], z[key + "/action"]
        st, ac, _, _ = sim._arrays(name)
        T = min(len(ac), len(A)) - 1
        eu = (np.abs(ac[1:T + 1] - A[1:T + 1]) / np.maximum(1.0, np.abs(A[1:T + 1]))).max(1) if T > 0 else np.zeros(0)
        ex = (np.abs(st[:T + 1] - S[:T + 1]) / np.maximum(1.0, np.abs(S[:T + 1]))).max(1)
        bad = np.flatnonzero(eu > 1e-6)
        first = int(bad[0]) if len(bad) else None
        nb = T if first is None else first
        it = np.array(sim.iters[name]) if len(sim.iters[name]) else np.zeros(1)
        rows[key] = {
            "track": m0["track"], "N": int(meta[key]["config"]["horizon"]), "M": int(meta[key]["config"].get("horizon_pm", 0)),
            "obstacles": bool(m0["config"].get("obstacles")), "order": list(names),
            "ticks": len(ac) - 1, "ref_ticks": len(A) - 1, "laptime": float(st[-1, 7]), "ref_laptime": float(S[-1, 7]),
            "finished": bool(st[-1, 4] > sim.track.length - 0.1) or bool(sim.finished()),
            "first_action_mismatch_tick": first, "ticks_matching": int(T if first is None else first),
            "max_action_err_before": float(eu[:nb].max()) if nb > 0 else 0.0,
            "max_state_err_before": float(ex[:nb + 1].max()),
            "final_state_err": float(ex[-1]), "mean_iters": float(it.mean()), "max_iters": int(it.max()),
            "wall_s": time.time() - t0,
        }
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/lap_parity_all.json")
    ap.add_argument("sims", nargs="*")
    a = ap.parse_args()
    z = np.load(os.path.join(ROOT, "tests", "golden", "runs.npz"))
    meta = json.load(open(os.path.join(ROOT, "tests", "golden", "runs.json")))
    groups = {}
    for key in meta:
        s, n = key.split("/")
        groups.setdefault(s, []).append(n)
    os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
    out = {}
    for s, names in groups.items():
        if a.sims and s not in a.sims:
            continue
        orders = [sorted(names, reverse=True)]            # singletrack first (the order of config/simconfig.yaml)
        if len(names) > 1:
            orders.append(sorted(names))
        best = None
        for od in orders:
            try:
                rows = run_group(s, od, meta, z)
            except Exception as e:                        # keep the survey going; the failure is part of the result
                rows = {f"{s}/{n}": {"error": repr(e), "order": od} for n in od}
            score = sum(r.get("ticks_matching", -1) for r in rows.values())
            if best is None or score > best[0]:
                best = (score, rows)
            if all(r.get("ticks_matching", 0) > 0 for r in rows.values()):
                break                                     # tick 0 reproduced for every controller: this is the order
        out.update(best[1])
        for k, r in best[1].items():
            print(k, json.dumps(r), flush=True)
        with open(a.out, "w") as f:
            json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
