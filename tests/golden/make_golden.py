"""Packs the reference's saved closed-loop runs (experiments/data/*) into small fixtures.

Run ONCE in the build container (needs /root/reference; the GPU box does not have it):
    python tests/golden/make_golden.py
Outputs (committed):
    tests/golden/runs.npz    state_traj / action_traj / elapsed of all 44 controller runs,
                             preds (T,H,3) for a few runs
    tests/golden/runs.json   the controller config saved next to each run + index
The arrays were written by the reference's RacingSimulator.save (simulation/racing.py:441-450).
"""
import glob
import json
import os

import numpy as np
import yaml

REF = "/root/reference/experiments/data"
HERE = os.path.dirname(os.path.abspath(__file__))
PREDS_FOR = {"cascaded3_ippodromo/cascaded", "singletrack3_ippodromo/singletrack",
             "race2_shoe/cascaded", "cascaded1_ippodromo/cascaded"}


def main():
    arrays, meta = {}, {}
    for cfgf in sorted(glob.glob(os.path.join(REF, "*", "*_config.yaml"))):
        run = os.path.basename(os.path.dirname(cfgf))
        ctrl = os.path.basename(cfgf)[: -len("_config.yaml")]
        key = f"{run}/{ctrl}"
        base = os.path.join(REF, run, ctrl)
        if not os.path.exists(base + "_state_traj.npy"):
            continue
        arrays[key + "/state"] = np.load(base + "_state_traj.npy")
        arrays[key + "/action"] = np.load(base + "_action_traj.npy")
        arrays[key + "/elapsed"] = np.load(base + "_elapsed.npy")
        if key in PREDS_FOR:
            arrays[key + "/preds"] = np.load(base + "_preds.npy")
        with open(cfgf) as f:
            meta[key] = {"config": yaml.safe_load(f),
                         "track": "shoe" if run.endswith("_shoe") else "ippodromo"}
    np.savez_compressed(os.path.join(HERE, "runs.npz"), **arrays)
    with open(os.path.join(HERE, "runs.json"), "w") as f:
        json.dump(meta, f, indent=1, sort_keys=True)
    print(len(meta), "runs;", sum(a.nbytes for a in arrays.values()) / 1e6, "MB raw")


if __name__ == "__main__":
    main()
