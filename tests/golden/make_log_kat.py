"""Generates tests/golden/log_kat.json: tyre-force known answers printed by the reference itself.

experiments/logs/<run>.log is written by simulation/racing.py:425-439 through models/dynamic_car.py:20-47 (`car.print`):
block k holds Fx_f, Fx_r, Fy_f, Fy_r of the state / action AFTER tick k, i.e. (state_traj[k+1], action_traj[k+1]) of the
same run (tests/golden/runs.npz), printed with 6 significant digits.  Every 10th block of two runs is kept.

    python tests/golden/make_log_kat.py      (reads /root/reference; output committed)
"""
import json
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
LOGS = "/root/reference/experiments/logs"
RUNS = {"cascaded4_ippodromo": "cascaded4_ippodromo/cascaded", "singletrack3_ippodromo": "singletrack3_ippodromo/singletrack"}


def main():
    out = {}
    for log, key in RUNS.items():
        txt = open(os.path.join(LOGS, log + ".log")).read()
        blocks = re.findall(r"Fx_f: (\S+)\nFx_r: (\S+)\nFy_f: (\S+)\nFy_r: (\S+)", txt)
        out[key] = {str(k): [float(v) for v in blocks[k]] for k in range(0, len(blocks), 10)}
        print(log, len(blocks), "blocks,", len(out[key]), "kept; block 200:", blocks[200] if len(blocks) > 200 else None)
    json.dump(out, open(os.path.join(HERE, "log_kat.json"), "w"), indent=0)


if __name__ == "__main__":
    main()
