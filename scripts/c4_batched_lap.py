"""BASELINE config C4: cascaded 20+35 NMPC on the shoe track, closed-loop lap of B vehicles in lock step on one GPU.

    python scripts/c4_batched_lap.py [B=16384] [max_ticks=1200] [lanes=1]

lanes > 1: the fleet is split into that many groups, each a closed loop of its own on a lane (pipeline.run_fleet_in_lanes).

Every tick: horizon parameters -> batched NMPC solve -> batched plant step -> termination mask, all device resident
(vehicle_control_b200.simulation.racing.BatchedRacingSimulator).  Prints one JSON line.
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from bench_common import ROOT, load_yaml
from vehicle_control_b200.simulation.racing import BatchedRacingSimulator


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
    max_ticks = int(sys.argv[2]) if len(sys.argv) > 2 else 1200
    meta = json.load(open(os.path.join(ROOT, "tests", "golden", "runs.json")))
    cfg = dict(meta["race2_shoe/cascaded"]["config"]); cfg["horizon_pm"] = 35      # race2_shoe config with horizon_pm 35
    rng = np.random.default_rng(1004)
    x0 = np.zeros((B, 8)); x0[:, 0] = 4.0; x0[:, 4] = 1.0 + rng.uniform(0, 5, B); x0[:, 5] = rng.uniform(-1, 1, B)
    np.random.seed(31)
    sim = BatchedRacingSimulator(load_yaml("models/dynamic_car.yaml"), load_yaml("environment/shoe.yaml"), cfg, x0, record=False)
    lanes = int(sys.argv[3]) if len(sys.argv) > 3 else 1
    if lanes > 1:
        from vehicle_control_b200.pipeline import run_fleet_in_lanes
        del sim
        torch.cuda.synchronize(); t0 = time.perf_counter()
        sims = run_fleet_in_lanes(load_yaml("models/dynamic_car.yaml"), load_yaml("environment/shoe.yaml"), cfg, x0, lanes,
                                  max_ticks=max_ticks)
        torch.cuda.synchronize(); el = time.perf_counter() - t0
        x = np.concatenate([s_.x.cpu().numpy() for s_ in sims])
        done = x[:, 4] > sims[0].track.length - 0.1
        lt = x[done, 7]
        solves = sum(int(s_.total_solves) for s_ in sims)
        retries = {k: sum(s_.controller.retries[k] for s_ in sims) for k in sims[0].controller.retries}
        print(json.dumps({"config": "C4: cascaded 20+35, shoe, closed-loop lap, %d groups of vehicles on lanes of their own" % lanes,
                          "vehicles": B, "lanes": lanes, "ticks": int(max(int(s_.ticks.max()) for s_ in sims)),
                          "finished": int(done.sum()), "failed": int(sum(int(s_.failed.sum()) for s_ in sims)),
                          "total_solves": solves, "wall_s": el, "solves_per_s": solves / el, "retries": retries,
                          "laptime_s": {"min": float(lt.min()) if len(lt) else None, "median": float(np.median(lt)) if len(lt) else None,
                                        "max": float(lt.max()) if len(lt) else None},
                          "note": "wall time includes building the lanes' simulators"}))
        return
    torch.cuda.synchronize(); t0 = time.perf_counter()
    n, it_sum, it_n = 0, 0.0, 0
    while n < max_ticks:
        act = sim.update()
        if act == 0:
            break
        n += 1
        if n % 50 == 0:
            it = sim.controller.info["iters"].float()
            print(f"tick {n}: active {act} failed {int(sim.failed.sum())} mean iters {it.mean().item():.1f} max {int(it.max())} "
                  f"elapsed {time.perf_counter() - t0:.0f}s", file=sys.stderr, flush=True)
    torch.cuda.synchronize(); el = time.perf_counter() - t0
    x = sim.x.cpu().numpy()
    done = x[:, 4] > sim.track.length - 0.1
    lt = x[done, 7]
    print(json.dumps({"config": "C4: cascaded 20+35, shoe, closed-loop lap, lock step", "vehicles": B, "ticks": n,
                      "finished": int(done.sum()), "failed": int(sim.failed.sum()), "total_solves": int(sim.total_solves),
                      "wall_s": el, "solves_per_s": sim.total_solves / el,
                      "retries": sim.controller.retries,
                      "laptime_s": {"min": float(lt.min()) if len(lt) else None, "median": float(np.median(lt)) if len(lt) else None,
                                    "max": float(lt.max()) if len(lt) else None},
                      "reference_laptime_s_20+25": 44.40}))


if __name__ == "__main__":
    main()
