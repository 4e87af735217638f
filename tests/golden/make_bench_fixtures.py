"""Generates tests/golden/bench_oracle.npz: instances drawn from the BENCHMARK batches themselves, solved by the CPU oracle.

C3 (bench.py default: cascaded 20+35 + obstacles, 65,536 perturbed instances) and C2 (cascaded 20+15, 4,096): the batch
is rebuilt with bench_nmpc.make_batch from the warm-start lap the engine replayed on the GPU box
(scripts/dump_c3_sample.py writes it to gpurun_out/{c3,c2}_lap.npz), with bench.py's seeds; an evenly spaced -- i.e.
unbiased -- sample of instance indices is taken, horizon parameters come from the oracle's track (oracle/nlp.py
horizon_params, cascaded_mpc.py:316-338), and every instance is solved by the C++ oracle (oracle/cpp, IPOPT's
algorithm with max_soc = 4).  A few are solved again by the independent Python oracle (oracle/ipm.py, no second-order
correction) to double-pin the answers.  The GPU parity test (tests/test_gpu_bench_parity.py) replays the stored inputs
through the C-ABI and compares controls and objective.

    python tests/golden/make_bench_fixtures.py [--n-c3 64] [--n-c2 16] [--py 4]
"""
import argparse
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import bench_common as bc  # noqa: E402
import bench_nmpc as bn  # noqa: E402
from oracle import cpp_oracle as co, ipm, models as om, nlp as on  # noqa: E402
from oracle.track import OracleTrack  # noqa: E402


class _TrackView:
    """what bench_nmpc.make_batch needs of a track (obstacle tuples), served by the oracle's track"""
    def __init__(self, tr):
        self.tr = tr

    def obstacle_array(self):
        return np.array([o[:3] for o in self.tr.obstacles], dtype=float)


def build(workload, lap_path, B, seed, n, n_py):
    cfg = bn.controller_config() if workload == "c3" else bn.controller_config(horizon_pm=15, obstacles=False)
    c = on.MpcCfg.from_config(cfg)
    tr = OracleTrack(bc.load_yaml("environment/ippodromo.yaml"))
    lap = dict(np.load(lap_path))
    x0, Xw, Uw, obst, tick = bn.make_batch(lap, _TrackView(tr), B, seed, cfg)
    sel = np.linspace(0, B - 1, n).astype(int)
    x0, Xw, Uw, tick = x0[sel], Xw[sel], Uw[sel], tick[sel]
    obst = obst[sel] if obst is not None else np.zeros((n, 0, 3))
    ds, kap = np.empty((n, c.H)), np.empty((n, c.H))
    for i in range(n):
        ds[i], kap[i] = on.horizon_params(c, tr, x0[i], Xw[i, 0])
    t0 = time.time()
    r = co.solve(c, x0, kap, ds, Xw, Uw, obstacles=obst if c.obstacles else None, max_soc=4)
    print(f"{workload}: {n} instances solved by the C++ oracle in {time.time() - t0:.1f} s; status "
          f"{dict(zip(*np.unique(r['status'], return_counts=True)))} iters mean {r['iters'].mean():.1f} max {r['iters'].max()}")
    U_py = np.full((n,) + r["U"].shape[1:], np.nan)
    for i in np.linspace(0, n - 1, n_py).astype(int) if n_py else []:
        p = on.NLP(c, om.DEFAULT, x0[i], kap[i], ds[i], obst[i] if c.obstacles else None)
        t1 = time.time()
        rp = ipm.solve(p, p.pack(Xw[i], Uw[i]))
        _, U_py[i] = p.unpack(rp.w)
        e = np.abs(U_py[i] - r["U"][i]).max() / max(1.0, np.abs(r["U"][i]).max())
        print(f"   python oracle on #{i}: {rp.status} it {rp.iters} (C++ {r['iters'][i]}) obj {rp.obj:.10f} vs {r['obj'][i]:.10f} "
              f"max|dU| rel {e:.2e}  {time.time() - t1:.0f} s", flush=True)
    out = {"index": sel, "tick": tick, "x0": x0, "Xw": Xw, "Uw": Uw, "obst": obst, "ds": ds, "kappa": kap, "X": r["X"],
           "U": r["U"], "obj": r["obj"], "status": r["status"], "iters": r["iters"], "U_py": U_py,
           "batch": np.array(B), "seed": np.array(seed)}
    return {f"{workload}/{k}": np.asarray(v) for k, v in out.items()}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n-c3", type=int, default=64)
    ap.add_argument("--n-c2", type=int, default=16)
    ap.add_argument("--py", type=int, default=4)
    ap.add_argument("--laps", default=os.path.join(ROOT, "gpurun_out"))
    a = ap.parse_args()
    out = {}
    out.update(build("c3", os.path.join(a.laps, "c3_lap.npz"), 65536, bn.SEED, a.n_c3, a.py))
    out.update(build("c2", os.path.join(a.laps, "c2_lap.npz"), 4096, 1002, a.n_c2, min(a.py, 2)))
    path = os.path.join(HERE, "bench_oracle.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
