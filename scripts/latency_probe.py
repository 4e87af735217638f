"""Latency of ONE NMPC iteration for a lone instance (the tail of a batch, and a single vehicle's tick): a few C3 bench
instances solved alone with profiling on.  VC_PASS_TRACE=1 prints the per-pass kernel times to stderr.
usage: latency_probe.py [n_instances] [max_iter] [wide_max]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from tests.test_gpu_bench_parity import _engine, _fx

f = _fx("c3")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1
mi = int(sys.argv[2]) if len(sys.argv) > 2 else 300
wide = int(sys.argv[3]) if len(sys.argv) > 3 else 296
eng, cfg = _engine("c3")
eng.set_config(cfg, wide_max=wide, max_iter=mi)
order = np.argsort(-f["iters"])[:n]          # the instances with the most iterations first
x0 = eng.tensor(f["x0"][order])
ob = eng.tensor(f["obst"][order])
for rep in range(3):
    X, U = eng.tensor(f["Xw"][order]), eng.tensor(f["Uw"][order])
    ds, kap = eng.horizon_params(x0, X)
    eng.set_profiling(rep == 2)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = eng.solve(x0, kap, ds, X, U, obstacles=ob)
    torch.cuda.synchronize()
    el = time.perf_counter() - t0
p = eng.profile()
it = out["iters"].cpu().numpy()
print(f"n={n} wide_max={wide}: iters {it.tolist()} passes {p['last_passes']} wall {el*1e3:.2f} ms "
      f"-> {el*1e3/p['last_passes']:.3f} ms/pass; eval_st {p['ms'][0]/p['last_passes']*1e3:.0f} us, "
      f"eval_pm {p['ms'][1]/p['last_passes']*1e3:.0f} us, step(a+b+update) {p['ms'][2]/p['last_passes']*1e3:.0f} us per pass")
