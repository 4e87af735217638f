"""Traces ONE instance of a dumped sample (gpurun_out/c3_sample.npz from scripts/dump_c3_sample.py) through the engine
with VC_DEBUG=1 (per-pass trace of instance 0 on stderr)."""
import os, sys
os.environ["VC_DEBUG"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tests.test_gpu_bench_parity import _engine
z = np.load(sys.argv[1])
i = int(sys.argv[2])
eng, cfg = _engine('c3')
eng.set_config(cfg, max_iter=int(sys.argv[3]) if len(sys.argv) > 3 else 3000)
one = lambda k: eng.tensor(np.ascontiguousarray(z[k][i:i + 1]))
x0, X, U = one('x0'), one('Xw'), one('Uw')
ds, kap = eng.horizon_params(x0, X)
out = eng.solve(x0, kap, ds, X, U, obstacles=one('obst'))
print("status", out["status"].cpu().numpy(), "iters", out["iters"].cpu().numpy(), "obj", out["obj"].cpu().numpy())
