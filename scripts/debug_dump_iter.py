import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tests.test_gpu_bench_parity import _engine
z = np.load(sys.argv[1]); i = int(sys.argv[2]); mi = int(sys.argv[3])
eng, cfg = _engine('c3'); eng.set_config(cfg, max_iter=mi)
one = lambda k: eng.tensor(np.ascontiguousarray(z[k][i:i + 1]))
x0, X, U = one('x0'), one('Xw'), one('Uw')
ds, kap = eng.horizon_params(x0, X)
out = eng.solve(x0, kap, ds, X, U, obstacles=one('obst'))
print("status", out["status"].cpu().numpy(), "iters", out["iters"].cpu().numpy(), "obj", out["obj"].cpu().numpy())
np.savez(sys.argv[4], X=X.cpu().numpy(), U=U.cpu().numpy(), kappa=kap.cpu().numpy(), ds=ds.cpu().numpy())
