"""Times the evaluation kernels at full load: the 64 C3 fixture instances tiled to a large batch, a few passes with
VC_PASS_TRACE (per-kernel CUDA-event times on stderr).  Used to compare build variants (VC_LIB_PATH)."""
import os, sys
os.environ["VC_PASS_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tests.test_gpu_bench_parity import _fx, _engine
f = _fx('c3')
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
rep = lambda a: np.ascontiguousarray(np.tile(a, (B // len(a),) + (1,) * (a.ndim - 1)))
eng, cfg = _engine('c3'); eng.set_config(cfg, max_iter=6)
x0, X, U = eng.tensor(rep(f['x0'])), eng.tensor(rep(f['Xw'])), eng.tensor(rep(f['Uw']))
ds, kap = eng.horizon_params(x0, X)
eng.set_profiling(True)
eng.solve(x0, kap, ds, X, U, obstacles=eng.tensor(rep(f['obst'])))
