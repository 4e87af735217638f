"""A few C3 bench instances solved for a handful of iterations by the block-per-instance (late pass) kernels: the
smallest program that exercises k_step_a_wide / k_step_b_wide (profiling target for the latency of one iteration)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests.test_gpu_bench_parity import _fx, _engine
f = _fx('c3')
wide = int(sys.argv[1]) if len(sys.argv) > 1 else (1 << 20)
eng, cfg = _engine('c3'); eng.set_config(cfg, wide_max=wide, max_iter=int(sys.argv[2]) if len(sys.argv) > 2 else 8)
n = 8
x0, X, U = eng.tensor(f['x0'][:n]), eng.tensor(f['Xw'][:n]), eng.tensor(f['Uw'][:n])
ds, kap = eng.horizon_params(x0, X)
out = eng.solve(x0, kap, ds, X, U, obstacles=eng.tensor(f['obst'][:n]))
print("iters", out["iters"].cpu().numpy(), "launches", eng.launch_count())
