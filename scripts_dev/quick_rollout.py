import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from bench_common import *
from vehicle_control_b200.engine import Engine
from vehicle_control_b200.environment.track import Track
tr = Track(load_yaml("environment/ippodromo.yaml"))
eng = Engine(load_yaml("models/dynamic_car.yaml"), load_yaml("controllers/cascaded.yaml"), tr)
B, N = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 18, 50
x0, U, kap, ds = (eng.tensor(a) for a in rollout_inputs(B, N, 1005, tr, golden_states()))
X = eng.empty(B, N + 1, 8); AB = eng.empty(B, N, 8, 10)
for jac in (False, True):
    for _ in range(3): eng.rollout_st_into(x0, U, kap, ds, X, AB if jac else None)
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(5): eng.rollout_st_into(x0, U, kap, ds, X, AB if jac else None)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(f"jac={jac} B={B} N={N}: {ms:.3f} ms  {B*N/ms*1e3/1e9:.3f} G state-steps/s  finite={bool(torch.isfinite(X).all())}"
          + (f" {B*N*736/ms*1e3/1e9:.1f} GB/s" if jac else ""))
