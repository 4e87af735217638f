import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from bench_common import golden_states, rollout_inputs, load_yaml
from vehicle_control_b200.engine import Engine
from vehicle_control_b200.environment.track import Track
tr = Track(load_yaml("environment/ippodromo.yaml"))
eng = Engine(load_yaml("models/dynamic_car.yaml"), load_yaml("controllers/cascaded.yaml"), tr)
B, N = 65536, 50
x0, U, kap, ds = (eng.tensor(a) for a in rollout_inputs(B, N, 1005, tr, golden_states()))
X = eng.rollout_st(x0, U, kap, ds)
X2 = eng.rollout_st(x0, U, kap, ds)
print("repeat equal", torch.equal(X, X2))
h = 25
Xa = eng.rollout_st(x0, U[:, :h].contiguous(), kap[:, :h].contiguous(), ds[:, :h].contiguous())
d = (Xa - X[:, :h + 1]).abs()
print("first half max diff", d.max().item(), "first differing step", (d.amax((0, 2)) > 0).nonzero()[:3].flatten().tolist(), "n traj differing", int((d.amax((1, 2)) > 0).sum()))
Xb = eng.rollout_st(Xa[:, -1].contiguous(), U[:, h:].contiguous(), kap[:, h:].contiguous(), ds[:, h:].contiguous())
d = (Xb - X[:, h:]).abs()
print("second half max diff", d.max().item(), "n traj differing", int((d.amax((1, 2)) > 0).sum()), "first differing step", (d.amax((0, 2)) > 0).nonzero()[:3].flatten().tolist())
b = int(d.amax((1, 2)).argmax())
print("worst traj", b, "x at h", X[b, h].tolist(), "\n", Xb[b, 0].tolist(), "\n step1", X[b, h + 1].tolist(), "\n", Xb[b, 1].tolist())
