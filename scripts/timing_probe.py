import os, sys
sys.path.insert(0, '/root/repo')
import numpy as np
from tests.test_gpu_bench_parity import _fx, _engine
f=_fx('c3')
for wide_max in (0, 1<<20):
    eng,cfg=_engine('c3'); eng.set_config(cfg, wide_max=wide_max, max_iter=6)
    n=4
    x0,X,U=eng.tensor(f['x0'][:n]),eng.tensor(f['Xw'][:n]),eng.tensor(f['Uw'][:n])
    ds,kap=eng.horizon_params(x0,X)
    out=eng.solve(x0,kap,ds,X,U,obstacles=eng.tensor(f['obst'][:n]))
    sc=eng.debug_read(0,-1,96)
    t=sc[84:90]
    print('wide_max',wide_max,'ticks: finish_grad %d  prelude(compl/mu) %d  riccati_bwd %d  riccati_fwd %d  slackdir %d  total %d cycles' % (t[1]-t[0], t[2]-t[1], t[3]-t[2], t[4]-t[3], t[5]-t[4], t[5]-t[0]))
