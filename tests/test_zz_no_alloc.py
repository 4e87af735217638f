"""GPU: vc_nmpc_solve_static -- the allocation-free variant of vc_nmpc_solve (include/vc_api.h; no counterpart in the
reference).  Called through the C-ABI, twice in a row, its outputs must
equal vc_nmpc_solve's on the same instances bit for bit (the same kernels run; only the scratch memory differs) and agree
with the oracle's answers within the 1e-6 bar of the other parity tests.  Collected last on purpose (file name)."""
import ctypes as C
import os

import numpy as np
import pytest

from tests.conftest import ROOT
from tests.test_gpu_bench_parity import TOL, _engine, _fx, _rel

pytestmark = pytest.mark.gpu


def _call_static(eng, x0, kap, ds, ob, X, U, obj, status, iters, kkt):
    """the raw C-ABI call: plain device pointers and sizes"""
    P = lambda t: C.c_void_p(t.data_ptr())
    B, n_obs = x0.shape[0], ob.shape[1]
    return eng.lib.vc_nmpc_solve_static(eng._h, B, P(x0), P(kap), P(ds), P(ob), n_obs, P(X), P(U), P(obj), P(status),
                                        P(iters), P(kkt))


def test_static_scratch_solve_equals_allocating_solve_twice_in_a_row():
    import torch
    f = _fx("c3")
    eng, cfg = _engine("c3")
    n = 48                                              # of the 63 instances the 16 MiB array holds at H = 55
    x0, ob = eng.tensor(f["x0"][:n]), eng.tensor(f["obst"][:n])
    Xr, Ur = eng.tensor(f["Xw"][:n]), eng.tensor(f["Uw"][:n])
    ds, kap = eng.horizon_params(x0, Xr)
    ref = eng.solve(x0, kap, ds, Xr, Ur, obstacles=ob)   # the existing entry point (cudaMalloc'ed workspace)
    assert (ref["status"].cpu().numpy() == 0).all()
    # every buffer of the static call exists before it; the call itself must not allocate
    X, U = eng.tensor(f["Xw"][:n]), eng.tensor(f["Uw"][:n])
    obj, kkt = eng.empty(n), eng.empty(n)
    status, iters = eng.empty(n, dtype=torch.int32), eng.empty(n, dtype=torch.int32)
    X2, U2 = eng.tensor(f["Xw"][n - 8:n]), eng.tensor(f["Uw"][n - 8:n])
    torch.cuda.synchronize()
    held = torch.cuda.memory_allocated()
    ws_before = eng.workspace_bytes()

    # call 1
    assert _call_static(eng, x0, kap, ds, ob, X, U, obj, status, iters, kkt) == 0
    torch.cuda.synchronize()
    assert torch.equal(U, Ur) and torch.equal(X, Xr)
    assert torch.equal(obj, ref["obj"]) and torch.equal(iters, ref["iters"]) and torch.equal(status, ref["status"])
    assert torch.equal(kkt, ref["kkt"])
    assert _rel(U.cpu().numpy(), f["U"][:n]).max() < TOL and _rel(obj.cpu().numpy(), f["obj"][:n]).max() < TOL

    # call 2, right behind it, on a different (smaller) batch: the array still holds call 1's records
    sl = slice(n - 8, n)
    assert _call_static(eng, x0[sl], kap[sl], ds[sl], ob[sl], X2, U2, obj[:8], status[:8], iters[:8], kkt[:8]) == 0
    torch.cuda.synchronize()
    assert torch.equal(U2, Ur[sl]) and torch.equal(X2, Xr[sl])
    assert torch.equal(obj[:8], ref["obj"][sl]) and torch.equal(iters[:8], ref["iters"][sl])
    assert (status[:8].cpu().numpy() == 0).all()
    assert _rel(U2.cpu().numpy(), f["U"][sl]).max() < TOL

    # nothing was allocated on the way: torch's pool and the handle's workspace are what they were
    assert torch.cuda.memory_allocated() == held
    assert eng.workspace_bytes() == ws_before


def test_static_scratch_rejects_batches_beyond_the_array():
    import torch
    f = _fx("c3")
    eng, cfg = _engine("c3")
    n = 64                                              # 64 x 263,144 B > 16 MiB at H = 55
    assert n * ((596 * eng.H + 112) * 8 + 8) > (16 << 20)
    x0, ob = eng.tensor(f["x0"][:n]), eng.tensor(f["obst"][:n])
    X, U = eng.tensor(f["Xw"][:n]), eng.tensor(f["Uw"][:n])
    U_in = U.clone()
    ds, kap = eng.horizon_params(x0, X)
    obj, kkt = eng.empty(n), eng.empty(n)
    status, iters = eng.empty(n, dtype=torch.int32), eng.empty(n, dtype=torch.int32)
    rc = _call_static(eng, x0, kap, ds, ob, X, U, obj, status, iters, kkt)
    assert rc != 0
    eng.lib.vc_last_error.restype = C.c_char_p
    msg = eng.lib.vc_last_error().decode()
    assert "vc_nmpc_solve_static" in msg and "at most 63 instances" in msg, msg
    torch.cuda.synchronize()
    assert torch.equal(U, U_in)                          # nothing was launched
    with pytest.raises(RuntimeError):                    # the host mirror reports it the same way
        eng.solve_static(x0, kap, ds, X, U, {"obj": obj, "status": status, "iters": iters, "kkt": kkt}, obstacles=ob)


def test_header_declares_the_limit():
    src = open(os.path.join(ROOT, "include", "vc_api.h")).read()
    assert "#define VC_STATIC_SCRATCH_BYTES (16u << 20)" in src and "vc_nmpc_solve_static" in src
