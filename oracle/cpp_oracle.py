"""ORACLE (test infrastructure only -- never imported by the product path).

ctypes binding of oracle/_ref/libvc_oracle.so, the C++ fp64 restatement of the reference's NMPC path for the CPU
(oracle/cpp/vc_oracle.cpp: exact derivatives by second-order jets, IPOPT's published algorithm, block-tridiagonal
LDL^T; OpenMP over instances).  Built by oracle/cpp/Makefile (`build()` below, also run by __graft_entry__.build()).

The parameter structs are the public ones of include/vc_api.h; their ctypes mirrors are generated from the header
text so that they cannot drift.  Array conventions are those of vc_nmpc_solve, with host (numpy) arrays.
"""
from __future__ import annotations

import ctypes as C
import os
import re
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIB_PATH = os.path.join(HERE, "_ref", "libvc_oracle.so")
STATUS_NAMES = {0: "solved", 1: "acceptable", 2: "max_iter", 3: "eval_error", 4: "linesearch_fail"}


def _struct_from_header(name: str):
    text = open(os.path.join(ROOT, "include", "vc_api.h")).read()
    body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (name, name), text, re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    fields = []
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl:
            continue
        ty, names = decl.split(None, 1)
        ct = {"double": C.c_double, "int": C.c_int}[ty]
        fields += [(n.strip(), ct) for n in names.split(",")]
    return type(name, (C.Structure,), {"_fields_": fields})


ModelParams = _struct_from_header("vc_model_params")
MpcConfig = _struct_from_header("vc_mpc_config")


class Options(C.Structure):
    _fields_ = [("max_iter", C.c_int), ("max_soc", C.c_int), ("max_restorations", C.c_int), ("verbose", C.c_int),
                ("obj_scale", C.c_double), ("watchdog", C.c_int)]


def build(force: bool = False) -> str:
    src = os.path.join(HERE, "cpp")
    deps = [os.path.join(src, f) for f in ("vc_oracle.cpp", "model.hpp", "jet.hpp", "Makefile")]
    deps.append(os.path.join(ROOT, "include", "vc_api.h"))
    stale = not os.path.exists(LIB_PATH) or any(os.path.getmtime(d) > os.path.getmtime(LIB_PATH) for d in deps)
    if force or stale:
        r = subprocess.run(["make", "-C", src] + (["-B"] if force else []), capture_output=True, text=True)
        if r.returncode:
            raise RuntimeError("oracle/cpp build failed:\n" + r.stdout + r.stderr)
    return LIB_PATH


_lib = None


def load() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        lib = C.CDLL(LIB_PATH)
        P = C.c_void_p
        lib.vco_nmpc_solve.restype = C.c_int
        lib.vco_nmpc_solve.argtypes = [C.POINTER(ModelParams), C.POINTER(MpcConfig), C.POINTER(Options), C.c_int, P, P, P, P,
                                       C.c_int, P, P, P, P, P, P, C.c_int]
        lib.vco_nmpc_eval.restype = C.c_int
        lib.vco_nmpc_eval.argtypes = [C.POINTER(ModelParams), C.POINTER(MpcConfig), C.c_int, P, P, P, P, C.c_int, P, P, P]
        lib.vco_rollout_st.restype = C.c_int
        lib.vco_rollout_st.argtypes = [C.POINTER(ModelParams), C.c_int, C.c_int, P, P, P, P, P, P, C.c_int]
        lib.vco_plant_step.restype = C.c_int
        lib.vco_plant_step.argtypes = [C.POINTER(ModelParams), C.c_int, P, P, P, C.c_double]
        _lib = lib
    return _lib


def model_params(car=None) -> ModelParams:
    """oracle.models.CarParams (config/models/dynamic_car.yaml) -> vc_model_params."""
    import math
    from . import models as om
    car = car or om.DEFAULT
    p = ModelParams()
    for k in ("dt", "l", "m", "Izz", "a", "b", "h", "eps", "Peng", "Xdf", "Xdr", "Xbf", "Xbr", "Caf", "Car", "Cd", "muf",
              "mur", "Av2", "Frr"):
        setattr(p, k, float(getattr(car, k)))
    p.g_eff = om.G_GRAV * math.cos(car.theta) * math.cos(car.phi)
    return p


def mpc_config(cfg, tol=1e-8) -> MpcConfig:
    """oracle.nlp.MpcCfg (or a controller yaml dict) -> vc_mpc_config with the IPOPT defaults the reference runs with."""
    from . import nlp as on
    if not isinstance(cfg, on.MpcCfg):
        cfg = on.MpcCfg.from_config(cfg)
    c = MpcConfig()
    for k in ("N", "M", "mpc_dt", "ds_pm", "w_time", "w_speed", "w_ey", "w_epsi", "w_w", "w_Fx", "w_Fy", "w_dev_st", "w_dev_pm",
              "w_boundary", "w_slip", "w_obstacles", "w_switch_F", "w_min", "w_max", "Ux_min", "max_speed", "delta_min",
              "delta_max", "ey_min", "ey_max", "V_min", "ey_min_pm", "ey_max_pm"):
        setattr(c, k, getattr(cfg, k))
    c.obstacles = 1 if cfg.obstacles else 0
    c.max_iter = 3000
    c.tol, c.acceptable_tol, c.acceptable_iter = tol, 1e-6, 15
    c.obj_scaling, c.mu_init, c.bound_relax = 1, 0.1, 1e-8
    return c


def _a(x, dtype=np.float64):
    return np.ascontiguousarray(x, dtype=dtype)


def solve(cfg, x0, kappa, ds, X, U, obstacles=None, car=None, max_iter=3000, max_soc=4, max_restorations=2, verbose=0,
          threads=0, obj_scale=0.0, watchdog=10):
    """Batched solve on the host cores.  x0[B,8], kappa[B,H], ds[B,H], X[B,8,H], U[B,2,H] (warm start), obstacles[B,n,3].
    Returns dict(X, U, obj, status, iters, kkt) of new arrays."""
    lib = load()
    c = cfg if isinstance(cfg, MpcConfig) else mpc_config(cfg)
    p = car if isinstance(car, ModelParams) else model_params(car)
    x0, kappa, ds = _a(x0), _a(kappa), _a(ds)
    single = x0.ndim == 1
    if single:
        x0, kappa, ds, X, U = x0[None], kappa[None], ds[None], np.asarray(X)[None], np.asarray(U)[None]
        obstacles = None if obstacles is None else np.asarray(obstacles)[None]
    B, H = x0.shape[0], c.N + c.M
    X, U = _a(X).copy(), _a(U).copy()
    assert X.shape == (B, 8, H) and U.shape == (B, 2, H) and kappa.shape == (B, H) and ds.shape == (B, H)
    n_obs = 0
    ob = None
    if obstacles is not None and c.obstacles and np.size(obstacles):
        ob = _a(obstacles)
        n_obs = ob.shape[1]
        assert ob.shape == (B, n_obs, 3)
    obj, kkt = np.empty(B), np.empty(B)
    status, iters = np.empty(B, np.int32), np.empty(B, np.int32)
    o = Options(int(max_iter), int(max_soc), int(max_restorations), int(verbose), float(obj_scale), int(watchdog))
    rc = lib.vco_nmpc_solve(C.byref(p), C.byref(c), C.byref(o), B, x0.ctypes.data, kappa.ctypes.data, ds.ctypes.data,
                            ob.ctypes.data if ob is not None else None, n_obs, X.ctypes.data, U.ctypes.data, obj.ctypes.data,
                            status.ctypes.data, iters.ctypes.data, kkt.ctypes.data, int(threads))
    if rc:
        raise RuntimeError(f"vco_nmpc_solve failed ({rc})")
    out = {"X": X, "U": U, "obj": obj, "status": status, "iters": iters, "kkt": kkt}
    return {k: v[0] for k, v in out.items()} if single else out


def evaluate(cfg, x0, kappa, ds, X, U, obstacles=None, car=None):
    """[B,3]: objective, max equality violation, min inequality margin of given trajectories."""
    lib = load()
    c = cfg if isinstance(cfg, MpcConfig) else mpc_config(cfg)
    p = car if isinstance(car, ModelParams) else model_params(car)
    x0, kappa, ds, X, U = _a(x0), _a(kappa), _a(ds), _a(X), _a(U)
    B = x0.shape[0]
    n_obs, ob = 0, None
    if obstacles is not None and c.obstacles and np.size(obstacles):
        ob = _a(obstacles)
        n_obs = ob.shape[1]
    out = np.empty((B, 3))
    lib.vco_nmpc_eval(C.byref(p), C.byref(c), B, x0.ctypes.data, kappa.ctypes.data, ds.ctypes.data,
                      ob.ctypes.data if ob is not None else None, n_obs, X.ctypes.data, U.ctypes.data, out.ctypes.data)
    return out


def rollout_st(x0, U, kappa, ds, jac=False, car=None, threads=0):
    lib = load()
    p = car if isinstance(car, ModelParams) else model_params(car)
    x0, U, kappa, ds = _a(x0), _a(U), _a(kappa), _a(ds)
    B, N = U.shape[0], U.shape[1]
    X = np.empty((B, N + 1, 8))
    AB = np.empty((B, N, 8, 10)) if jac else None
    lib.vco_rollout_st(C.byref(p), B, N, x0.ctypes.data, U.ctypes.data, kappa.ctypes.data, ds.ctypes.data, X.ctypes.data,
                       AB.ctypes.data if jac else None, int(threads))
    return (X, AB) if jac else X


def plant_step(x, u, kappa, dt, car=None):
    lib = load()
    p = car if isinstance(car, ModelParams) else model_params(car)
    x = _a(x).copy()
    u, kappa = _a(u), _a(kappa)
    lib.vco_plant_step(C.byref(p), x.shape[0], x.ctypes.data, u.ctypes.data, kappa.ctypes.data, float(dt))
    return x


def num_threads() -> int:
    return int(load().vco_num_threads())
