"""ctypes binding of libvc_b200.so (include/vc_api.h).  There is no CPU fallback: if the shared
library is missing it is built with nvcc; if that fails, or no CUDA device exists when a handle is
created, an exception is raised."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VC_LIB_PATH") or os.path.join(_HERE, "libvc_b200.so")   # VC_LIB_PATH: tuning builds

D = C.c_double
PD = C.c_void_p  # device or host pointers are passed as raw addresses


class ModelParams(C.Structure):
    """vc_model_params"""
    _fields_ = [(n, D) for n in (
        "dt", "l", "m", "Izz", "a", "b", "h", "eps", "Peng", "Xdf", "Xdr", "Xbf", "Xbr", "Caf", "Car",
        "Cd", "muf", "mur", "Av2", "Frr", "g_eff")]


class MpcConfig(C.Structure):
    """vc_mpc_config"""
    _fields_ = ([("N", C.c_int), ("M", C.c_int), ("obstacles", C.c_int), ("max_iter", C.c_int),
                 ("mpc_dt", D), ("ds_pm", D)]
                + [(n, D) for n in ("w_time", "w_speed", "w_ey", "w_epsi", "w_w", "w_Fx", "w_Fy", "w_dev_st",
                                    "w_dev_pm", "w_boundary", "w_slip", "w_obstacles", "w_switch_F",
                                    "w_min", "w_max", "Ux_min", "max_speed", "delta_min", "delta_max",
                                    "ey_min", "ey_max", "V_min", "ey_min_pm", "ey_max_pm",
                                    "tol", "acceptable_tol")]
                + [("acceptable_iter", C.c_int), ("obj_scaling", C.c_int), ("mu_init", D), ("bound_relax", D),
                   ("warm_rungs", C.c_int), ("max_soc", C.c_int), ("watchdog_trigger", C.c_int), ("watchdog_trials", C.c_int),
                   ("max_damp", C.c_int), ("wide_max", C.c_int)])


EXPORTS = {
    # name: (restype, argtypes)
    "vc_last_error": (C.c_char_p, []),
    "vc_version": (C.c_int, []),
    "vc_create": (C.c_int, [C.POINTER(ModelParams), C.POINTER(MpcConfig), C.c_int, C.c_void_p, C.POINTER(C.c_void_p)]),
    "vc_destroy": (C.c_int, [C.c_void_p]),
    "vc_set_config": (C.c_int, [C.c_void_p, C.POINTER(MpcConfig)]),
    "vc_default_config": (C.c_int, [C.POINTER(MpcConfig)]),
    "vc_set_track": (C.c_int, [C.c_void_p, PD, C.c_int, D, D, PD, PD, C.c_int, D]),
    "vc_track_curvature": (C.c_int, [C.c_void_p, C.c_int, PD, PD]),
    "vc_rel2glob": (C.c_int, [C.c_void_p, C.c_int, PD, PD, PD, PD]),
    "vc_horizon_params": (C.c_int, [C.c_void_p, C.c_int, PD, PD, PD, PD]),
    "vc_rollout_st": (C.c_int, [C.c_void_p, C.c_int, C.c_int, PD, PD, PD, PD, PD, PD]),
    "vc_rollout_pm": (C.c_int, [C.c_void_p, C.c_int, C.c_int, PD, PD, PD, PD, PD, PD]),
    "vc_plant_step": (C.c_int, [C.c_void_p, C.c_int, PD, PD, D]),
    "vc_set_track_atlas": (C.c_int, [C.c_void_p, C.c_int, PD, PD, PD, PD]),
    "vc_horizon_params_tracks": (C.c_int, [C.c_void_p, C.c_int, PD, PD, PD, PD, PD]),
    "vc_plant_step_tracks": (C.c_int, [C.c_void_p, C.c_int, PD, PD, PD, D]),
    "vc_build_tracks": (C.c_int, [C.c_void_p, C.c_int, C.c_int, PD, PD, PD, PD]),
    "vc_built_track_info": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(D), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "vc_built_track_read": (C.c_int, [C.c_void_p, C.c_int, C.c_int, PD, C.c_int]),
    "vc_built_track_rel2glob": (C.c_int, [C.c_void_p, C.c_int, C.c_int, PD, PD, PD]),
    "vc_tyre_forces": (C.c_int, [C.c_void_p, C.c_int, PD, PD, PD]),
    "vc_nmpc_solve": (C.c_int, [C.c_void_p, C.c_int, PD, PD, PD, PD, C.c_int, PD, PD, PD, PD, PD, PD]),
    "vc_nmpc_solve_static": (C.c_int, [C.c_void_p, C.c_int, PD, PD, PD, PD, C.c_int, PD, PD, PD, PD, PD, PD]),
    "vc_nmpc_eval": (C.c_int, [C.c_void_p, C.c_int, PD, PD, PD, PD, C.c_int, PD, PD, PD]),
    "vc_nmpc_solve_host": (C.c_int, [C.c_void_p, C.c_int, PD, PD, PD, PD, C.c_int, PD, PD, PD, PD, PD, PD]),
    "vc_set_blocking_sync": (C.c_int, [C.c_void_p, C.c_int]),
    "vc_nmpc_command_host": (C.c_int, [C.c_void_p, C.c_int, PD, PD, C.c_int, PD, PD, PD, PD, PD, PD]),
    "vc_set_profiling": (C.c_int, [C.c_void_p, C.c_int]),
    "vc_get_profile": (C.c_int, [C.c_void_p, PD, PD, PD]),
    "vc_get_profile_instance_passes": (C.c_longlong, [C.c_void_p]),
    "vc_nmpc_debug_read": (C.c_int, [C.c_void_p, C.c_int, C.c_int, PD, C.c_int]),
    "vc_workspace_bytes": (C.c_longlong, [C.c_void_p]),
    "vc_launch_count": (C.c_longlong, [C.c_void_p]),
}

_lib = None


def load(build_if_missing: bool = True) -> C.CDLL:
    """Load (building first if needed) the CUDA library; raises if that is impossible."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        if not build_if_missing:
            raise RuntimeError(f"{LIB_PATH} is missing and there is no CPU fallback; run vehicle_control_b200/build.py")
        from .build import build
        build()
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in EXPORTS.items():
        fn = getattr(lib, name)   # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class VcError(RuntimeError):
    pass


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = load().vc_last_error()
        raise VcError(f"{what} failed ({rc}): {msg.decode() if msg else ''}")
