"""CPU: the C-ABI library builds, loads and exports every symbol include/vc_api.h declares."""
import os
import re

import pytest

from tests.conftest import ROOT


def _declared():
    src = open(os.path.join(ROOT, "include", "vc_api.h")).read()
    return sorted(set(re.findall(r"\b(vc_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from vehicle_control_b200 import _capi
    lib = _capi.load()
    names = _declared()
    assert "vc_nmpc_solve" in names and "vc_rollout_st" in names
    for n in names:
        assert hasattr(lib, n), n
    assert set(names) == set(_capi.EXPORTS), set(names) ^ set(_capi.EXPORTS)
    assert lib.vc_version() >= 100


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from vehicle_control_b200.engine import Engine
    from tests.conftest import load_yaml
    with pytest.raises(RuntimeError):
        Engine(load_yaml("models/dynamic_car.yaml"))


def test_product_does_not_import_oracle():
    bad = []
    for d, _, fs in os.walk(os.path.join(ROOT, "vehicle_control_b200")):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(d, f)).read()
                if re.search(r"^\s*(from|import)\s+oracle\b", txt, re.M):
                    bad.append(f)
    assert not bad, bad


def test_handwritten_adjoints_match_hyperdual_templates(tmp_path):
    """vc_adj.cuh (forward-over-reverse stage Lagrangian, what k_eval_st runs) compiled for the host against the
    hyper-dual templates of vc_model.cuh / vc_nlp.cuh: values exact, gradient and Hessian columns to 1e-10."""
    import shutil
    import subprocess
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else shutil.which("g++")
    exe = str(tmp_path / "adj_check")
    r = subprocess.run([cxx, "-O1", "-std=c++17", "-x", "c++", os.path.join(ROOT, "tests", "host", "adj_check.cpp"), "-o", exe],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-2000:]
    assert "cases 600" in r.stdout


def test_handderived_rk4_jacobian_matches_dual_numbers(tmp_path):
    """vc_jac.cuh (hand-derived Jacobian of the spatial SingleTrack right-hand side chained through RK4, what
    k_rollout_st_jac_an runs) compiled for the host against the forward-mode dual numbers of vc_ad.cuh / vc_model.cuh
    (dynamic_car.py:78-191, integrators.py:26-37): values to 1e-13, all 80 entries of [A|B] to 1e-11, both Fiala branches."""
    import shutil
    import subprocess
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else shutil.which("g++")
    exe = str(tmp_path / "jac_check")
    r = subprocess.run([cxx, "-O1", "-std=c++17", "-x", "c++", os.path.join(ROOT, "tests", "host", "jac_check.cpp"), "-o", exe],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-2000:]
    assert "cases 2000" in r.stdout
