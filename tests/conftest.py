import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden():
    """Reference closed-loop runs packed by tests/golden/make_golden.py."""
    z = np.load(os.path.join(ROOT, "tests", "golden", "runs.npz"))
    with open(os.path.join(ROOT, "tests", "golden", "runs.json")) as f:
        meta = json.load(f)
    return z, meta


def load_yaml(rel):
    import yaml
    with open(os.path.join(ROOT, "vehicle_control_b200", "config", rel)) as f:
        return yaml.safe_load(f)


@pytest.fixture(scope="session")
def oracle_tracks():
    from oracle.track import OracleTrack
    return {n: OracleTrack(load_yaml(f"environment/{n}.yaml")) for n in ("ippodromo", "shoe")}
