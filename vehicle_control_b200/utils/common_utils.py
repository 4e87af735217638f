"""Host glue: YAML loading with attribute access, angle wrapping, project root.

Mirrors vehicle_control/utils/common_utils.py:16-47 (load_config, wrap, project_root).  The reference
wraps configs in OmegaConf (scripts/main.py:6-10); omegaconf is optional here -- ``AttrDict`` gives the
same attribute / item access on plain dicts.
"""
from __future__ import annotations

import copy
import os

import numpy as np
import yaml


class AttrDict(dict):
    """dict with attribute access, recursive; stand-in for OmegaConf.create()."""

    def __init__(self, data=None):
        super().__init__()
        for k, v in (data or {}).items():
            self[k] = _wrap_value(v)

    def __getattr__(self, key):
        try:
            return self[key]
        except KeyError as e:
            raise AttributeError(key) from e

    def __setattr__(self, key, value):
        self[key] = _wrap_value(value)

    def __deepcopy__(self, memo):
        return AttrDict(copy.deepcopy(dict(self), memo))

    def to_dict(self):
        return {k: (v.to_dict() if isinstance(v, AttrDict) else copy.deepcopy(v)) for k, v in self.items()}


def _wrap_value(v):
    if isinstance(v, dict) and not isinstance(v, AttrDict):
        return AttrDict(v)
    return v


def as_config(obj):
    """Accept a dict, AttrDict or OmegaConf node and return something with attribute access."""
    if isinstance(obj, AttrDict):
        return obj
    if isinstance(obj, dict):
        return AttrDict(obj)
    try:  # OmegaConf node
        from omegaconf import OmegaConf
        return AttrDict(OmegaConf.to_container(obj, resolve=True))
    except Exception:
        return obj


def load_config(file_path):
    with open(file_path, "r") as f:
        return AttrDict(yaml.safe_load(f))


def config_dir() -> str:
    return os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "config")


def project_root() -> str:
    return os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def wrap(angle):
    """Wrap into [-pi, pi] by one turn (scalar or array)."""
    a = np.asarray(angle, dtype=float)
    out = np.where(a < -np.pi, 2 * np.pi + a, np.where(a > np.pi, a - 2 * np.pi, a))
    return out if out.ndim else float(out)
