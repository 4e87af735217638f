"""Named state / action vectors (mirror of vehicle_control/utils/fancy_vector.py:7-72).

The reference's vectors also carry CasADi symbols; here ``syms`` are just the key names, because
derivatives are produced on the device (csrc/vc_ad.cuh), not by a symbolic graph.
"""
from __future__ import annotations

from typing import Union

import numpy as np


class FancyVector:
    _keys: list = []
    _labels: list = []

    def __init__(self, *args, **kwargs):
        vals = [0.0] * len(self._keys)
        for i, a in enumerate(args):
            vals[i] = a
        for k, v in kwargs.items():
            vals[self._keys.index(k)] = v
        self._values = np.array(vals, dtype=float)

    @classmethod
    def create(cls, *args, **kwargs):
        return cls(*args, **kwargs)

    @property
    def values(self) -> np.ndarray:
        return self._values

    @property
    def keys(self) -> list:
        return self._keys

    @property
    def syms(self):
        return list(self._keys)

    @property
    def labels(self):
        return self._labels

    @property
    def variables(self):
        return [self.syms[i] for i in range(len(self.values))]

    def index(self, key):
        return self.keys.index(key)

    def __getitem__(self, key: Union[int, str]):
        if isinstance(key, (int, np.integer)):
            return self.values[key]
        return self.values[self.keys.index(key)]

    def __setitem__(self, key: Union[int, str], value):
        if isinstance(key, (int, np.integer)):
            self.values[key] = value
        else:
            self.values[self.keys.index(key)] = value

    def __getattr__(self, name):
        keys = type(self)._keys
        if name in keys:
            return self._values[keys.index(name)]
        raise AttributeError(name)

    def __setattr__(self, name, value):
        if name in type(self)._keys:
            self._values[type(self)._keys.index(name)] = value
        else:
            object.__setattr__(self, name, value)

    def __iter__(self):
        return iter(self.values)

    def __len__(self):
        return len(self.values)

    def __str__(self):
        return str({k: f"{v:.2f}" for k, v in zip(self.keys, self.values)})

    def __add__(self, other):
        assert isinstance(other, self.__class__), "You can only sum two same states"
        return self.__class__.create(*(self.values + other.values))
