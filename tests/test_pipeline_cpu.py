"""Host logic of vehicle_control_b200/pipeline.py (several batches in flight): round-robin step -> lane assignment, per-lane
order, the global order of `ordered` sections, first/last hooks and error propagation -- with stream-less lanes (no GPU)."""
import threading
import time

import pytest

from vehicle_control_b200.pipeline import InFlight, Lane


def _fl(n):
    return InFlight([Lane(i, engine=None, stream=None) for i in range(n)])


def test_round_robin_and_ordered_sections():
    fl = _fl(3)
    seen, ordered, lock = [], [], threading.Lock()

    def step(lane, i):
        time.sleep(0.002 * ((7 * i) % 5))          # steps finish out of order
        with lock:
            seen.append((lane.index, i, threading.current_thread().name))
        with fl.ordered(i):
            ordered.append(i)

    hooks = []
    fl.run(step, 11, first=lambda ln: hooks.append(("first", ln.index)), last=lambda ln: hooks.append(("last", ln.index)))
    assert sorted(i for _, i, _ in seen) == list(range(11))
    assert all(lane == i % 3 for lane, i, _ in seen)                       # step i on lane i mod L
    for ln in range(3):                                                     # a lane executes its steps in order, on one thread
        mine = [(i, t) for lane, i, t in seen if lane == ln]
        assert [i for i, _ in mine] == sorted(i for i, _ in mine) and len({t for _, t in mine}) == 1
    assert ordered == list(range(11))                                       # ordered sections follow the step order
    assert sorted(hooks) == sorted([("first", i) for i in range(3)] + [("last", i) for i in range(3)])
    fl.run(step, 2)                                                         # fewer steps than lanes, and a second run
    assert ordered[-2:] == [0, 1]


def test_lane_error_is_raised_and_does_not_deadlock():
    fl = _fl(2)

    def step(lane, i):
        if i == 3:
            raise ValueError("boom")
        with fl.ordered(i):
            pass

    with pytest.raises((ValueError, RuntimeError)):
        fl.run(step, 8)
    with pytest.raises(ValueError):
        InFlight([])
