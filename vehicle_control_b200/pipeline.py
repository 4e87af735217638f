"""Several batches in flight on one GPU.

A batched NMPC solve ends in a latency-bound tail: after ~50 lock-step passes a few dozen hard instances are left, each
pass of them costs about a millisecond whatever the batch size, and the GPU idles.  Instances of
different batches are as independent as instances of one batch (simulation/racing.py:230-241 of the reference steps its
controllers one after the other with no interaction), so the tail of one batch can run under the full passes of the
next: every LANE owns a handle (include/vc_api.h: one vc_handle per stream), a non-blocking CUDA stream and its solver
workspace, and is driven by its own host thread (the C-ABI calls release the GIL).  Step i goes to lane i mod L; a lane
executes its steps in order.  Results are bit-identical to solving the batches one after the other: a handle never
shares state with another handle.

`ordered(i)` gives the steps a global order for the sections that need one (a collective over the ranks must be issued
in the same order by every rank).
"""
from __future__ import annotations

import threading
import time
from typing import Callable, List, Optional, Sequence


class Lane:
    """One handle + stream.  `stream` is a torch.cuda.Stream (None in host-only tests)."""

    def __init__(self, index: int, engine, stream=None):
        self.index, self.eng, self.stream = index, engine, stream
        self.buf = {}     # per-lane work buffers of the caller

    def __enter__(self):
        if self.stream is not None:
            import torch
            self._ctx = torch.cuda.stream(self.stream)
            self._ctx.__enter__()
        return self

    def __exit__(self, *a):
        if self.stream is not None:
            self._ctx.__exit__(*a)

    def synchronize(self):
        if self.stream is not None:
            self.stream.synchronize()


class InFlight:
    def __init__(self, lanes: Sequence[Lane]):
        if not lanes:
            raise ValueError("InFlight needs at least one lane")
        self.lanes: List[Lane] = list(lanes)
        self._cv = threading.Condition()
        self._next = 0

    @classmethod
    def create(cls, car_config, mpc_config, track, device: int, n_lanes: int, blocking_sync: bool = True):
        """n_lanes engines on `device`, each on a non-blocking stream of its own.  blocking_sync: the lanes' host threads
        sleep while they wait for the GPU instead of spinning (there may be more lanes x ranks than cores)."""
        import torch
        from .engine import Engine
        lanes = []
        for i in range(n_lanes):
            st = torch.cuda.Stream(device=device)
            lanes.append(Lane(i, Engine(car_config, mpc_config, track, device=device, stream=st.cuda_stream), st))
            lanes[-1].eng.set_blocking_sync(blocking_sync)
        return cls(lanes)

    def ordered(self, i: int):
        """Context manager: the body of step i runs after the bodies of steps 0..i-1 (of this run) have finished."""
        return _Ordered(self, i)

    def run(self, step_fn: Callable[[Lane, int], None], n_steps: int, first: Optional[Callable[[Lane], None]] = None,
            last: Optional[Callable[[Lane], None]] = None, stagger_s: float = 0.0):
        """step_fn(lane, i) for i in range(n_steps), step i on lane i mod L, one host thread per lane; `first` / `last`
        run on every lane (in its stream context) before its first / after its last step.  Returns when every lane has
        issued all its work (not a device synchronisation).  The first exception of a lane is re-raised.
        stagger_s: lane l starts l * stagger_s seconds late (lanes that start together reach their tails together; on
        the C3 bench this made no measurable difference -- the lanes drift apart by themselves)."""
        self._next = 0
        errors: List[BaseException] = []
        L = len(self.lanes)

        def loop(lane: Lane):
            try:
                if stagger_s > 0.0 and lane.index > 0:
                    time.sleep(lane.index * stagger_s)
                with lane:
                    if first:
                        first(lane)
                    for i in range(lane.index, n_steps, L):
                        step_fn(lane, i)
                    if last:
                        last(lane)
            except BaseException as e:   # noqa: BLE001
                errors.append(e)
                with self._cv:            # do not leave the other lanes waiting for this lane's ordered sections
                    self._next = 1 << 60
                    self._cv.notify_all()

        threads = [threading.Thread(target=loop, args=(ln,), name=f"vc-lane-{ln.index}") for ln in self.lanes[:max(1, min(L, n_steps))]]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        if errors:
            raise errors[0]

    def synchronize(self):
        for ln in self.lanes:
            ln.synchronize()


class _Ordered:
    def __init__(self, owner: InFlight, i: int):
        self.o, self.i = owner, i

    def __enter__(self):
        with self.o._cv:
            self.o._cv.wait_for(lambda: self.o._next >= self.i)
            if self.o._next > self.i:
                raise RuntimeError("another lane failed")
        return self

    def __exit__(self, et, ev, tb):
        with self.o._cv:
            if et is None:
                self.o._next = self.i + 1
            self.o._cv.notify_all()
        return False


def run_fleet_in_lanes(carconfig, trackconfig, controller_config, x0, n_lanes: int, device: int = 0, max_ticks: int = 3000,
                       obstacles=None, record: bool = False, seed: Optional[int] = 31):
    """Closed-loop lap of a fleet (BASELINE config C4) with the vehicles split into `n_lanes` contiguous groups, each a
    BatchedRacingSimulator on a lane of its own.  A closed loop needs every action of a group before its plant step, but
    the groups do not interact (simulation/racing.py:230-241 of the reference), so while one group's tick is in its
    latency-bound tail the others' full passes fill the GPU.  Every vehicle's trajectory is what the single-handle
    simulator gives (instances never influence each other).  `seed`: numpy's legacy generator is seeded with it before
    each group's controller draws its cold guess (cascaded_mpc.py:13 seeds 31 at import), so every group starts from the
    guess a single simulator would use.  Returns the list of simulators (group order = vehicle order)."""
    import numpy as np
    from .environment.track import Track
    from .simulation.racing import BatchedRacingSimulator
    from .utils.common_utils import as_config
    x0 = np.ascontiguousarray(x0, dtype=float)
    track = trackconfig if isinstance(trackconfig, Track) else Track(trackconfig)
    n_lanes = max(1, min(int(n_lanes), len(x0)))
    fl = InFlight.create(as_config(carconfig), as_config(controller_config), track, device, n_lanes)
    bounds = [len(x0) * i // n_lanes for i in range(n_lanes + 1)]
    sims = []
    for i, lane in enumerate(fl.lanes):          # built one after the other (the cold guess comes from a global generator)
        ob = None if obstacles is None else obstacles[bounds[i]:bounds[i + 1]]
        with lane:
            if seed is not None:
                np.random.seed(seed)
            sims.append(BatchedRacingSimulator(carconfig, track, controller_config, x0[bounds[i]:bounds[i + 1]], obstacles=ob,
                                               record=record, engine=lane.eng))
        lane.synchronize()
    fl.run(lambda lane, i: sims[i].run(max_ticks), n_lanes)
    fl.synchronize()
    return sims
