"""Minimal discrete-event core with simpy 4.0.1's semantics, sufficient for the reference's coordsim
simulator (reference coordsim/simulation/flowsimulator.py:24-139, controller/duration_controller.py:
10-77): ``Environment`` with ``now / process / timeout / event / run(until) / step``, events with
callbacks, generator-driven processes.  Event order is simpy's: (time, priority, insertion id) with
URGENT before NORMAL.  Stand-in for the un-installable dependency; no part of the CUDA hot path."""
from heapq import heappop, heappush
from itertools import count

from . import events  # noqa: F401  (simpy.events.Event is referenced by the reference)
from .events import NORMAL, URGENT, Event, Process, Timeout

__all__ = ["Environment", "Event", "Timeout", "Process", "Interrupt"]
__version__ = "4.0.1-compat"


class Interrupt(Exception):
    pass


class EmptySchedule(Exception):
    pass


class StopSimulation(Exception):
    @classmethod
    def callback(cls, event):
        if event._ok:
            raise cls(event._value)
        raise event._value


class Environment:
    def __init__(self, initial_time=0):
        self._now = initial_time
        self._queue = []
        self._eid = count()
        self._active_proc = None

    @property
    def now(self):
        return self._now

    @property
    def active_process(self):
        return self._active_proc

    def schedule(self, event, priority=NORMAL, delay=0):
        heappush(self._queue, (self._now + delay, priority, next(self._eid), event))

    def peek(self):
        return self._queue[0][0] if self._queue else float("inf")

    def event(self):
        return Event(self)

    def timeout(self, delay=0, value=None):
        return Timeout(self, delay, value)

    def process(self, generator):
        return Process(self, generator)

    def step(self):
        try:
            self._now, _, _, event = heappop(self._queue)
        except IndexError:
            raise EmptySchedule()
        callbacks, event.callbacks = event.callbacks, None
        for cb in callbacks:
            cb(event)
        if not event._ok and not getattr(event, "_defused", False):
            raise event._value

    def run(self, until=None):
        if until is not None:
            if not isinstance(until, Event):
                at = float(until)
                if at <= self.now:
                    raise ValueError(f"until(={at}) must be > the current simulation time.")
                until = Event(self)
                until._ok = True
                until._value = None
                self.schedule(until, URGENT, at - self.now)
            elif until.callbacks is None:
                return until.value
            until.callbacks.append(StopSimulation.callback)
        try:
            while True:
                self.step()
        except StopSimulation as exc:
            return exc.args[0]
        except EmptySchedule:
            if until is not None and not until.triggered:
                raise RuntimeError(f'No scheduled events left but "until" event was not triggered: {until}')
        return None
