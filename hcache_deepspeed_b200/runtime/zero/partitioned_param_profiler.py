"""Event counters / timers of the ZeRO-3 parameter coordinator (``partitioned_param_profiler.py:10-63``)."""
from __future__ import annotations

import time
from dataclasses import dataclass
from typing import Dict


class PartitionedParameterProfiler:

    @dataclass
    class EventCounter:
        name: str
        count: int = 0
        num_elem: int = 0

        def reset(self):
            self.count = self.num_elem = 0

        def increment(self, numel: int):
            self.count += 1
            self.num_elem += numel

    def __init__(self, timers=None):
        self.timers = timers
        self.event_counters: Dict[str, "PartitionedParameterProfiler.EventCounter"] = {}
        self._started: Dict[str, float] = {}
        self.elapsed: Dict[str, float] = {}

    def reset_events(self):
        for c in self.event_counters.values():
            c.reset()

    def start_event(self, name: str):
        if self.timers is not None:
            self.timers(name).start()
        self._started[name] = time.perf_counter()
        self.event_counters.setdefault(name, self.EventCounter(name))

    def stop_event(self, name: str, num_elem: int):
        if self.timers is not None:
            self.timers(name).stop()
        self.elapsed[name] = self.elapsed.get(name, 0.0) + time.perf_counter() - self._started.pop(name, time.perf_counter())
        self.event_counters.setdefault(name, self.EventCounter(name)).increment(num_elem)

    def log_events(self):
        from ...utils.logging import log_dist
        for name, c in self.event_counters.items():
            log_dist(f"{name}: count = {c.count}, numel = {c.num_elem}, seconds = {self.elapsed.get(name, 0.0):.4f}", [0])
