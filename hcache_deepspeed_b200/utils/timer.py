"""Timers (ref: deepspeed/utils/timer.py:44-290): CUDA-event based wall-clock timers that do not synchronise
until read, a no-op stand-in, and a throughput timer."""
from __future__ import annotations

import time
from typing import Dict, List, Optional

import torch

from .logging import log_dist

FORWARD_MICRO_TIMER = "fwd_microstep"
FORWARD_GLOBAL_TIMER = "fwd"
BACKWARD_MICRO_TIMER = "bwd_microstep"
BACKWARD_GLOBAL_TIMER = "bwd"
STEP_MICRO_TIMER = "step_microstep"
STEP_GLOBAL_TIMER = "step"


class CudaEventTimer:

    def __init__(self, start_event, end_event):
        self.start_event, self.end_event = start_event, end_event

    def get_elapsed_msec(self) -> float:
        self.end_event.synchronize()
        return self.start_event.elapsed_time(self.end_event)


class SynchronizedWallClockTimer:
    """Group of named timers; on a GPU they record CUDA events and only synchronise when ``elapsed`` is read."""

    class Timer:

        def __init__(self, name: str):
            self.name_ = name
            self.started_ = False
            self.event_timers: List[CudaEventTimer] = []
            self.use_host_timer = not torch.cuda.is_available()
            self.start_event = None
            self.elapsed_records: Optional[List[float]] = None
            self.start_time = 0.0
            self.end_time = 0.0

        def start(self):
            assert not self.started_, f"{self.name_} timer has already been started"
            if self.use_host_timer:
                self.start_time = time.time()
            else:
                self.start_event = torch.cuda.Event(enable_timing=True)
                self.start_event.record()
            self.started_ = True

        def stop(self, reset: bool = False, record: bool = False):
            assert self.started_, "timer is not started"
            if self.use_host_timer:
                self.end_time = time.time()
                self.event_timers.append(self.end_time - self.start_time)
            else:
                end = torch.cuda.Event(enable_timing=True)
                end.record()
                self.event_timers.append(CudaEventTimer(self.start_event, end))
                self.start_event = None
            self.started_ = False

        def _get_elapsed_msec(self) -> float:
            if self.use_host_timer:
                self.elapsed_records = [et * 1000.0 for et in self.event_timers]
            else:
                self.elapsed_records = [et.get_elapsed_msec() for et in self.event_timers]
            self.event_timers.clear()
            return sum(self.elapsed_records)

        def reset(self):
            self.started_ = False
            self.start_event = None
            self.elapsed_records = None
            self.event_timers.clear()

        def elapsed(self, reset: bool = True) -> float:
            started = self.started_
            if started:
                self.stop()
            elapsed = self._get_elapsed_msec()
            if reset:
                self.reset()
            if started:
                self.start()
            return elapsed

        def mean(self) -> float:
            self.elapsed(reset=False)
            recs = self.elapsed_records or []
            return sum(recs) / len(recs) if recs else 0.0

    def __init__(self):
        self.timers: Dict[str, "SynchronizedWallClockTimer.Timer"] = {}

    def get_timers(self):
        return self.timers

    def __call__(self, name: str) -> "SynchronizedWallClockTimer.Timer":
        if name not in self.timers:
            self.timers[name] = self.Timer(name)
        return self.timers[name]

    @staticmethod
    def memory_usage() -> str:
        if not torch.cuda.is_available():
            return "no accelerator"
        g = 1024 ** 3
        return (f" | mem alloc (GB) {torch.cuda.memory_allocated() / g:.4f} | max mem alloc (GB) "
                f"{torch.cuda.max_memory_allocated() / g:.4f} | mem reserved (GB) {torch.cuda.memory_reserved() / g:.4f}")

    def log(self, names, normalizer: float = 1.0, reset: bool = True, memory_breakdown: bool = False, ranks=None):
        assert normalizer > 0.0
        string = "time (ms)"
        for name in names:
            if name in self.timers:
                string += " | {}: {:.2f}".format(name, self.timers[name].elapsed(reset=reset) / normalizer)
        if memory_breakdown:
            string += self.memory_usage()
        log_dist(string, ranks=ranks or [0])

    def get_mean(self, names, normalizer: float = 1.0, reset: bool = True) -> Dict[str, float]:
        assert normalizer > 0.0
        means = {}
        for name in names:
            if name in self.timers:
                means[name] = self.timers[name].mean() * 1000.0 / normalizer
        return means


class NoopTimer:

    class Timer:

        def start(self):
            ...

        def reset(self):
            ...

        def stop(self, **kwargs):
            ...

        def elapsed(self, **kwargs):
            return 0

        def mean(self):
            return 0

    def __init__(self):
        self.timer = self.Timer()

    def __call__(self, name):
        return self.timer

    def get_timers(self):
        return {}

    def log(self, names, normalizer=1.0, reset=True, memory_breakdown=False, ranks=None):
        ...

    def get_mean(self, names, normalizer=1.0, reset=True):
        ...


class ThroughputTimer:
    """Samples (or tokens) per second with warm-up steps excluded (ref: timer.py:199-290)."""

    def __init__(self, batch_size: int, start_step: int = 2, steps_per_output: Optional[int] = None,
                 monitor_memory: bool = False, logging_fn=None):
        self.start_time = 0.0
        self.end_time = 0.0
        self.started = False
        self.batch_size = max(1, batch_size)
        self.start_step = start_step
        self.epoch_count = 0
        self.micro_step_count = 0
        self.global_step_count = 0
        self.total_elapsed_time = 0.0
        self.step_elapsed_time = 0.0
        self.steps_per_output = steps_per_output
        self.monitor_memory = monitor_memory
        self.logging = logging_fn or (lambda msg: log_dist(msg, ranks=[0]))
        self.initialized = False

    def update_epoch_count(self):
        self.epoch_count += 1
        self.micro_step_count = 0

    def _init_timer(self):
        self.initialized = True

    def start(self):
        self._init_timer()
        self.started = True
        if self.global_step_count >= self.start_step:
            if torch.cuda.is_available():
                torch.cuda.synchronize()
            self.start_time = time.time()

    def stop(self, global_step: bool = False, report_speed: bool = True):
        if not self.started:
            return
        self.started = False
        self.micro_step_count += 1
        if global_step:
            self.global_step_count += 1
        if self.start_time > 0:
            if torch.cuda.is_available():
                torch.cuda.synchronize()
            self.end_time = time.time()
            duration = self.end_time - self.start_time
            self.total_elapsed_time += duration
            self.step_elapsed_time += duration
            if global_step:
                if report_speed and self.steps_per_output and self.global_step_count % self.steps_per_output == 0:
                    self.logging("epoch={}/micro_step={}/global_step={}, RunningAvgSamplesPerSec={}, CurrSamplesPerSec={}".format(
                        self.epoch_count, self.micro_step_count, self.global_step_count, self.avg_samples_per_sec(),
                        self.batch_size / self.step_elapsed_time))
                self.step_elapsed_time = 0.0

    def avg_samples_per_sec(self) -> float:
        if self.global_step_count > self.start_step and self.total_elapsed_time > 0:
            return self.batch_size / (self.total_elapsed_time / (self.global_step_count - self.start_step))
        return float("-inf")
