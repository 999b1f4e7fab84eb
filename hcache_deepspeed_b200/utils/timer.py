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


class _Interval:
    """One measured span. On a GPU it is a pair of CUDA events on the current stream (nothing is synchronised until the
    duration is read — the way every kernel in this repository is timed); on the host it is two ``perf_counter`` reads."""
    __slots__ = ("t0", "t1", "on_device")

    def __init__(self, on_device: bool):
        self.on_device = on_device
        if on_device:
            self.t0 = torch.cuda.Event(enable_timing=True)
            self.t0.record()
        else:
            self.t0 = time.perf_counter()
        self.t1 = None

    def close(self) -> None:
        if self.on_device:
            self.t1 = torch.cuda.Event(enable_timing=True)
            self.t1.record()
        else:
            self.t1 = time.perf_counter()

    def milliseconds(self) -> float:
        if self.on_device:
            self.t1.synchronize()
            return self.t0.elapsed_time(self.t1)
        return (self.t1 - self.t0) * 1e3


class CudaEventTimer:
    """A closed span between two recorded CUDA events (kept for API parity: ``get_elapsed_msec``)."""

    def __init__(self, start_event, end_event):
        self.start_event, self.end_event = start_event, end_event

    def get_elapsed_msec(self) -> float:
        self.end_event.synchronize()
        return self.start_event.elapsed_time(self.end_event)


class SynchronizedWallClockTimer:
    """Named timers for the training engine's forward / backward / step breakdown (ref role:
    ``deepspeed/utils/timer.py:44-197``). A timer accumulates the spans between its ``start`` / ``stop`` calls;
    ``elapsed`` sums them in milliseconds (and by default forgets them), ``mean`` averages them."""

    class Timer:

        def __init__(self, name: str):
            self.name = name
            self._open: Optional[_Interval] = None
            self._closed: List[_Interval] = []
            self._last_ms: List[float] = []

        @property
        def running(self) -> bool:
            return self._open is not None

        def start(self):
            if self._open is not None:
                raise AssertionError(f"{self.name} timer has already been started")
            self._open = _Interval(torch.cuda.is_available())

        def stop(self, reset: bool = False, record: bool = False):
            if self._open is None:
                raise AssertionError(f"{self.name} timer is not started")
            self._open.close()
            self._closed.append(self._open)
            self._open = None

        def _spans_ms(self) -> List[float]:
            """Milliseconds of every closed span (device events are resolved once and cached as plain numbers)."""
            self._closed = [iv if isinstance(iv, float) else iv.milliseconds() for iv in self._closed]
            self._last_ms = list(self._closed)
            return self._last_ms

        def reset(self):
            self._open, self._closed, self._last_ms = None, [], []

        def elapsed(self, reset: bool = True) -> float:
            """Milliseconds accumulated so far; a running timer is paused for the reading and resumed afterwards. With
            ``reset=False`` the accumulated spans are kept, so a later reading still includes them."""
            was_running = self.running
            if was_running:
                self.stop()
            total = sum(self._spans_ms())
            if reset:
                self.reset()
            if was_running:
                self.start()
            return total

        def mean(self) -> float:
            was_running = self.running
            if was_running:
                self.stop()
            ms = self._spans_ms()
            if was_running:
                self.start()
            return sum(ms) / len(ms) if ms else 0.0

    def __init__(self):
        self.timers: Dict[str, "SynchronizedWallClockTimer.Timer"] = {}

    def get_timers(self):
        return self.timers

    def has_timer(self, name: str) -> bool:
        return name in self.timers

    def __call__(self, name: str) -> "SynchronizedWallClockTimer.Timer":
        timer = self.timers.get(name)
        if timer is None:
            timer = self.timers[name] = self.Timer(name)
        return timer

    @staticmethod
    def memory_usage() -> str:
        if not torch.cuda.is_available():
            return "no accelerator"
        gib = float(1 << 30)
        parts = (("mem alloc", torch.cuda.memory_allocated()), ("max mem alloc", torch.cuda.max_memory_allocated()),
                 ("mem reserved", torch.cuda.memory_reserved()))
        return "".join(f" | {label} (GB) {value / gib:.4f}" for label, value in parts)

    def log(self, names, normalizer: float = 1.0, reset: bool = True, memory_breakdown: bool = False, ranks=None):
        if normalizer <= 0.0:
            raise AssertionError("normalizer must be positive")
        fields = [f"{n}: {self.timers[n].elapsed(reset=reset) / normalizer:.2f}" for n in names if n in self.timers]
        line = " | ".join(["time (ms)"] + fields)
        if memory_breakdown:
            line += self.memory_usage()
        log_dist(line, ranks=ranks or [0])

    def get_mean(self, names, normalizer: float = 1.0, reset: bool = True) -> Dict[str, float]:
        if normalizer <= 0.0:
            raise AssertionError("normalizer must be positive")
        return {n: self.timers[n].mean() * 1000.0 / normalizer for n in names if n in self.timers}


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


def trim_mean(data, trim_percent: float) -> float:
    """Mean after dropping the ``trim_percent`` smallest and largest share of ``data`` (0 for an empty list)."""
    assert 0.0 <= trim_percent <= 1.0
    n = len(data)
    if n == 0:
        return 0
    ordered = sorted(data)
    k = int(round(n * trim_percent))
    kept = ordered[k:n - k] or ordered
    return sum(kept) / len(kept)
