"""The "graph" the schedule passes work on.

The reference's DeepCompile (``deepspeed/compile/backend.py``, ``fx.py``, ``graph_param.py``, ``profilers/``) captures an
FX graph through ``torch.compile`` and rewrites it. This runtime has no tracing compiler on its path: the unit of
scheduling is the MODULE, and the graph is the order in which the ZeRO-3 hooks saw modules run during one profiled
iteration (:class:`~hcache_deepspeed_b200.runtime.zero.stage3.DeepSpeedZeRoOffload` records it). Each event knows the
partitioned parameters it needs and how long it ran, which is all the passes (prefetch, selective gather, offload of
optimizer state) consume."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List

import torch


@dataclass
class TraceEvent:
    index: int
    module_name: str
    phase: str                       # "fwd" | "bwd"
    params: List[torch.nn.Parameter]
    seconds: float = 0.0             # wall time until the next event (compute of this module, roughly)

    @property
    def numel(self) -> int:
        return sum(p.ds_numel if hasattr(p, "ds_numel") else p.numel() for p in self.params)


@dataclass
class ModuleTrace:
    events: List[TraceEvent] = field(default_factory=list)

    @classmethod
    def from_offload(cls, offload) -> "ModuleTrace":
        """Snapshot of the trace the ZeRO-3 hook installer recorded (it must have completed one iteration)."""
        if not offload.trace_complete:
            raise RuntimeError("no complete trace yet: run one forward/backward before compiling the schedule")
        names = {id(m): n for n, m in offload.module.named_modules()}
        times = offload.event_times
        ev = []
        for i, (m, phase) in enumerate(offload.trace):
            dt = times[i + 1] - times[i] if i + 1 < len(times) else 0.0
            ev.append(TraceEvent(i, names.get(id(m), type(m).__name__), phase, offload._own_params(m), dt))
        return cls(ev)

    def uses(self) -> Dict[int, List[int]]:
        """param id -> indices of the events that need it."""
        out: Dict[int, List[int]] = {}
        for e in self.events:
            for p in e.params:
                out.setdefault(id(p), []).append(e.index)
        return out

    def param_table(self) -> Dict[int, torch.nn.Parameter]:
        return {id(p): p for e in self.events for p in e.params}

    def summary(self) -> str:
        rows = [f"{e.index:4d} {e.phase} {e.module_name:40s} numel={e.numel:>12d} {e.seconds * 1e3:8.3f} ms" for e in self.events]
        return "\n".join(rows)
