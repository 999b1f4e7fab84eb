"""List scheduling of gathers under a memory cap (reference ``deepspeed/compile/list_schedule.py`` ``fast_free_schedule``
orders graph nodes so that memory is freed as early as possible).

On the module trace the compute order is fixed; what can move is WHEN each parameter's gather is launched. The
scheduler walks the trace once and launches every gather as early as the live-parameter cap allows, releasing a
parameter right after its last use inside the same phase."""
from __future__ import annotations

from typing import Dict, List

from .trace import ModuleTrace


def schedule_gathers(trace: ModuleTrace, max_live_numel: int, max_ahead: int = 8) -> Dict[int, List]:
    """event index -> parameters to launch there. Memory model: a parameter is live from its launch to the event that
    uses it (it is released after that event)."""
    ev = trace.events
    launch: Dict[int, List] = {e.index: [] for e in ev}
    live_until: List[tuple] = []          # (use_index, numel) of launched, not yet consumed gathers
    claimed = set()
    for i, e in enumerate(ev):
        live_until = [(u, n) for u, n in live_until if u >= i]
        live = sum(n for _, n in live_until)
        for j in range(i + 1, min(len(ev), i + 1 + max_ahead)):
            for p in ev[j].params:
                key = (id(p), j)
                if key in claimed or getattr(p, "ds_persist", False) or any(p is q for q in e.params):
                    continue
                if live + p.ds_numel > max_live_numel:
                    break
                claimed.add(key)
                launch[i].append(p)
                live_until.append((j, p.ds_numel))
                live += p.ds_numel
            else:
                continue
            break
    return launch


def peak_live_numel(trace: ModuleTrace, launch: Dict[int, List]) -> int:
    """Peak of (own parameters of the running event + launched-ahead parameters) over the trace."""
    ahead: List[tuple] = []
    peak = 0
    for i, e in enumerate(trace.events):
        ahead = [(p, n) for p, n in ahead if not any(p is q for q in e.params)]
        ahead += [(p, p.ds_numel) for p in launch.get(i, [])]
        peak = max(peak, e.numel + sum(n for _, n in ahead))
    return peak
