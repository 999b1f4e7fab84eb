"""Selective gather (reference ``deepspeed/compile/passes/selective_gather.py``): with memory to spare after the
profiled iteration, keep some parameters gathered for good instead of re-gathering them in every forward AND backward.

Cost model: a gather of ``n`` bytes costs ``alpha + n / bandwidth``; keeping the parameter saves that once per use and
costs ``n * (1 - 1/world)`` bytes of memory. Benefit per byte is therefore highest for SMALL, OFTEN-USED parameters
(the launch latency dominates them), so the greedy order is by ``uses * (alpha + n/bw) / n``."""
from __future__ import annotations

from typing import List

from ..trace import ModuleTrace

NAME = "selective_gather"


def selective_gather(trace: ModuleTrace, budget_bytes: int, world_size: int, alpha_s: float = 12e-6,
                     bandwidth_Bps: float = 600e9) -> List:
    """Parameters to make persistent, best first, within ``budget_bytes`` of extra resident memory."""
    if budget_bytes <= 0 or world_size <= 1:
        return []
    uses, table = trace.uses(), trace.param_table()
    cand = []
    for pid, idx in uses.items():
        p = table[pid]
        if getattr(p, "ds_persist", False):
            continue
        nbytes = p.ds_numel * p.element_size()
        extra = nbytes - nbytes // world_size
        saved = len(idx) * (alpha_s + nbytes / bandwidth_Bps)
        cand.append((saved / max(extra, 1), extra, p))
    cand.sort(key=lambda c: -c[0])
    chosen, left = [], budget_bytes
    for _, extra, p in cand:
        if extra <= left:
            chosen.append(p)
            left -= extra
    return chosen
