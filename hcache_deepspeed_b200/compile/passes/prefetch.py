"""Prefetch scheduling (reference ``deepspeed/compile/passes/prefetch.py`` ``schedule_prefetch``: move all-gathers
earlier in the graph and fuse them up to a byte limit).

Here: for event ``i`` of the module trace, the list of parameters whose gathers should be LAUNCHED when ``i`` starts, so
they fly while ``i`` computes. Walking forward from ``i + 1`` collects parameters until the bucket is full."""
from __future__ import annotations

from typing import List, Sequence, Tuple

NAME = "prefetch"


def _numel(p) -> int:
    return p.ds_numel if hasattr(p, "ds_numel") else p.numel()


def schedule_prefetch(events: Sequence[Tuple[list, str]], bucket_numel: int, lookahead: int = 8) -> List[list]:
    """``events``: ``[(params, phase)]`` in execution order. Returns ``plan[i]`` = parameters to launch at event ``i``.
    Persistent parameters are skipped (they are always resident); at run time, parameters that are already gathered or
    in flight are skipped too, so overlapping windows of neighbouring events cost nothing."""
    n = len(events)
    plan: List[list] = [[] for _ in range(n)]
    if bucket_numel <= 0:
        return plan
    for i in range(n):
        budget = bucket_numel
        mine = {id(p) for p in events[i][0]}
        seen = set()
        full = False
        for j in range(i + 1, min(n, i + 1 + lookahead)):
            for p in events[j][0]:
                if id(p) in mine or id(p) in seen or getattr(p, "ds_persist", False):
                    continue
                if _numel(p) > budget and plan[i]:
                    full = True          # gathers are launched in use order: do not jump over one that does not fit
                    break
                seen.add(id(p))
                plan[i].append(p)
                budget -= _numel(p)
            if full or budget <= 0:
                break
    return plan
