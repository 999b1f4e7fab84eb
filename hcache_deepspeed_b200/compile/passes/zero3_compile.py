"""ZeRO-3 gather/release insertion (reference ``deepspeed/compile/passes/zero3_compile.py`` ``add_z3_gather_release``
rewrites the FX graph with allgather / release / reduce nodes). Here the same three operations hang on module and
parameter hooks, installed by :class:`~hcache_deepspeed_b200.runtime.zero.stage3.DeepSpeedZeRoOffload`."""
from __future__ import annotations

NAME = "zero3_compile"


def add_z3_gather_release(module, dp_process_group=None, prefetch_bucket_size: int = 0, **kw):
    from ...runtime.zero.stage3 import DeepSpeedZeRoOffload
    return DeepSpeedZeRoOffload(module, dp_process_group=dp_process_group, prefetch_bucket_size=prefetch_bucket_size, **kw)
