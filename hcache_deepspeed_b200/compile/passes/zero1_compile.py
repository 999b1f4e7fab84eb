"""ZeRO-1 gradient reduction scheduling (reference ``deepspeed/compile/passes/zero1_compile.py`` inserts a
reduce-scatter after each gradient is produced). :class:`~hcache_deepspeed_b200.runtime.zero.stage_1_and_2.DeepSpeedZeroOptimizer`
already reduces bucket by bucket from gradient hooks when ``overlap_comm`` is on; this pass turns that on."""
from __future__ import annotations

NAME = "zero1_compile"


def add_z1_reduce(optimizer) -> None:
    if hasattr(optimizer, "overlap_comm"):
        optimizer.overlap_comm = True
