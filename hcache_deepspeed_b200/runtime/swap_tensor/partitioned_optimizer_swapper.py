"""Synchronous optimizer-state swapper (``deepspeed/runtime/swap_tensor/partitioned_optimizer_swapper.py:27-220``):
every swap-in / swap-out completes before it returns."""
from __future__ import annotations

from .optimizer_utils import OptimizerSwapper


class PartitionedOptimizerSwapper(OptimizerSwapper):

    def swap_in_optimizer_state(self, parameter, async_parameter=None):
        super().swap_in_optimizer_state(parameter, None)

    def swap_out_optimizer_state(self, parameter, async_swap: bool = False):
        super().swap_out_optimizer_state(parameter, False)
