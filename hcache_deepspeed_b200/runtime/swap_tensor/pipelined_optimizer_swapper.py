"""Pipelined optimizer-state swapper (``deepspeed/runtime/swap_tensor/pipelined_optimizer_swapper.py:52-280``): while
sub-group ``i`` is being updated, the read of ``i + 1`` and the write of ``i - 1`` are in flight on separate aio
handles (``pipeline_read`` / ``pipeline_write`` of ``offload_optimizer``)."""
from __future__ import annotations

from .optimizer_utils import OptimizerSwapper


class PipelinedOptimizerSwapper(OptimizerSwapper):

    def __init__(self, *args, pipeline_read: bool = True, pipeline_write: bool = True, **kw):
        swap_config = args[0] if args else kw.get("swap_config")
        self.pipeline_read = bool(getattr(swap_config, "pipeline_read", pipeline_read)) if swap_config is not None else pipeline_read
        self.pipeline_write = bool(getattr(swap_config, "pipeline_write", pipeline_write)) if swap_config is not None else pipeline_write
        super().__init__(*args, **kw)

    def swap_in_optimizer_state(self, parameter, async_parameter=None):
        super().swap_in_optimizer_state(parameter, async_parameter if self.pipeline_read else None)

    def swap_out_optimizer_state(self, parameter, async_swap: bool = True):
        super().swap_out_optimizer_state(parameter, async_swap and self.pipeline_write)
