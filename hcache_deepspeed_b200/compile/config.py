"""``"compile"`` section of the training config (field names of ``deepspeed/compile/config.py:9-48``)."""
from __future__ import annotations

from ..config.config_utils import DeepSpeedConfigModel


class CompileConfig(DeepSpeedConfigModel):
    deepcompile: bool = False             # run the schedule passes below on the recorded module trace
    free_activation: bool = False
    offload_activation: bool = False      # checkpointed activations -> pinned host memory on a side stream
    offload_opt_states: bool = False      # Adam moments live on the host between optimizer steps
    double_buffer: bool = True            # reload of optimizer states overlaps the backward pass
    symmetric_memory: bool = False
    debug_log: bool = False
    offload_parameters: bool = False
    sync_before_reduce: bool = False
    sync_after_reduce: bool = False
    sync_before_allgather: bool = False
    sync_after_allgather: bool = False
    # knobs of the passes
    selective_gather_budget_bytes: int = -1      # -1: derive from free device memory after the profiled step
    prefetch_bucket_size: int = 50_000_000
