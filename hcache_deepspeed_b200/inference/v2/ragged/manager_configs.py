"""Configuration of the ragged state manager and the paged KV cache.

Field names, defaults and validation follow the reference so existing configs load unchanged
(``deepspeed/inference/v2/ragged/manager_configs.py:15-181``). ``offload`` is honoured here (the
reference accepts it but its KV offload raises NotImplementedError, ``ragged/kv_cache.py:166-184``).
"""
from __future__ import annotations

from enum import Enum
from typing import Tuple

from pydantic import PositiveInt, model_validator

from ....config import DeepSpeedConfigModel, DtypeEnum


class KVCacheType(Enum):
    DENSE = "dense"
    """Every token's K/V is kept (the only type implemented, as in the reference)."""
    LOCAL = "local"
    """Sliding-window cache (declared for config compatibility)."""


class KVCacheConfig(DeepSpeedConfigModel):
    type: KVCacheType = KVCacheType.DENSE
    block_size: int = 128
    """Tokens per KV block."""
    num_allocation_groups: PositiveInt = 1
    cache_shape: Tuple[PositiveInt, PositiveInt, PositiveInt]
    """(num_layers, num_kv_heads on this rank, head_dim)."""
    cache_dtype: DtypeEnum = DtypeEnum.fp16
    max_blocks_per_allocation_group: PositiveInt = 64
    """Upper bound of blocks one sequence may own (= ceil(max_context / block_size))."""


class AllocationMode(Enum):
    RESERVE = "reserve"
    """Use all free device memory except ``size`` bytes."""
    ALLOCATE = "allocate"
    """Use exactly ``size`` bytes."""


class MemoryConfig(DeepSpeedConfigModel):
    mode: AllocationMode = AllocationMode.RESERVE
    size: PositiveInt = 1_000_000_000


class DSStateManagerConfig(DeepSpeedConfigModel):
    max_tracked_sequences: PositiveInt = 2048
    """Sequences whose KV state may be resident at once."""
    max_ragged_batch_size: PositiveInt = 768
    """Tokens per forward."""
    max_ragged_sequence_count: PositiveInt = 512
    """Sequences per forward."""
    max_context: PositiveInt = 8192
    """Longest sequence (prompt + generation) the engine will hold."""
    memory_config: MemoryConfig = MemoryConfig()
    offload: bool = False
    """Allow whole-sequence KV offload to pinned host memory (``DSStateManager.offload_sequence``)."""

    @model_validator(mode="after")
    def _max_ragged_sequence_count_validator(self):
        if self.max_ragged_sequence_count > self.max_tracked_sequences:
            raise ValueError("max_ragged_sequence_count must be less than max_tracked_sequences")
        if self.max_ragged_sequence_count > self.max_ragged_batch_size:
            raise ValueError("max_ragged_sequence_count must be less than max_ragged_batch_size")
        return self
