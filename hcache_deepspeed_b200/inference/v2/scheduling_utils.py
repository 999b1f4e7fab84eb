"""Scheduling outcome codes (same members as ``deepspeed/inference/v2/scheduling_utils.py:9-54``)."""
from enum import Enum


class SchedulingResult(Enum):
    Success = 0
    """The proposed batch is valid and can be scheduled."""
    EngineSequenceLimitExceeded = 1
    """More sequences than the engine can track (``max_tracked_sequences``)."""
    BatchSequenceLimitExceeded = 2
    """More sequences than one ragged batch may hold (``max_ragged_sequence_count``)."""
    BatchTokenLimitExceeded = 3
    """More tokens than one ragged batch may hold (``max_ragged_batch_size``)."""
    KVCacheLimitExceeded = 4
    """Not enough free KV blocks."""
    SequenceTokenLimitExceeded = 5
    """A sequence would exceed ``max_context``."""


class SchedulingError(RuntimeError):

    result: SchedulingResult

    def __init__(self, result: SchedulingResult) -> None:
        self.result = result
        super().__init__(f"Batch scheduling failed with result {result}")
