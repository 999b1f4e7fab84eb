"""Host-side description of one tracked sequence.

Behavioural counterpart of ``DSSequenceDescriptor`` / ``PlaceholderSequenceDescriptor``
(``deepspeed/inference/v2/ragged/sequence_descriptor.py:11-280``). Differences: block ids live in one
host int32 array read directly by the native batch builder (no per-sequence device tables, no raw
device pointers), and ``free_kv_cache`` is implemented.
"""
from __future__ import annotations

from typing import List, Optional

import torch


class BaseSequenceDescriptor:

    @property
    def seen_tokens(self) -> int:
        raise NotImplementedError()

    @property
    def cur_allocated_blocks(self) -> int:
        raise NotImplementedError()

    def kv_blocks_ptr(self, cache_group: int = 0) -> int:
        raise NotImplementedError()


class PlaceholderSequenceDescriptor(BaseSequenceDescriptor):
    """Stand-in for a sequence the engine does not know yet (schedulability dry-runs)."""

    def __init__(self, seen_tokens: int = 0, cur_allocated_blocks: int = 0) -> None:
        self._seen_tokens = seen_tokens
        self._cur_allocated_blocks = cur_allocated_blocks

    @property
    def seen_tokens(self) -> int:
        return self._seen_tokens

    @property
    def cur_allocated_blocks(self, cache_group: int = 0) -> int:
        return self._cur_allocated_blocks

    @property
    def in_flight_tokens(self) -> int:
        return 0


class DSSequenceDescriptor(BaseSequenceDescriptor):

    def __init__(self, tracking_id: int, max_blocks: int) -> None:
        self._tracking_id = tracking_id
        self._seen_tokens = 0
        self._in_flight_tokens = 0
        self._block_ids = torch.zeros(max_blocks, dtype=torch.int32)  # host; read by hds_build_batch
        self._n_blocks = 0
        self._max_blocks = max_blocks
        #: set while the sequence's KV lives in host memory (whole-sequence KV offload)
        self.offloaded_kv: Optional[torch.Tensor] = None

    @property
    def max_context(self) -> int:
        """Tokens this sequence's block table can address (blocks x block size is the engine's business; here: blocks)."""
        return self._max_blocks

    # ---- token accounting -------------------------------------------------------------------
    @property
    def seen_tokens(self) -> int:
        return self._seen_tokens

    @property
    def in_flight_tokens(self) -> int:
        return self._in_flight_tokens

    @property
    def tracking_id(self) -> int:
        return self._tracking_id

    def pre_forward(self, num_tokens: int) -> None:
        self._in_flight_tokens = num_tokens

    def post_forward(self) -> None:
        self._seen_tokens += self._in_flight_tokens
        self._in_flight_tokens = 0

    # ---- KV blocks --------------------------------------------------------------------------
    @property
    def cur_allocated_blocks(self) -> int:
        return self._n_blocks

    @property
    def max_blocks(self) -> int:
        return self._max_blocks

    @property
    def all_block_ids(self) -> torch.Tensor:
        return self._block_ids[:self._n_blocks]

    def kv_cache_ids(self, cache_group: int = 0) -> torch.Tensor:
        return self.all_block_ids

    def kv_blocks_ptr(self, cache_group: int = 0) -> int:
        return self._block_ids.data_ptr()

    def extend_kv_cache(self, new_ids: torch.Tensor) -> None:
        n = new_ids.numel()
        if self._n_blocks + n > self._max_blocks:
            raise ValueError(f"sequence would own {self._n_blocks + n} KV blocks, limit is {self._max_blocks}")
        self._block_ids[self._n_blocks:self._n_blocks + n] = new_ids.to(torch.int32)
        self._n_blocks += n

    def free_kv_cache(self, free_ids: Optional[List[int]] = None) -> torch.Tensor:
        """Drop every block (or the listed ones, which must be a suffix) and return the ids to release."""
        if free_ids is None:
            ids = self.all_block_ids.clone()
            self._n_blocks = 0
            return ids
        k = len(free_ids)
        tail = self._block_ids[self._n_blocks - k:self._n_blocks].tolist()
        if sorted(tail) != sorted(int(i) for i in free_ids):
            raise ValueError("only the most recently allocated blocks of a sequence can be freed individually")
        self._n_blocks -= k
        return torch.tensor(list(free_ids), dtype=torch.int32)
