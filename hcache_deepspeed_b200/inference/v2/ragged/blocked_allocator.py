"""KV block id allocator backed by the native host runtime.

The reference keeps a linked list inside a pinned int32 tensor and touches it with one ``.item()``
per block (``deepspeed/inference/v2/ragged/blocked_allocator.py:11-105``); here allocate/free are
single C calls (``hds_alloc_*`` in ``include/hds_host.h``) with double-free detection.
"""
from __future__ import annotations

from typing import Iterable, Union

import torch

from ....ops import native


class BlockedAllocator:

    def __init__(self, num_blocks: int) -> None:
        if num_blocks < 1:
            raise ValueError(f"Blocked KV-cache must have at least 1 block, provided {num_blocks}")
        self._num_blocks = num_blocks
        self._lib = native.host_lib()
        self._h = self._lib.hds_alloc_create(num_blocks)
        if not self._h:
            raise RuntimeError("could not create block allocator")

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                self._lib.hds_alloc_destroy(h)
            except Exception:  # interpreter shutdown
                pass

    def allocate(self, num_blocks: int) -> torch.Tensor:
        """Return ``num_blocks`` block ids as an int32 CPU tensor; raises ValueError if not enough are free."""
        if num_blocks < 0:
            raise ValueError("cannot allocate a negative number of blocks")
        out = torch.empty(num_blocks, dtype=torch.int32)
        if num_blocks and self._lib.hds_alloc_allocate(self._h, num_blocks, out.data_ptr()) != 0:
            raise ValueError(f"Not enough free blocks in the KV-cache to allocate {num_blocks} blocks")
        return out

    def free(self, blocks: Union[Iterable[int], int, torch.Tensor]) -> None:
        """Return blocks to the pool; raises ValueError on an invalid or already-free id (nothing is freed)."""
        if isinstance(blocks, int):
            blocks = [blocks]
        t = blocks if isinstance(blocks, torch.Tensor) else torch.tensor(list(blocks), dtype=torch.int32)
        t = t.to(dtype=torch.int32, device="cpu").contiguous()
        if t.numel() and self._lib.hds_alloc_free(self._h, t.data_ptr(), t.numel()) != 0:
            raise ValueError(f"Invalid free: out-of-range or already-free block id in {t.tolist()}")

    @property
    def free_blocks(self) -> int:
        return int(self._lib.hds_alloc_free_blocks(self._h))

    @property
    def total_blocks(self) -> int:
        return self._num_blocks
