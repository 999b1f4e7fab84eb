"""Fire-and-forget tensor swap-out through a pool of pinned buffers (gradients / optimizer state to NVMe).
Reference: ``deepspeed/runtime/swap_tensor/async_swapper.py:18-174``."""
from __future__ import annotations

from typing import List

import torch

from .utils import SwapBufferManager, SwapBufferPool


class AsyncTensorSwapper:

    def __init__(self, aio_handle, numel_alignment: int, timers=None):
        self.aio_handle, self.numel_alignment = aio_handle, numel_alignment
        self.free_buffer_index, self.swapping_buffer_index, self.ready_buffer_index = [], [], []
        self.all_buffers: List = []
        self.current_buffer_index = -1
        self.max_numel = 0
        self.num_pending_swaps = 0
        self.num_elements_swapped = 0
        self.dtype = None

    def has_buffers(self) -> bool:
        return len(self.all_buffers) > 0

    def add_buffers(self, buffer_list: List[torch.Tensor]):
        assert not self.all_buffers and all(b.device.type == "cpu" for b in buffer_list)
        self.dtype = buffer_list[0].dtype
        self.all_buffers = [SwapBufferPool([b]) for b in buffer_list]
        self.free_buffer_index = list(range(len(self.all_buffers)))
        self.max_numel = max(b.numel() for b in buffer_list)

    def get_timer_names(self):
        return []

    def release_buffers(self):
        self._flush_buffers_until_complete()
        out = [p.buffers[0].buffer for p in self.all_buffers]
        self.all_buffers, self.free_buffer_index, self.current_buffer_index = [], [], -1
        return out

    def _aligned(self, numel: int) -> int:
        a = self.numel_alignment
        return (numel + a - 1) // a * a

    def swap_out_tensors(self, tensor_list: List[torch.Tensor], path_list: List[str]):
        for t, p in zip(tensor_list, path_list):
            self._swap_out_tensor(t, p)

    def _swap_out_tensor(self, tensor: torch.Tensor, swap_path: str):
        assert self.all_buffers, "add_buffers() first"
        aligned = self._aligned(tensor.numel())
        assert aligned <= self.max_numel
        if self.current_buffer_index < 0 or not self.all_buffers[self.current_buffer_index].has_space(aligned):
            self._make_swap_space()
        self.all_buffers[self.current_buffer_index].insert_tensor(tensor, swap_path, aligned)
        self.num_elements_swapped += tensor.numel()

    def _make_swap_space(self):
        if self.current_buffer_index >= 0:
            self._flush_current()
        if not self.free_buffer_index:
            self._wait_for_swap_complete()
        self.current_buffer_index = self.free_buffer_index.pop()

    def _flush_current(self):
        pool = self.all_buffers[self.current_buffer_index]
        pool.swap_out(self.aio_handle, async_op=True)
        self.swapping_buffer_index.append(self.current_buffer_index)
        self.num_pending_swaps += len(pool.get_swap_tensors())
        self.current_buffer_index = -1

    def _wait_for_swap_complete(self):
        self.aio_handle.wait()
        self.num_pending_swaps = 0
        for i in self.swapping_buffer_index:
            self.all_buffers[i].reset()
            self.free_buffer_index.append(i)
        self.swapping_buffer_index = []

    def _flush_buffers_until_complete(self):
        if self.current_buffer_index >= 0:
            self._flush_current()
        self._wait_for_swap_complete()
