"""Pinned swap buffers carved into aligned slots. Reference: ``deepspeed/runtime/swap_tensor/utils.py:37-240``."""
from __future__ import annotations

from typing import Dict, List, Optional

import torch

MIN_AIO_BYTES = 1024 ** 2
AIO_ALIGNED_BYTES = 1024


def _pinned(numel: int, dtype) -> torch.Tensor:
    t = torch.empty(numel, dtype=dtype)
    return t.pin_memory() if torch.cuda.is_available() else t


def get_sized_buffer(buffer: torch.Tensor, num_elems: int) -> torch.Tensor:
    assert num_elems <= buffer.numel(), f"num_elems {num_elems} > buffer {buffer.numel()}"
    return buffer.narrow(0, 0, num_elems) if num_elems < buffer.numel() else buffer


def get_sized_buffers(buffer_list, num_elems_list):
    return [get_sized_buffer(b, n) for b, n in zip(buffer_list, num_elems_list)]


class SwapBuffer:
    def __init__(self, buffer: torch.Tensor):
        self.buffer = buffer
        self.reset()

    def reset(self):
        self.offset = 0
        self.swap_tensors: Dict[int, torch.Tensor] = {}
        self.compute_tensors: Dict[int, torch.Tensor] = {}
        self.swap_paths: Dict[int, str] = {}
        self.num_elem = 0

    def insert_tensor(self, tensor: torch.Tensor, swap_path: str, aligned_numel: int):
        swap_t, compute_t = self.allocate_tensor(swap_path, tensor.numel(), aligned_numel)
        compute_t.data.copy_(tensor.data.reshape(-1))
        return swap_t, compute_t

    def allocate_tensor(self, swap_path: str, numel: int, aligned_numel: int):
        assert self.has_space(aligned_numel)
        idx = len(self.swap_tensors)
        swap_t = self.buffer.narrow(0, self.offset, aligned_numel)
        self.swap_tensors[idx] = swap_t
        self.compute_tensors[idx] = swap_t.narrow(0, 0, numel)
        self.swap_paths[idx] = swap_path
        self.offset += aligned_numel
        self.num_elem += numel
        return swap_t, self.compute_tensors[idx]

    def has_space(self, numel: int) -> bool:
        return self.offset + numel <= self.buffer.numel()

    def get_swap_tensors(self):
        return list(self.swap_tensors.values())

    def get_swap_paths(self):
        return list(self.swap_paths.values())

    def get_compute_tensors(self):
        return list(self.compute_tensors.values())

    def get_num_elem(self):
        return self.num_elem


class SwapBufferPool:
    def __init__(self, buffers: List[torch.Tensor]):
        self.buffers = [SwapBuffer(b) for b in buffers]
        self.current_index = 0

    def reset(self):
        self.current_index = 0
        for b in self.buffers:
            b.reset()

    def allocate_tensor(self, numel, swap_path, aligned_numel):
        if self.has_space(aligned_numel):
            return self._current().allocate_tensor(swap_path, numel, aligned_numel)
        return None, None

    def insert_tensor(self, tensor, swap_path, aligned_numel):
        if self.has_space(aligned_numel):
            return self._current().insert_tensor(tensor, swap_path, aligned_numel)
        return None, None

    def has_space(self, numel) -> bool:
        if self._current().has_space(numel):
            return True
        if self.current_index == len(self.buffers) - 1:
            return False
        self.current_index += 1
        return self._current().has_space(numel)

    def _current(self) -> SwapBuffer:
        return self.buffers[self.current_index]

    def get_swap_tensors(self):
        return [t for b in self.buffers for t in b.get_swap_tensors()]

    def get_swap_paths(self):
        return [p for b in self.buffers for p in b.get_swap_paths()]

    def get_compute_tensors(self):
        return [t for b in self.buffers for t in b.get_compute_tensors()]

    def swap_out(self, aio_handle, async_op: bool = False):
        for t, p in zip(self.get_swap_tensors(), self.get_swap_paths()):
            aio_handle.async_pwrite(t, p)
        if not async_op:
            aio_handle.wait()

    def swap_in(self, aio_handle, async_op: bool = False):
        for t, p in zip(self.get_swap_tensors(), self.get_swap_paths()):
            aio_handle.async_pread(t, p)
        if not async_op:
            aio_handle.wait()


class SwapBufferManager:
    """``count`` pinned buffers of ``num_elems`` each, handed out and returned by id."""

    def __init__(self, num_elems: int, count: int, dtype):
        self.num_elems, self.count, self.dtype = num_elems, count, dtype
        self.all_buffers = [_pinned(num_elems, dtype) for _ in range(count)]
        self.free_buffer_index = list(range(count))
        self.used_buffer_index: Dict[int, int] = {}

    def allocate(self, num_elems: int, count: int, dtype) -> Optional[List[torch.Tensor]]:
        assert dtype == self.dtype and num_elems <= self.num_elems
        if count > len(self.free_buffer_index):
            return None
        idxs, self.free_buffer_index = self.free_buffer_index[-count:], self.free_buffer_index[:-count]
        out = []
        for i in idxs:
            b = self.all_buffers[i].narrow(0, 0, num_elems)
            self.used_buffer_index[id(b)] = i
            out.append(b)
        return out

    def allocate_all(self, num_elems: int, dtype):
        return self.allocate(num_elems, len(self.free_buffer_index), dtype)

    def free(self, buffers: List[torch.Tensor]):
        for b in buffers:
            i = self.used_buffer_index.pop(id(b))
            self.free_buffer_index.append(i)


def swap_in_tensors(swap_handle, tensor_buffers, swap_paths) -> None:
    """Queue one read per (buffer, file) on an AIO handle; the caller waits on the handle."""
    for buf, path in zip(tensor_buffers, swap_paths):
        rc = swap_handle.async_pread(buf, path)
        assert rc in (0, None), f"async_pread({path}) failed: {rc}"


def swap_out_tensors(swap_handle, tensor_buffers, swap_paths) -> None:
    for buf, path in zip(tensor_buffers, swap_paths):
        rc = swap_handle.async_pwrite(buf, path)
        assert rc in (0, None), f"async_pwrite({path}) failed: {rc}"


def print_object(obj, name: str, exclude_list=()) -> None:
    from ...utils.logging import logger
    logger.info(f"{name}:")
    for arg in sorted(vars(obj)):
        if arg not in exclude_list:
            logger.info(f"  {arg} {'.' * max(1, 29 - len(arg))} {getattr(obj, arg)}")
