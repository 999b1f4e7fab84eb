"""Asynchronous file I/O handle over the native thread-pool engine (``libhds_host.so``).

API of the reference's ``aio_handle`` (``csrc/aio/py_lib/py_ds_aio.cpp:15-116``,
``deepspeed_py_io_handle.cpp``): ``sync_pread / sync_pwrite / async_pread / async_pwrite / wait`` on tensors,
plus ``get_block_size / get_queue_depth / get_intra_op_parallelism`` and pinned-tensor helpers. The engine is
pread/pwrite worker threads with optional O_DIRECT instead of libaio (its headers are absent from the target
image; io_uring would need a newer kernel guarantee) — same semantics: a request is split into ``block_size``
pieces spread over ``intra_op_parallelism`` threads.
"""
from __future__ import annotations

import os
from typing import Dict, Optional

import torch

from ..ops import native
from ..utils.memory import pinned_empty


class AsyncIOHandle:

    def __init__(self, block_size: int = 1 << 20, queue_depth: int = 32, single_submit: bool = False,
                 overlap_events: bool = True, intra_op_parallelism: int = 4, use_direct: bool = False) -> None:
        self._lib = native.host_lib()
        self._block_size, self._queue_depth = block_size, queue_depth
        self._single_submit, self._overlap_events = single_submit, overlap_events
        self._threads = max(1, intra_op_parallelism)
        self._h = self._lib.hds_aio_create(block_size, queue_depth, self._threads, 1 if use_direct else 0)
        self._inflight: Dict[int, torch.Tensor] = {}  # keeps buffers alive until their request completes
        self._pinned = []

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                self._lib.hds_aio_destroy(h)
            except Exception:
                pass

    # ---- reference getters ---------------------------------------------------------------------------
    def get_block_size(self) -> int:
        return self._block_size

    def get_queue_depth(self) -> int:
        return self._queue_depth

    def get_single_submit(self) -> bool:
        return self._single_submit

    def get_overlap_events(self) -> bool:
        return self._overlap_events

    def get_intra_op_parallelism(self) -> int:
        return self._threads

    def get_alignment(self) -> int:
        return 4096

    # ---- I/O -------------------------------------------------------------------------------------------
    @staticmethod
    def _check(buffer: torch.Tensor) -> int:
        if buffer.device.type != "cpu" or not buffer.is_contiguous():
            raise ValueError("aio buffers must be contiguous host tensors")
        return buffer.numel() * buffer.element_size()

    def async_pread(self, buffer: torch.Tensor, filename: str, file_offset: int = 0, nbytes: Optional[int] = None) -> int:
        n = self._check(buffer) if nbytes is None else nbytes
        rid = self._lib.hds_aio_pread(self._h, buffer.data_ptr(), os.fsencode(filename), n, file_offset)
        if rid < 0:
            raise OSError(f"cannot open {filename} for reading")
        self._inflight[rid] = buffer
        return rid

    def async_pwrite(self, buffer: torch.Tensor, filename: str, file_offset: int = 0, nbytes: Optional[int] = None) -> int:
        n = self._check(buffer) if nbytes is None else nbytes
        rid = self._lib.hds_aio_pwrite(self._h, buffer.data_ptr(), os.fsencode(filename), n, file_offset)
        if rid < 0:
            raise OSError(f"cannot open {filename} for writing")
        self._inflight[rid] = buffer
        return rid

    def wait(self, request_id: int = 0) -> int:
        """Wait for one request (or all when 0). Returns bytes moved; raises OSError on I/O failure."""
        got = self._lib.hds_aio_wait(self._h, request_id)
        if request_id == 0:
            self._inflight.clear()
        else:
            self._inflight.pop(request_id, None)
        if got < 0:
            raise OSError(-got, os.strerror(-got))
        return int(got)

    def poll(self, request_id: int) -> bool:
        return self._lib.hds_aio_poll(self._h, request_id) == 1

    def stats(self) -> dict:
        """How the pieces were executed so far: through kernel AIO (``io_submit``, true device queue depth) or
        synchronously (``pread`` / ``pwrite``: unaligned pieces, no O_DIRECT on this filesystem), and the deepest
        in-flight queue one worker reached."""
        import ctypes
        out = (ctypes.c_int64 * 3)()
        self._lib.hds_aio_stats(self._h, out)
        return {"kernel_aio_pieces": int(out[0]), "sync_pieces": int(out[1]), "max_inflight_per_worker": int(out[2])}

    def pending(self) -> int:
        return int(self._lib.hds_aio_pending(self._h))

    def sync_pread(self, buffer: torch.Tensor, filename: str, file_offset: int = 0) -> int:
        return self.wait(self.async_pread(buffer, filename, file_offset))

    def sync_pwrite(self, buffer: torch.Tensor, filename: str, file_offset: int = 0) -> int:
        return self.wait(self.async_pwrite(buffer, filename, file_offset))

    # reference spellings
    read = sync_pread
    write = sync_pwrite
    pread = sync_pread
    pwrite = sync_pwrite

    # ---- pinned buffers (ref: new_cpu_locked_tensor / free_cpu_locked_tensor) --------------------------------
    def new_cpu_locked_tensor(self, num_elem: int, example: torch.Tensor) -> torch.Tensor:
        t = pinned_empty((num_elem, ), example.dtype)
        self._pinned.append(t)
        return t

    def free_cpu_locked_tensor(self, tensor: torch.Tensor) -> bool:
        for i, t in enumerate(self._pinned):
            if t.data_ptr() == tensor.data_ptr():
                self._pinned.pop(i)
                return True
        return False


def aio_handle(block_size: int = 1 << 20, queue_depth: int = 32, single_submit: bool = False, overlap_events: bool = True,
               intra_op_parallelism: int = 4, use_direct: bool = False) -> AsyncIOHandle:
    """Factory with the reference's positional signature (``AsyncIOBuilder().load().aio_handle(...)``)."""
    return AsyncIOHandle(block_size, queue_depth, single_submit, overlap_events, intra_op_parallelism, use_direct)
