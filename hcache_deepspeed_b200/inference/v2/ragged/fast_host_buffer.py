"""Host staging memory allocated by the native library with explicit flags (``hds_host_buffer_alloc``) and viewed as a
torch tensor. Role of the reference's ``allocate_fast_host_buffer`` (``ragged/csrc/fast_host_buffer.cu:9-18``,
``ragged_ops.cpp``): pinned + mapped + write-combined memory for the ragged batch's metadata.

The ragged batch wrapper uses plain pinned torch memory by default, because its C++ batch builder and the CPU back end
READ the staging area back (write-combined memory makes those reads uncached); set ``HDS_FAST_HOST_BUFFER=wc`` (or
``mapped``) to stage through one of these buffers instead."""
from __future__ import annotations

import ctypes
import os
from typing import Optional

import torch

from ....ops import native


class FastHostBuffer:

    def __init__(self, numel: int, dtype: torch.dtype = torch.int32, mapped: bool = True, write_combined: bool = True):
        self.numel, self.dtype = int(numel), dtype
        self.nbytes = self.numel * torch.empty((), dtype=dtype).element_size()
        self.mapped, self.write_combined = bool(mapped), bool(write_combined)
        self._lib = native.cuda_lib()
        ptr = ctypes.c_void_p()
        native.check(self._lib.hds_host_buffer_alloc(self.nbytes, int(mapped), int(write_combined), ctypes.byref(ptr)),
                     "hds_host_buffer_alloc")
        self.ptr = ptr.value
        raw = (ctypes.c_char * self.nbytes).from_address(self.ptr)
        self.tensor = torch.frombuffer(raw, dtype=dtype, count=self.numel)
        self.tensor.zero_()

    def device_ptr(self) -> int:
        """Address kernels can use to read the buffer in place (mapped buffers only)."""
        assert self.mapped, "the buffer was not allocated with mapped=True"
        out = ctypes.c_void_p()
        native.check(self._lib.hds_host_buffer_device_ptr(ctypes.c_void_p(self.ptr), ctypes.byref(out)), "hds_host_buffer_device_ptr")
        return out.value

    def free(self) -> None:
        if getattr(self, "ptr", None):
            self.tensor = None
            self._lib.hds_host_buffer_free(ctypes.c_void_p(self.ptr))
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def staging_mode() -> Optional[str]:
    """``HDS_FAST_HOST_BUFFER``: unset / ``0`` = pinned torch memory, ``mapped`` = pinned + mapped, ``wc`` = pinned +
    mapped + write-combined (the reference's flag set)."""
    v = os.environ.get("HDS_FAST_HOST_BUFFER", "0").lower()
    return v if v in ("mapped", "wc") else None


def allocate_fast_host_buffer(like: torch.Tensor) -> torch.Tensor:
    """Reference-compatible helper: a pinned + mapped + write-combined host tensor with the shape / dtype of ``like``
    (the returned tensor keeps its allocation alive)."""
    buf = FastHostBuffer(like.numel(), like.dtype)
    t = buf.tensor.view(like.shape)
    t._hds_fast_host_buffer = buf
    return t
