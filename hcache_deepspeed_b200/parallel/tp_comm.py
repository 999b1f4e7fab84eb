"""Tensor-parallel reduction of row-parallel GEMM results over NVLink 5 / NVSwitch, fused into the GEMM.

The reference finishes every attention-out / MLP-down GEMM with a separate NCCL all-reduce kernel
(``deepspeed/inference/v2/model_implementations/llama_v2/model.py:159-160, 168-169``). Here the fp32 partial-sum
buffer that the next norm kernel consumes lives in *symmetric memory* (same virtual offset on every rank, peers
mapped into each other's address space, plus an NVLS multicast alias when the switch offers one). The GEMM
epilogue then reduces its tile straight into every rank's buffer:

* ``multimem.red.add.v4.f32`` on the multicast address — one instruction, the switch applies the reduction to
  all GPUs' copies (in-switch reduction, NVLS); or
* ``red.relaxed.sys.add.v4.f32`` to each peer mapping (P2P atomics over NVLink) when no multicast object exists.

What is left of the "collective" is a device-side barrier (flags in symmetric memory, release/acquire at system
scope) between the GEMM and the consuming norm kernel. Two buffers alternate so that a fast rank's next GEMM can
never add into a buffer a slow rank is still consuming.

Symmetric allocation and handle exchange are plumbing and use ``torch.distributed._symmetric_memory``; the data
path is this library's kernels.
"""
from __future__ import annotations

from typing import List, Optional

import torch

from ..models.transformer import TPComm
from ..utils.logging import logger


class FusedTPComm(TPComm):
    """``mode``:

    * ``"pull"`` (default) — every rank accumulates its partial sums locally (split-K atomics stay on-device) and
      the consuming norm kernel adds up all ranks' buffers with plain loads over NVLink. Measured on 2xB200: remote
      fp32 atomics made a 64-token forward ~2x slower than plain peer loads.
    * ``"push"`` — the GEMM epilogue reduces its tile into every rank's buffer (``multimem.red`` on an NVLS multicast
      address when the switch provides one, else ``red.sys`` to each peer mapping): GEMM -> all-reduce in one kernel.
    """
    fused = True

    def __init__(self, group, size: int, rank: int, device: torch.device, mode: str = "pull") -> None:
        super().__init__(group, size, rank)
        self.device = device
        self.mode = mode
        self._symm = None
        self._bufs: List[torch.Tensor] = []
        self._handles = []
        import torch.distributed._symmetric_memory as symm_mem
        self._symm = symm_mem
        self._group_name = group.group_name
        try:
            symm_mem.enable_symm_mem_for_group(self._group_name)
        except Exception:
            pass

    # ---- buffers ---------------------------------------------------------------------------------------
    def alloc_delta(self, n_tokens: int, hidden: int, count: int = 2) -> List[torch.Tensor]:
        """``count`` zeroed fp32 ``[n_tokens, hidden]`` buffers in symmetric memory."""
        out = []
        for _ in range(count):
            t = self._symm.empty((n_tokens, hidden), dtype=torch.float32, device=self.device)
            hdl = self._symm.rendezvous(t, group=self._group_name)
            t.zero_()
            self._bufs.append(t)
            self._handles.append(hdl)
            out.append(t)
        torch.cuda.synchronize(self.device)
        hdl.barrier(channel=0)
        self.use_multicast = bool(getattr(self._handles[0], "multicast_ptr", 0))
        logger.info(f"TP rank {self.rank}: symmetric fp32 reduction buffers ready "
                    f"({'NVLS multicast' if self.use_multicast else 'P2P atomics'})")
        return out

    def alloc_symmetric(self, shape, dtype: torch.dtype) -> torch.Tensor:
        """A zeroed tensor in symmetric memory (peers can address it: see :meth:`peer_ptrs_raw`)."""
        t = self._symm.empty(tuple(shape), dtype=dtype, device=self.device)
        hdl = self._symm.rendezvous(t, group=self._group_name)
        t.zero_()
        self._bufs.append(t)
        self._handles.append(hdl)
        torch.cuda.synchronize(self.device)
        hdl.barrier(channel=0)
        return t

    def peer_ptrs_raw(self, t: torch.Tensor, byte_offset: int = 0) -> List[int]:
        """Address of ``t`` (+offset) in every rank's copy of its symmetric allocation, this rank's mapping."""
        base = t.data_ptr()
        for buf, hdl in zip(self._bufs, self._handles):
            b0 = buf.data_ptr()
            if b0 <= base < b0 + buf.numel() * buf.element_size():
                return [int(p) + (base - b0) + byte_offset for p in hdl.buffer_ptrs]
        raise ValueError("tensor is not a symmetric allocation of this communicator")

    def peer_view(self, t: torch.Tensor, peer: int) -> torch.Tensor:
        """Tensor aliasing ``t``'s storage in rank ``peer``'s copy of the symmetric allocation (same shape)."""
        base = t.data_ptr()
        for buf, hdl in zip(self._bufs, self._handles):
            b0 = buf.data_ptr()
            if b0 <= base < b0 + buf.numel() * buf.element_size():
                off = (base - b0) // buf.element_size()
                return hdl.get_buffer(peer, tuple(t.shape), t.dtype, off)
        raise ValueError("tensor is not a symmetric allocation of this communicator")

    def barrier(self) -> None:
        """Device-side barrier of the TP group on the current stream (flags in symmetric memory)."""
        self._handles[0].barrier(channel=1)

    def _find(self, t: torch.Tensor):
        base = t.data_ptr()
        for buf, hdl in zip(self._bufs, self._handles):
            b0 = buf.data_ptr()
            if b0 <= base < b0 + buf.numel() * buf.element_size():
                return hdl, base - b0
        raise ValueError("tensor is not part of the symmetric reduction buffers")

    def peer_ptrs(self, t: torch.Tensor) -> Optional[List[int]]:
        """Addresses of ``t``'s storage in every rank's buffer (own rank included), for ``red.sys`` epilogues."""
        if getattr(self, "use_multicast", False):
            return None
        hdl, off = self._find(t)
        return [int(p) + off for p in hdl.buffer_ptrs]

    def multicast_ptr(self, t: torch.Tensor) -> Optional[int]:
        if not getattr(self, "use_multicast", False):
            return None
        hdl, off = self._find(t)
        return int(hdl.multicast_ptr) + off

    # ---- the collective that remains ---------------------------------------------------------------------
    def all_reduce_delta(self, delta32: torch.Tensor) -> None:
        """Partial sums were already reduced into every rank's buffer by the GEMM epilogues; wait until all
        ranks' contributions have landed (device-side barrier on the current stream)."""
        hdl, _ = self._find(delta32)
        hdl.barrier(channel=0)


def make_tp_comm(group, size: int, rank: int, device: torch.device, mode: str = "pull") -> TPComm:
    """Fused communicator when symmetric memory works on this system, else the NCCL baseline."""
    try:
        import os
        return FusedTPComm(group, size, rank, device, mode=os.environ.get("HDS_TP_REDUCE", mode))
    except Exception as e:  # no symmetric memory support (driver / topology): keep serving over NCCL
        logger.warning(f"fused tensor-parallel collectives unavailable ({type(e).__name__}: {e}); using NCCL all-reduce")
        return TPComm(group=group, size=size, rank=rank)
