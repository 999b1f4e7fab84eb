"""NVMe tier for optimizer state (ZeRO-Infinity): the fp32 master tensor of a sub-group and every optimizer-state tensor
that belongs to it (Adam moments, ...) live in files between optimizer steps and are staged through pinned buffers.

Reference: ``deepspeed/runtime/swap_tensor/optimizer_utils.py:27-520`` (``OptimizerStateSwapInfo``, ``OptimizerSwapper``).
The reference tracks swap state in per-parameter info objects with parallel lists; here one :class:`_SwapRecord` per
master tensor holds ``name -> (path, numel)`` and the buffers currently lent to it, and the buffer pool falls back to
plain host allocations when every pinned buffer is taken (so ``state_dict()`` can bring everything in at once)."""
from __future__ import annotations

import os
import shutil
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import torch

from ...io.aio import aio_handle as AioHandle
from .utils import AIO_ALIGNED_BYTES, MIN_AIO_BYTES, _pinned

MASTER = "__master__"


@dataclass
class _SwapRecord:
    key: int
    numel: int
    files: Dict[str, str] = field(default_factory=dict)          # tensor name -> path (exists once written)
    buffers: Dict[str, torch.Tensor] = field(default_factory=dict)   # name -> staging buffer while swapped in
    resident: bool = True
    inflight: bool = False        # a read was issued and not yet waited for


class OptimizerStateSwapInfo(_SwapRecord):
    """Name kept for code written against the reference."""


class OptimizerSwapper:

    def __init__(self, swap_config=None, aio_config: Optional[dict] = None, base_folder: Optional[str] = None, optimizer=None,
                 largest_numel: int = 0, device="cpu", dtype=torch.float32, timers=None, buffer_count: Optional[int] = None):
        aio_config = {"block_size": 1048576, "queue_depth": 8, "intra_op_parallelism": 1, "single_submit": False,
                      "overlap_events": True, **(aio_config or {})}
        base_folder = base_folder or str(getattr(swap_config, "nvme_path", None) or ".")
        self.optimizer = getattr(optimizer, "optimizer", optimizer)
        self.dtype, self.device, self.timers = dtype, torch.device(device), timers
        self.element_size = torch.tensor([], dtype=dtype).element_size()
        rank = int(os.environ.get("RANK", 0))
        self.swap_folder = os.path.join(base_folder, "zero_stage_3", "optimizer", f"rank{rank}")
        shutil.rmtree(self.swap_folder, ignore_errors=True)
        os.makedirs(self.swap_folder, exist_ok=True)
        self.min_aio_bytes = max(MIN_AIO_BYTES, aio_config["block_size"])
        self.numel_alignment = AIO_ALIGNED_BYTES * aio_config["intra_op_parallelism"] // self.element_size
        self.largest_numel = self._aligned(largest_numel)
        count = buffer_count if buffer_count is not None else int(getattr(swap_config, "buffer_count", 4) or 4)
        self._free: List[torch.Tensor] = [_pinned(self.largest_numel, dtype) for _ in range(count)] if self.largest_numel else []
        self._pool_ids = {b.data_ptr() for b in self._free}
        self.records: Dict[int, _SwapRecord] = {}
        mk = lambda: AioHandle(aio_config["block_size"], aio_config["queue_depth"], aio_config["single_submit"],  # noqa: E731
                               aio_config["overlap_events"], aio_config["intra_op_parallelism"])
        self.read_handle, self.write_handle = mk(), mk()
        self._pending_write: List[_SwapRecord] = []
        self.stats = {"bytes_read": 0, "bytes_written": 0, "pool_misses": 0}

    # ---- helpers ---------------------------------------------------------------------------------------------------
    def _aligned(self, numel: int) -> int:
        a = max(self.numel_alignment, 1)
        return (numel + a - 1) // a * a

    def swappable_tensor(self, param=None, numel: Optional[int] = None) -> bool:
        numel = numel if numel is not None else param.numel()
        return self.min_aio_bytes <= numel * self.element_size

    def _record(self, param: torch.Tensor) -> _SwapRecord:
        rec = self.records.get(id(param))
        if rec is None:
            rec = self.records[id(param)] = OptimizerStateSwapInfo(key=len(self.records), numel=param.numel())
        return rec

    def _path(self, rec: _SwapRecord, name: str) -> str:
        return os.path.join(self.swap_folder, f"group{rec.key}_{name.strip('_')}.tensor.swp")

    def _take(self, numel: int) -> torch.Tensor:
        need = self._aligned(numel)
        for i, b in enumerate(self._free):
            if b.numel() >= need:
                return self._free.pop(i)
        self.stats["pool_misses"] += 1
        return torch.empty(need, dtype=self.dtype)

    def _give(self, buf: torch.Tensor) -> None:
        if buf.data_ptr() in self._pool_ids:
            self._free.append(buf)

    def _state_tensors(self, param: torch.Tensor) -> Dict[str, torch.Tensor]:
        """The optimizer-state tensors that are as long as the parameter (moments); scalars like ``step`` stay put."""
        st = self.optimizer.state.get(param, {}) if self.optimizer is not None else {}
        return {k: v for k, v in st.items() if torch.is_tensor(v) and v.numel() == param.numel() and v.numel() > 1}

    # ---- out -------------------------------------------------------------------------------------------------------
    def _issue_writes(self, param: torch.Tensor, rec: _SwapRecord) -> None:
        tensors = {MASTER: param, **self._state_tensors(param)}
        for name, t in tensors.items():
            buf = rec.buffers.get(name)
            if buf is None or buf.data_ptr() != t.data_ptr():
                buf = rec.buffers[name] = self._take(rec.numel) if buf is None else buf
                buf.narrow(0, 0, rec.numel).copy_(t.detach().reshape(-1))
            rec.files[name] = self._path(rec, name)
            self.write_handle.async_pwrite(buf, rec.files[name])
            self.stats["bytes_written"] += buf.numel() * self.element_size
        empty = torch.empty(0, dtype=self.dtype)
        for name, t in tensors.items():
            t.data = empty
        rec.resident = False

    def _finish_writes(self) -> None:
        if not self._pending_write:
            return
        self.write_handle.wait()
        for rec in self._pending_write:
            for buf in rec.buffers.values():
                self._give(buf)
            rec.buffers = {}
        self._pending_write = []

    def swap_out_optimizer_state(self, parameter: torch.Tensor, async_swap: bool = False) -> None:
        rec = self._record(parameter)
        if not rec.resident:
            return
        self._finish_writes()            # at most one sub-group's writes are in flight
        self._issue_writes(parameter, rec)
        self._pending_write.append(rec)
        if not async_swap:
            self._finish_writes()

    def initialize_parameters(self, parameters: List[torch.Tensor], src_tensors: List[torch.Tensor]) -> None:
        """Write the initial fp32 values to their files; the parameters end up swapped out."""
        for p, src in zip(parameters, src_tensors):
            if p.numel() == 0 and src.numel():
                p.data = src.detach().to(self.dtype).reshape(-1).clone()
            self._record(p).numel = p.numel()
            self.swap_out_optimizer_state(p)

    # ---- in --------------------------------------------------------------------------------------------------------
    def _issue_reads(self, rec: _SwapRecord) -> None:
        for name, path in rec.files.items():
            buf = rec.buffers[name] = self._take(rec.numel)
            self.read_handle.async_pread(buf, path)
            self.stats["bytes_read"] += buf.numel() * self.element_size
        rec.inflight = True

    def _finish_reads(self, parameter: torch.Tensor, rec: _SwapRecord) -> None:
        self.read_handle.wait()
        rec.inflight = False
        state = self.optimizer.state.get(parameter, {}) if self.optimizer is not None else {}
        for name, buf in rec.buffers.items():
            view = buf.narrow(0, 0, rec.numel)
            if name == MASTER:
                parameter.data = view
            elif name in state:
                state[name].data = view
        rec.resident = True

    def swap_in_optimizer_state(self, parameter: torch.Tensor, async_parameter: Optional[torch.Tensor] = None) -> None:
        """Make ``parameter`` and its optimizer state resident; with ``async_parameter`` also START reading that one."""
        rec = self._record(parameter)
        if not rec.resident:
            if rec in self._pending_write:
                self._finish_writes()    # its own files are still being written
            if not rec.inflight:
                self._issue_reads(rec)
            self._finish_reads(parameter, rec)
        if async_parameter is not None:
            nxt = self._record(async_parameter)
            if not nxt.resident and not nxt.inflight and nxt.files:
                if nxt in self._pending_write:
                    self._finish_writes()
                self._issue_reads(nxt)

    # ---- bulk ------------------------------------------------------------------------------------------------------
    def swap_in_all(self, parameters: List[torch.Tensor]) -> None:
        for p in parameters:
            self.swap_in_optimizer_state(p)

    def swap_out_all(self, parameters: List[torch.Tensor]) -> None:
        for p in parameters:
            self.swap_out_optimizer_state(p)

    def purge_state(self) -> None:
        self._finish_writes()
        for rec in self.records.values():
            for buf in rec.buffers.values():
                self._give(buf)
            rec.buffers = {}
