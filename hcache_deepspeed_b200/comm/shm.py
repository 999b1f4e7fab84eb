"""Shared-memory all-reduce for CPU tensors of one node — the low-latency collective of the CPU configuration's tensor
parallelism (C++ engine: ``ops/csrc_host/shm_comm.cpp``). ``inference_all_reduce`` uses it for CPU fp32 / bf16 tensors
once a group has been attached; everything else stays on torch.distributed.

Reference: ``deepspeed/comm/torch.py:160-166`` + ``csrc/cpu/comm/shm_interface.cpp`` (``torch.ops.deepspeed.inference_all_reduce_``)."""
from __future__ import annotations

import os
from typing import Dict, Optional

import torch
import torch.distributed as tdist

from ..ops import native

_DT = {torch.bfloat16: 0, torch.float32: 2}
_groups: Dict[int, "ShmAllReduce"] = {}


class ShmAllReduce:
    """One attached segment. Every rank of ``group`` must construct it (collective) and call :meth:`all_reduce` in
    the same order."""

    def __init__(self, group=None, slot_bytes: int = 1 << 20, name: Optional[str] = None):
        assert tdist.is_available() and tdist.is_initialized(), "attach the shared-memory group after init_distributed()"
        self.rank, self.world = tdist.get_rank(group), tdist.get_world_size(group)
        if name is None:
            # a name nobody else (or a crashed earlier job) can be using: rank 0's pid + a random token
            token = [f"hds_shm_{os.getpid()}_{int.from_bytes(os.urandom(4), 'little')}" if self.rank == 0 else None]
            tdist.broadcast_object_list(token, src=tdist.get_global_rank(group, 0) if group is not None else 0, group=group)
            name = token[0]
        self.name, self.slot_bytes = name, slot_bytes
        self._lib = native.host_lib()
        self._h = self._lib.hds_shm_open(name.encode(), self.rank, self.world, slot_bytes)
        if not self._h:
            raise OSError(f"could not attach shared-memory segment /{name}")

    def all_reduce(self, tensor: torch.Tensor) -> torch.Tensor:
        assert tensor.device.type == "cpu" and tensor.dtype in _DT and tensor.is_contiguous()
        rc = self._lib.hds_shm_allreduce(self._h, tensor.data_ptr(), tensor.numel(), _DT[tensor.dtype])
        if rc != 0:
            raise RuntimeError("shared-memory all-reduce failed")
        return tensor

    def close(self) -> None:
        if self._h:
            tdist.barrier()
            self._lib.hds_shm_close(self._h, 1 if self.rank == 0 else 0)
            self._h = None


def attach(group=None, slot_bytes: int = 1 << 20) -> ShmAllReduce:
    """Collective: set up the shared-memory path for ``group`` (single node only)."""
    key = id(group) if group is not None else 0
    if key not in _groups:
        _groups[key] = ShmAllReduce(group, slot_bytes)
    return _groups[key]


def attached(group=None) -> Optional[ShmAllReduce]:
    return _groups.get(id(group) if group is not None else 0)


def detach(group=None) -> None:
    g = _groups.pop(id(group) if group is not None else 0, None)
    if g is not None:
        g.close()


def usable(tensor: torch.Tensor, group=None) -> bool:
    return tensor.device.type == "cpu" and tensor.dtype in _DT and tensor.is_contiguous() and attached(group) is not None
