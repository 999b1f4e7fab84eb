"""Shared-memory collectives for CPU tensors of one node — the low-latency transport of the CPU configuration's tensor
parallelism (C++ engine: ``ops/csrc_host/shm_comm.cpp``: all-reduce, broadcast, all-gather, reduce-scatter, barrier;
AVX-512 sums where the CPU has them). ``inference_all_reduce`` uses it for CPU tensors once a group has been attached,
``CCLBackend`` (:mod:`.ccl`) routes every collective it covers through it; everything else stays on torch.distributed.

Reference: ``deepspeed/comm/torch.py:160-166`` + ``csrc/cpu/comm/shm_interface.cpp`` (``torch.ops.deepspeed.inference_all_reduce_``)."""
from __future__ import annotations

import os
from typing import Dict, Optional

import torch
import torch.distributed as tdist

from ..ops import native

_DT = {torch.bfloat16: 0, torch.float16: 1, torch.float32: 2, torch.int32: 3, torch.int64: 4}
_OPS = {"sum": 0, "max": 1, "min": 2}
_groups: Dict[int, "ShmAllReduce"] = {}
#: set by ``CCLBackend``: the general collectives of :mod:`.comm` (not only ``inference_all_reduce``) use attached groups
route_all_collectives = False


def op_code(op) -> Optional[int]:
    """Engine code of a reduction (a ``ReduceOp`` of this package or of torch, or its name); None if the engine has none."""
    name = getattr(op, "name", None) or str(op)
    return _OPS.get(name.split(".")[-1].lower())


class ShmAllReduce:
    """One attached segment. Every rank of ``group`` must construct it (collective) and call its collectives in the
    same order. (The class keeps its first name; it now carries broadcast / all-gather / reduce-scatter / barrier too.)"""

    def __init__(self, group=None, slot_bytes: int = 1 << 20, name: Optional[str] = None):
        assert tdist.is_available() and tdist.is_initialized(), "attach the shared-memory group after init_distributed()"
        self.group = group
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

    @staticmethod
    def _check(rc: int, what: str) -> None:
        if rc == -2:
            raise TimeoutError(f"shared-memory {what}: a rank of the group did not arrive (HDS_SHM_TIMEOUT_S)")
        if rc != 0:
            raise RuntimeError(f"shared-memory {what} failed (rc={rc})")

    @staticmethod
    def supports(tensor: torch.Tensor, op="sum") -> bool:
        return tensor.device.type == "cpu" and tensor.dtype in _DT and tensor.is_contiguous() and op_code(op) is not None

    def all_reduce(self, tensor: torch.Tensor, op="sum") -> torch.Tensor:
        assert self.supports(tensor, op)
        self._check(self._lib.hds_shm_allreduce_op(self._h, tensor.data_ptr(), tensor.numel(), _DT[tensor.dtype], op_code(op)),
                    "all-reduce")
        return tensor

    def broadcast(self, tensor: torch.Tensor, src: int) -> torch.Tensor:
        """``src`` is the rank WITHIN the group."""
        assert tensor.device.type == "cpu" and tensor.is_contiguous()
        self._check(self._lib.hds_shm_broadcast(self._h, tensor.data_ptr(), tensor.numel() * tensor.element_size(), src), "broadcast")
        return tensor

    def all_gather_into_tensor(self, out: torch.Tensor, inp: torch.Tensor) -> torch.Tensor:
        assert out.device.type == "cpu" and inp.device.type == "cpu" and out.is_contiguous() and inp.is_contiguous()
        assert out.numel() == self.world * inp.numel() and out.dtype == inp.dtype
        self._check(self._lib.hds_shm_allgather(self._h, inp.data_ptr(), out.data_ptr(), inp.numel() * inp.element_size()), "all-gather")
        return out

    def reduce_scatter_tensor(self, out: torch.Tensor, inp: torch.Tensor, op="sum") -> torch.Tensor:
        assert self.supports(inp, op) and out.is_contiguous() and out.dtype == inp.dtype and inp.numel() == self.world * out.numel()
        self._check(self._lib.hds_shm_reduce_scatter(self._h, inp.data_ptr(), out.data_ptr(), out.numel(), _DT[inp.dtype], op_code(op)),
                    "reduce-scatter")
        return out

    def barrier(self) -> None:
        self._check(self._lib.hds_shm_barrier(self._h), "barrier")

    def info(self) -> Dict[str, int]:
        import ctypes
        buf = (ctypes.c_int64 * 4)()
        self._lib.hds_shm_info(self._h, buf)
        return {"avx512": int(buf[0]), "rounds": int(buf[1]), "allreduce_one_exchange": int(buf[2]),
                "allreduce_shard_per_rank": int(buf[3])}

    def close(self) -> None:
        if self._h:
            tdist.barrier(group=self.group)
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
