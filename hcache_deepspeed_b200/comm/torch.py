"""``TorchBackend`` (``deepspeed/comm/torch.py:96-420``): the ``torch.distributed`` transport as an object. Every
collective of the functional API is available as a method (delegating to :mod:`.comm`, so the comms logger sees them),
plus the capability probes the ZeRO code asks before picking a collective."""
from __future__ import annotations

import os
from typing import Optional

import torch
import torch.distributed as tdist

from . import comm as _c
from .backend import Backend

_DELEGATED = ["broadcast", "all_gather", "all_gather_into_tensor", "all_gather_coalesced", "reduce_scatter_tensor", "reduce_scatter",
              "all_to_all_single", "all_to_all", "send", "recv", "isend", "irecv", "gather", "scatter", "barrier",
              "monitored_barrier", "reduce", "all_reduce", "inference_all_reduce", "all_reduce_coalesced", "get_rank",
              "get_world_size", "get_global_rank", "get_world_group", "get_all_ranks_from_group", "new_group",
              "destroy_process_group", "is_available"]


def get_coalescing_manager(group, device, reqs, async_op: bool = False):
    return tdist._coalescing_manager(group, device=device, async_ops=async_op)


class TorchBackend(Backend):

    def __init__(self, backend: str = "nccl", timeout=None, init_method: Optional[str] = None, rank: int = -1, world_size: int = -1,
                 name: str = "torch"):
        super().__init__(name=name)
        self.using_mpi = False
        if not tdist.is_initialized():
            kw = {"timeout": timeout} if timeout is not None else {}
            tdist.init_process_group(backend, init_method=init_method, rank=rank, world_size=world_size, **kw)
        self.initialized = True
        self.world_rank, self.world_size = tdist.get_rank(), tdist.get_world_size()
        self.world_group = tdist.group.WORLD
        self.backend_name = tdist.get_backend()

    # ---- capability probes -----------------------------------------------------------------------------------------
    @classmethod
    def has_all_gather_into_tensor(cls) -> bool:
        return hasattr(tdist, "all_gather_into_tensor")

    @classmethod
    def has_reduce_scatter_tensor(cls) -> bool:
        return hasattr(tdist, "reduce_scatter_tensor")

    @classmethod
    def get_all_gather_function(cls):
        return getattr(tdist, "all_gather_into_tensor", None) or getattr(tdist, "_all_gather_base", None)

    @classmethod
    def get_reduce_scatter_function(cls):
        return getattr(tdist, "reduce_scatter_tensor", None) or getattr(tdist, "_reduce_scatter_base", None)

    def all_gather_base(self, output_tensor, input_tensor, group=None, async_op=False):
        """Older name of ``all_gather_into_tensor``."""
        return self.get_all_gather_function()(output_tensor, input_tensor, group=group, async_op=async_op)

    def broadcast_object_list(self, object_list, src, group=None, device=None):
        return tdist.broadcast_object_list(object_list=object_list, src=src, group=group, device=device)

    def has_all_reduce_coalesced(self) -> bool:
        return hasattr(tdist, "all_reduce_coalesced")

    def has_coalescing_manager(self) -> bool:
        return hasattr(tdist, "_coalescing_manager")

    def is_initialized(self) -> bool:
        return tdist.is_initialized()

    def get_backend(self, group=None) -> str:
        return tdist.get_backend(group)

    def get_local_rank(self) -> int:
        return int(os.environ.get("LOCAL_RANK", 0))

    def init_device_mesh(self, mesh_shape, mesh_dim_names):
        from torch.distributed.device_mesh import init_device_mesh
        return init_device_mesh("cuda" if torch.cuda.is_available() else "cpu", tuple(mesh_shape), mesh_dim_names=tuple(mesh_dim_names))

    def enable_symm_mem_for_group(self, group_name: str):
        return _c.enable_symm_mem_for_group(group_name)


# ---- switches that turn whole families of collectives into no-ops on this backend object (ref: comm/torch.py:20-110): a
#      way to measure how much of a step is communication. A switched-off call returns ``Noop``, whose ``wait()`` returns.
DS_COMM_ALL_GATHER_OFF = False
DS_COMM_REDUCE_SCATTER_OFF = False
DS_COMM_BROADCAST_OFF = False
DS_COMM_ALL_REDUCE_OFF = False
DS_COMM_REDUCE_OFF = False


class Noop:

    def wait(self, *args, **kwargs):
        return None

    def is_completed(self) -> bool:
        return True


def all_gather_comm_off(flag: bool = False) -> None:
    global DS_COMM_ALL_GATHER_OFF
    DS_COMM_ALL_GATHER_OFF = flag


def reduce_scatter_comm_off(flag: bool = False) -> None:
    global DS_COMM_REDUCE_SCATTER_OFF
    DS_COMM_REDUCE_SCATTER_OFF = flag


def broadcast_comm_off(flag: bool = False) -> None:
    global DS_COMM_BROADCAST_OFF
    DS_COMM_BROADCAST_OFF = flag


def all_reduce_comm_off(flag: bool = False) -> None:
    global DS_COMM_ALL_REDUCE_OFF
    DS_COMM_ALL_REDUCE_OFF = flag


def reduce_comm_off(flag: bool = False) -> None:
    global DS_COMM_REDUCE_OFF
    DS_COMM_REDUCE_OFF = flag


def backward_comm_off(flag: bool = False) -> None:
    """The two collectives of a data-parallel backward pass (gradient reduce-scatter / parameter all-gather) off or on."""
    all_gather_comm_off(flag)
    reduce_scatter_comm_off(flag)


def _switched_off(name: str) -> bool:
    if name.startswith("all_gather"):
        return DS_COMM_ALL_GATHER_OFF
    if name.startswith("reduce_scatter"):
        return DS_COMM_REDUCE_SCATTER_OFF
    if name == "broadcast":
        return DS_COMM_BROADCAST_OFF
    if name in ("all_reduce", "inference_all_reduce", "all_reduce_coalesced"):
        return DS_COMM_ALL_REDUCE_OFF
    if name == "reduce":
        return DS_COMM_REDUCE_OFF
    return False


def disable_compiler_collective(func):
    """No tracing compiler runs over this package's collectives: the decorator is the identity."""
    return func


def has_coalescing_manager() -> bool:
    return hasattr(tdist, "distributed_c10d") and hasattr(tdist.distributed_c10d, "_coalescing_manager")


def has_all_reduce_coalesced() -> bool:
    return hasattr(tdist, "all_reduce_coalesced")


def build_shm_op():
    """The host collectives engine (C++ over POSIX shared memory) as a loaded library, or ``None`` when it cannot be
    built here."""
    try:
        from ..ops.op_builder import ALL_OPS
        b = ALL_OPS.get("shm_comm") or ALL_OPS.get("deepspeed_shm_comm")
        return b.load() if b is not None and not b.compatibility_problems() else None
    except Exception:
        return None


def _delegate(name):
    fn = getattr(_c, name)

    def method(self, *a, **k):
        if _switched_off(name):
            if int(os.environ.get("RANK", "0")) == 0:
                from ..utils.logging import logger
                logger.warning(f"{name} is OFF")
            return Noop()
        return fn(*a, **k)

    method.__name__ = name
    method.__doc__ = fn.__doc__
    return method


for _n in _DELEGATED:
    setattr(TorchBackend, _n, _delegate(_n))
