"""``CCLBackend`` (``deepspeed/comm/ccl.py:35-194``): the host-side collectives back end.

The reference binds Intel oneCCL plus a shared-memory all-reduce for the ranks of one node. This build has no oneCCL; the
same role — low-latency collectives between the CPU ranks of one node, with ``torch.distributed`` as the fallback for
everything else — is filled by this framework's own C++ engine (``ops/csrc_host/shm_comm.cpp``: all-reduce,
broadcast, all-gather, reduce-scatter and barrier over one POSIX shared-memory segment, AVX-512 sums). ``gloo`` carries
rendezvous, group creation, point-to-point and any call the engine does not cover (other dtypes / ops, non-contiguous
tensors, groups that span nodes).

    dist.init_distributed(dist_backend="ccl")          # or CCLBackend() directly
    dist.all_reduce(cpu_tensor)                         # -> shared memory
"""
from __future__ import annotations

import socket
from collections import Counter
from typing import Dict, List, Optional

import torch.distributed as tdist

from . import comm as _c
from . import shm
from .torch import TorchBackend

#: collectives the native engine implements (``run_collective`` names)
NATIVE_COLLECTIVES = ("all_reduce", "inference_all_reduce", "broadcast", "all_gather", "all_gather_into_tensor",
                      "reduce_scatter_tensor", "barrier")


class CCLBackend(TorchBackend):

    def __init__(self, name: str = "ccl", rank: int = -1, world_size: int = -1, mpu=None, timeout=None, init_method=None,
                 slot_bytes: int = 4 << 20):
        super().__init__(backend="gloo", timeout=timeout, init_method=init_method, rank=rank, world_size=world_size, name=name)
        self.mpu = mpu
        self.slot_bytes = slot_bytes
        self.calls: Counter = Counter()          # "<collective>.native" / "<collective>.torch"
        self._hosts = self._gather_hosts(None)
        self.single_node = len(set(self._hosts)) == 1
        shm.route_all_collectives = True
        if self.single_node and self.world_size > 1:
            shm.attach(None, slot_bytes)

    # ---- groups ----------------------------------------------------------------------------------------------------
    @staticmethod
    def _gather_hosts(group) -> List[str]:
        hosts: List[Optional[str]] = [None] * tdist.get_world_size(group)
        tdist.all_gather_object(hosts, socket.gethostname(), group=group)
        return hosts

    def new_group(self, ranks=None, **kwargs):
        """Collective over ALL ranks (like ``torch.distributed.new_group``); the members whose ranks share a node also
        attach a shared-memory segment for the group."""
        group = tdist.new_group(ranks=ranks, **kwargs)
        members = list(range(self.world_size)) if ranks is None else list(ranks)
        self.process_groups.append(group)
        if self.world_rank in members and len(members) > 1 and len({self._hosts[r] for r in members}) == 1:
            shm.attach(group, self.slot_bytes)
        return group

    def get_all_ranks_from_group(self, group=None) -> List[int]:
        return _c.get_all_ranks_from_group(group)

    def native_engine(self, group=None):
        return shm.attached(group)

    # ---- collectives ---------------------------------------------------------------------------------------------------
    def run_collective(self, name: str, **kwargs):
        """Dispatch by name (the reference's ``CCLBackend.run_collective``): the native engine when it covers the call,
        ``torch.distributed`` otherwise. ``self.calls`` records which one ran."""
        if name not in NATIVE_COLLECTIVES:
            self.calls[f"{name}.torch"] += 1
            return getattr(_c, name)(**kwargs)
        before = self._rounds(kwargs.get("group"))
        out = getattr(_c, name)(**kwargs)
        self.calls[f"{name}.{'native' if self._rounds(kwargs.get('group')) > before else 'torch'}"] += 1
        return out

    def _rounds(self, group) -> int:
        eng = shm.attached(group)
        return eng.info()["rounds"] if eng is not None else -1

    def all_reduce(self, tensor, op=_c.ReduceOp.SUM, group=None, async_op=False):
        return self.run_collective("all_reduce", tensor=tensor, op=op, group=group, async_op=async_op)

    def inference_all_reduce(self, tensor, op=_c.ReduceOp.SUM, group=None):
        return self.run_collective("inference_all_reduce", tensor=tensor, op=op, group=group)

    def broadcast(self, tensor, src, group=None, async_op=False):
        return self.run_collective("broadcast", tensor=tensor, src=src, group=group, async_op=async_op)

    def all_gather(self, tensor_list, tensor, group=None, async_op=False):
        return self.run_collective("all_gather", tensor_list=tensor_list, tensor=tensor, group=group, async_op=async_op)

    def all_gather_into_tensor(self, output_tensor, input_tensor, group=None, async_op=False):
        return self.run_collective("all_gather_into_tensor", output_tensor=output_tensor, tensor=input_tensor, group=group,
                                   async_op=async_op)

    def reduce_scatter_tensor(self, output_tensor, input_tensor, op=_c.ReduceOp.SUM, group=None, async_op=False):
        return self.run_collective("reduce_scatter_tensor", output_tensor=output_tensor, tensor=input_tensor, op=op, group=group,
                                   async_op=async_op)

    def barrier(self, group=None, async_op=False):
        return self.run_collective("barrier", group=group, async_op=async_op)

    def stats(self) -> Dict[str, object]:
        eng = shm.attached(None)
        return {"single_node": self.single_node, "calls": dict(self.calls), "engine": eng.info() if eng is not None else None}

    def close(self) -> None:
        """Detach every segment (collective per group) and hand the general collectives back to torch.distributed."""
        for g in [None] + list(self.process_groups):
            if shm.attached(g) is not None:
                shm.detach(g)
        shm.route_all_collectives = False


def build_ccl_op():
    """The native host-collectives library behind ``CCLBackend`` (oneCCL in the reference, the shared-memory engine here)."""
    from .torch import build_shm_op
    return build_shm_op()


class CCLHandler:
    """Completion handle of a host collective: they complete before returning, so ``wait`` has nothing left to do."""

    def __init__(self, ccl_comm_op=None):
        self.ccl_comm_op = ccl_comm_op

    def wait(self):
        return None
