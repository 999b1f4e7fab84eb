"""MiCS: ZeRO-3 with parameters sharded inside a *sub-group* of ``mics_shard_size`` ranks and replicated across
sub-groups (gathers stay inside a node / NVSwitch domain; gradients are reduced inside the shard group and then
all-reduced across replicas). Reference: ``deepspeed/runtime/zero/mics.py`` (MiCS_Init :62, MiCS_Optimizer :358),
``mics_utils.py`` (group creation)."""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional

import torch

from ... import comm as dist
from .partition_parameters import Init
from .partition_parameters import AllGatherCoalescedHandle
from .stage3 import DeepSpeedZeRoOffload, DeepSpeedZeroOptimizer_Stage3


@dataclass
class MiCS_CommGroups:
    param_shard_group: object = None
    param_shard_size: int = -1
    param_shard_rank: int = -1
    param_repli_group: object = None
    param_repli_size: int = -1
    param_repli_rank: int = -1
    param_intra_node_group: object = None
    param_inter_node_shard_group: object = None


def mics_group_ranks(world_size: int, shard_size: int):
    """(shard groups, replicate groups): consecutive ranks shard, same-offset ranks replicate."""
    assert world_size % shard_size == 0, "world size must be divisible by mics_shard_size"
    shard = [list(range(i, i + shard_size)) for i in range(0, world_size, shard_size)]
    repl = [list(range(o, world_size, shard_size)) for o in range(shard_size)]
    return shard, repl


def create_mics_comm_groups(shard_size: int, dp_group=None, hierarchical_allgather: bool = False, mpu=None) -> MiCS_CommGroups:
    ws, me = dist.get_world_size(), dist.get_rank()
    shard, repl = mics_group_ranks(ws, shard_size)
    g = MiCS_CommGroups()
    for ranks in shard:
        grp = dist.new_group(ranks)
        if me in ranks:
            g.param_shard_group, g.param_shard_size, g.param_shard_rank = grp, len(ranks), ranks.index(me)
    for ranks in repl:
        grp = dist.new_group(ranks)
        if me in ranks:
            g.param_repli_group, g.param_repli_size, g.param_repli_rank = grp, len(ranks), ranks.index(me)
    node = _node_size()
    if hierarchical_allgather and shard_size > node and shard_size % node == 0:
        # a shard group that spans nodes: gather across nodes among ranks of equal local index first (few large messages
        # over the network), then inside each node (NVLink) - ref: mics_utils.py create_mics_comm_groups
        for ranks in shard:
            for lo in range(0, shard_size, node):
                intra = ranks[lo:lo + node]
                grp = dist.new_group(intra)
                if me in intra:
                    g.param_intra_node_group = grp
            for off in range(node):
                inter = ranks[off::node]
                grp = dist.new_group(inter)
                if me in inter:
                    g.param_inter_node_shard_group = grp
    return g


def _node_size() -> int:
    import os
    for key in ("HDS_MICS_NODE_SIZE", "LOCAL_WORLD_SIZE"):
        if os.environ.get(key):
            return int(os.environ[key])
    return max(1, torch.cuda.device_count()) if torch.cuda.is_available() else 1


def hierarchical_all_gather(local: torch.Tensor, comm_groups: MiCS_CommGroups) -> torch.Tensor:
    """All-gather of equal slices over a shard group in two hops (inter-node among equal local indices, then intra-node),
    returned in shard-rank order. Falls back to one all-gather without the hierarchical groups."""
    flat = local.contiguous().view(-1)
    if not has_hierarchical_all_gather_groups(comm_groups):
        out = torch.empty(flat.numel() * comm_groups.param_shard_size, dtype=flat.dtype, device=flat.device)
        dist.all_gather_into_tensor(out, flat, group=comm_groups.param_shard_group)
        return out
    n_nodes = dist.get_world_size(group=comm_groups.param_inter_node_shard_group)
    node = dist.get_world_size(group=comm_groups.param_intra_node_group)
    across = torch.empty(flat.numel() * n_nodes, dtype=flat.dtype, device=flat.device)
    dist.all_gather_into_tensor(across, flat, group=comm_groups.param_inter_node_shard_group)   # [node j] slices of my local index
    both = torch.empty(across.numel() * node, dtype=flat.dtype, device=flat.device)
    dist.all_gather_into_tensor(both, across, group=comm_groups.param_intra_node_group)          # [local i][node j]
    return both.view(node, n_nodes, -1).transpose(0, 1).reshape(-1)                             # -> [node j][local i] = rank order


class MiCS_Init(Init):

    def __init__(self, module=None, data_parallel_group=None, sequence_data_parallel_group=None, mem_efficient_linear=True,
                 remote_device=None, pin_memory=False, config_dict_or_path=None, config=None, enabled=True, dtype=None,
                 mpu=None):
        cfg = config_dict_or_path if config_dict_or_path is not None else config
        shard = int(((cfg or {}).get("zero_optimization") or {}).get("mics_shard_size", -1)) if isinstance(cfg, dict) else -1
        self.mics_comm_groups = None
        group = data_parallel_group
        if enabled and shard > 0 and dist.is_initialized():
            self.mics_comm_groups = create_mics_comm_groups(shard, data_parallel_group, mpu=mpu)
            group = self.mics_comm_groups.param_shard_group
        super().__init__(module=module, data_parallel_group=group, remote_device=remote_device, pin_memory=pin_memory,
                         config_dict_or_path=cfg, enabled=enabled, dtype=dtype, mpu=mpu)


class MiCS_Optimizer(DeepSpeedZeroOptimizer_Stage3):
    """Stage-3 optimizer over the shard group, plus a gradient all-reduce across replica groups before the step."""

    def __init__(self, module, init_optimizer, *args, mics_comm_groups: Optional[MiCS_CommGroups] = None, **kwargs):
        self.mics_comm_groups = mics_comm_groups
        if mics_comm_groups is not None:
            kwargs["dp_process_group"] = mics_comm_groups.param_shard_group
        super().__init__(module, init_optimizer, *args, **kwargs)

    @torch.no_grad()
    def allreduce_mics_shard_grads(self, partitioned_grads=None) -> None:
        """Average this rank's gradient slices with the ranks that hold the same slices in the other replica groups."""
        g = self.mics_comm_groups
        if g is None or g.param_repli_size <= 1:
            return
        grads = list(partitioned_grads) if partitioned_grads is not None else list(self.grad_partitions.values())
        for t in grads:
            dist.all_reduce(t, group=g.param_repli_group)
        scale_tensors(grads, g.param_repli_size)

    @torch.no_grad()
    def step(self, closure=None):
        self.allreduce_mics_shard_grads()
        return super().step(closure)


class MiCS_Offload(DeepSpeedZeRoOffload):
    """The stage-3 hook installer over MiCS-partitioned parameters (their process group is the shard group, so the fetch
    / release logic is the base class's)."""


def scale_tensors(tensors, scale) -> None:
    """Divide every tensor in place (averaging after a summed collective)."""
    for t in tensors:
        t.div_(scale)


def has_hierarchical_all_gather_groups(comm_groups) -> bool:
    """Were both the intra-node group and the inter-node shard group created (the two-hop gather is possible)?"""
    return getattr(comm_groups, "param_intra_node_group", None) is not None and \
        getattr(comm_groups, "param_inter_node_shard_group", None) is not None


class MiCS_AllGatherCoalescedHandle(AllGatherCoalescedHandle):   # noqa: F811
    """Handle of a (possibly two-hop) MiCS gather; completion is the base handle's."""
