"""AutoTP for TRAINING: shard an HF-style model over a tensor-parallel group and keep data loading consistent inside
the group (every TP rank must see the same batch). Reference: ``deepspeed/runtime/tensor_parallel/tp_manager.py:12-68``,
``deepspeed/__init__.py:369 (tp_model_init)``."""
from __future__ import annotations

import torch

from ... import comm as dist
from ...module_inject.auto_tp import AutoTP
from ...utils import groups


class TpTrainingManager:

    def __init__(self, model: torch.nn.Module, tp_size: int, dtype=None):
        self.module = model
        self.tp_size = tp_size
        if dtype is not None:
            self.module.to(dtype)
        if tp_size > 1:
            assert dist.is_initialized(), "tensor-parallel training needs an initialised process group"
            groups._create_model_parallel(tp_size)
            g = groups._get_model_parallel_group()
            AutoTP.apply(self.module, tp_size, dist.get_rank(group=g), g)
            for p in self.module.parameters():
                p.requires_grad_(p.dtype.is_floating_point)
                p.tensor_model_parallel = True
            self._broadcast_replicated(g)

    def _broadcast_replicated(self, group):
        """Norms / embeddings are replicated over TP ranks: make them identical."""
        from ...module_inject.layers import LinearAllreduce, LinearLayer, LmHeadLinearAllreduce
        sharded = set()
        for m in self.module.modules():
            if isinstance(m, (LinearLayer, LinearAllreduce, LmHeadLinearAllreduce)):
                sharded.add(id(m.weight))
                if isinstance(m, LinearLayer) and m.bias is not None:
                    sharded.add(id(m.bias))
        src = dist.get_global_rank(group, 0)
        for p in self.module.parameters():
            if id(p) not in sharded:
                dist.broadcast(p.data, src, group=group)
                p.tensor_model_parallel = False

    @staticmethod
    def broadcast_batch(batch, group=None):
        """Make TP rank 0's batch the batch of the whole TP group."""
        group = group if group is not None else groups._get_model_parallel_group()
        if group is None or dist.get_world_size(group=group) == 1:
            return batch
        src = dist.get_global_rank(group, 0)
        if torch.is_tensor(batch):
            dist.broadcast(batch, src, group=group)
            return batch
        if isinstance(batch, dict):
            return {k: TpTrainingManager.broadcast_batch(v, group) for k, v in batch.items()}
        if isinstance(batch, (list, tuple)):
            return type(batch)(TpTrainingManager.broadcast_batch(v, group) for v in batch)
        return batch


def tp_model_init(model, tp_size: int, dtype=None):
    """``deepspeed.tp_model_init``: returns the sharded model."""
    return TpTrainingManager(model, tp_size, dtype).module
