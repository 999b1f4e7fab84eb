"""Tensor-parallel building blocks used by AutoTP and the policies.

``LinearLayer``            column-parallel: owns a slice of the OUTPUT features, no communication
``LinearAllreduce``        row-parallel: owns a slice of the INPUT features, output all-reduced (bias added after)
``LmHeadLinearAllreduce``  row-parallel LM head whose input is the full hidden state (it slices it itself)
``EmbeddingLayer`` / ``OPTEmbedding`` / ``Normalize`` / ``RMSNormalize``   thin parameter holders for policies
``RowParallel`` / ``ColumnParallel``  autograd functions for training with AutoTP
Reference: ``deepspeed/module_inject/layers.py`` (RowParallel :51, ColumnParallel :80, LinearAllreduce :565,
LinearLayer :419, LmHeadLinearAllreduce :660, EmbeddingLayer :730, OPTEmbedding :750, RMSNormalize :775).
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F

from .. import comm as dist
from .tp_shard import get_shard_size_list


def _ws(group) -> int:
    return dist.get_world_size(group=group) if (group is not None and dist.is_initialized()) else 1


class RowParallel(torch.autograd.Function):
    """Forward: all-reduce the partial outputs; backward: identity."""

    @staticmethod
    def forward(ctx, group, x: torch.Tensor, is_inference_mode: bool = False):
        ctx.group = group
        if group is None or _ws(group) == 1:
            return x
        out = x.contiguous().clone() if x.requires_grad else x.contiguous()
        if is_inference_mode:
            dist.inference_all_reduce(out, group=group)
        else:
            dist.all_reduce(out, group=group)
        return out

    @staticmethod
    def backward(ctx, grad):
        return None, grad, None


class ColumnParallel(torch.autograd.Function):
    """Forward: identity; backward: all-reduce the input gradient."""

    @staticmethod
    def forward(ctx, group, x: torch.Tensor):
        ctx.group = group
        return x

    @staticmethod
    def backward(ctx, grad):
        if ctx.group is None or _ws(ctx.group) == 1:
            return None, grad
        g = grad.contiguous()
        dist.all_reduce(g, group=ctx.group)
        return None, g


class LinearAllreduce(torch.nn.Module):

    def __init__(self, weight: torch.Tensor, bias: Optional[torch.Tensor] = None, mp_group=None):
        super().__init__()
        self.weight = weight if isinstance(weight, torch.nn.Parameter) else torch.nn.Parameter(weight, requires_grad=False)
        self.bias = None if bias is None else (bias if isinstance(bias, torch.nn.Parameter) else
                                               torch.nn.Parameter(bias, requires_grad=False))
        self.mp_group = mp_group

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        out = torch.matmul(x, self.weight.transpose(-1, -2))
        out = RowParallel.apply(self.mp_group, out, not self.training)
        return out if self.bias is None else out + self.bias

    def extra_repr(self):
        return f"in_features(local)={self.weight.shape[1]}, out_features={self.weight.shape[0]}, bias={self.bias is not None}"


class LmHeadLinearAllreduce(torch.nn.Module):

    def __init__(self, weight, rank: int, world_size: int, bias=None, mp_group=None):
        super().__init__()
        self.weight = weight if isinstance(weight, torch.nn.Parameter) else torch.nn.Parameter(weight, requires_grad=False)
        self.bias = None if bias is None else torch.nn.Parameter(bias, requires_grad=False)
        self.mp_group, self.rank, self.world_size = mp_group, rank, world_size

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        sizes = get_shard_size_list(x.shape[-1], self.world_size, "lm_head")
        start = sum(sizes[:self.rank])
        if self.training and x.requires_grad:
            # each rank back-propagates into ITS slice of x only: the input gradient is complete after a sum over the group
            x = ColumnParallel.apply(self.mp_group, x)
        out = torch.matmul(x[..., start:start + sizes[self.rank]], self.weight.transpose(-1, -2))
        out = RowParallel.apply(self.mp_group, out, not self.training)
        return out if self.bias is None else out + self.bias


class LinearLayer(torch.nn.Module):

    def __init__(self, weight_shape=None, dtype=torch.half, weight=None, bias=None, mp_group=None):
        super().__init__()
        if weight is not None:
            self.weight = weight if isinstance(weight, torch.nn.Parameter) else torch.nn.Parameter(weight, requires_grad=False)
            self.bias = None if bias is None else (bias if isinstance(bias, torch.nn.Parameter) else
                                                   torch.nn.Parameter(bias, requires_grad=False))
        else:
            self.weight = torch.nn.Parameter(torch.empty(weight_shape, dtype=dtype), requires_grad=False)
            self.bias = torch.nn.Parameter(torch.empty(weight_shape[0], dtype=dtype), requires_grad=False) if bias is not None else None
        self.mp_group = mp_group

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = ColumnParallel.apply(self.mp_group, x) if (self.training and x.requires_grad) else x
        out = torch.matmul(x, self.weight.transpose(-1, -2))
        return out if self.bias is None else out + self.bias


class Normalize(torch.nn.Module):

    def __init__(self, dim=None, dtype=torch.float, eps: float = 1e-5, weight=None, bias=None):
        super().__init__()
        if weight is not None:
            self.weight, self.bias = weight, bias
        else:
            self.norm = torch.nn.LayerNorm(dim, eps=eps).to(dtype)
            self.weight, self.bias = self.norm.weight, self.norm.bias
        self.eps = eps

    def forward(self, x):
        return F.layer_norm(x, x.shape[-1:], self.weight, self.bias, eps=self.eps)


class EmbeddingLayer(torch.nn.Module):

    def __init__(self, weight_shape=None, dtype=torch.half, weight=None, bias=None):
        super().__init__()
        self.weight = weight if weight is not None else torch.nn.Parameter(torch.empty(weight_shape, dtype=dtype))

    def forward(self, ids):
        return F.embedding(ids, self.weight)


class OPTEmbedding(EmbeddingLayer):
    """OPT's learned positions are offset by 2 and derived from the attention mask."""

    def __init__(self, weight_shape=None, weight=None, bias=None):
        self.offset = 2
        super().__init__(weight_shape, weight=weight)

    def forward(self, attention_mask: torch.LongTensor, past_key_values_length: int = 0, position_ids=None):
        attention_mask = attention_mask.long()
        positions = (torch.cumsum(attention_mask, dim=1).type_as(attention_mask) * attention_mask).long() - 1
        positions = positions[:, past_key_values_length:]
        return super().forward(positions + self.offset)


class RMSNormalize(torch.nn.Module):

    def __init__(self, dim=None, dtype=torch.float, eps: float = 1e-5, weight=None):
        super().__init__()
        self.weight = weight if weight is not None else torch.nn.Parameter(torch.ones(dim, dtype=dtype))
        self.eps = eps

    def forward(self, x):
        var = x.to(torch.float32).pow(2).mean(-1, keepdim=True)
        x = x * torch.rsqrt(var + self.eps)
        if self.weight.dtype in (torch.float16, torch.bfloat16):
            x = x.to(self.weight.dtype)
        return x * self.weight


class GatherReplacedLayerParams:
    """Context manager that reassembles the full weight of TP-sharded layers (e.g. for saving) and re-shards on exit."""

    def __init__(self, params, module, enabled: bool = True):
        self.enabled = enabled
        self.module = module
        self.layers = [m for m in module.modules() if isinstance(m, (LinearLayer, LinearAllreduce))] if params is None else params

    def __enter__(self):
        if not self.enabled:
            return
        self._saved = []
        for l in self.layers:
            g = l.mp_group
            ws = _ws(g)
            if ws == 1:
                continue
            dim = 1 if isinstance(l, LinearAllreduce) else 0
            parts = [torch.empty_like(l.weight.data) for _ in range(ws)]
            dist.all_gather(parts, l.weight.data.contiguous(), group=g)
            self._saved.append((l, l.weight.data, dim))
            l.weight.data = torch.cat(parts, dim=dim)

    def __exit__(self, *exc):
        if not self.enabled:
            return False
        for l, local, _ in self._saved:
            l.weight.data = local
        return False
