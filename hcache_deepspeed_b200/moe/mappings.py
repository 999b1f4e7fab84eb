"""Token gather / drop across the tensor-parallel group around an MoE layer (so each TP rank routes a distinct
slice of tokens). Reference: ``deepspeed/moe/mappings.py:28-117``."""
from __future__ import annotations

import torch

from .. import comm as dist
from ..utils import groups


def _tp():
    g = groups._get_model_parallel_group() if hasattr(groups, "_get_model_parallel_group") else None
    ws = groups._get_model_parallel_world_size()
    return g, ws, groups._get_model_parallel_rank()


def _gather_tokens(x: torch.Tensor, dim: int = 0) -> torch.Tensor:
    g, ws, _ = _tp()
    if ws == 1:
        return x
    x = x.contiguous()
    parts = [torch.empty_like(x) for _ in range(ws)]
    dist.all_gather(parts, x, group=g)
    return torch.cat(parts, dim=dim).contiguous()


def _drop_tokens(x: torch.Tensor, dim: int = 0) -> torch.Tensor:
    g, ws, rank = _tp()
    if ws == 1:
        return x
    assert x.shape[dim] % ws == 0, f"input dimension {dim} ({x.shape[dim]}) is not divisible by tensor parallel world size ({ws})"
    n = x.shape[dim] // ws
    return torch.narrow(x, dim, rank * n, n).contiguous()


class _GatherTokens(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, dim):
        ctx.dim = dim
        return _gather_tokens(x, dim)

    @staticmethod
    def backward(ctx, grad):
        return _drop_tokens(grad, ctx.dim), None


class _DropTokens(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, dim):
        ctx.dim = dim
        return _drop_tokens(x, dim)

    @staticmethod
    def backward(ctx, grad):
        return _gather_tokens(grad, ctx.dim), None


def gather_tokens(x, dim=0):
    if groups._get_model_parallel_world_size() == 1:
        return x
    return _GatherTokens.apply(x, dim)


def drop_tokens(x, dim=0):
    if groups._get_model_parallel_world_size() == 1:
        return x
    return _DropTokens.apply(x, dim)
