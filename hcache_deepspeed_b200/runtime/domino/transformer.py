"""Domino: hide tensor-parallel all-reduces behind compute by splitting the batch into two micro-batches and
pipelining them through each layer — while micro-batch 1's attention output is being all-reduced (async), micro-batch
0's MLP runs, and so on. Reference: ``deepspeed/runtime/domino/transformer.py:20-605``, ``async_linear.py:20-140``."""
from __future__ import annotations

from typing import Callable, Optional

import torch

from ... import comm as dist


class DominoAsyncLinear(torch.nn.Module):
    """Row-parallel linear whose all-reduce is launched asynchronously; ``wait()`` on the returned handle before use."""

    def __init__(self, in_features: int, out_features: int, group=None, bias: bool = False, dtype=None):
        super().__init__()
        self.linear = torch.nn.Linear(in_features, out_features, bias=False, dtype=dtype)
        self.bias = torch.nn.Parameter(torch.zeros(out_features, dtype=dtype)) if bias else None
        self.group = group

    def forward(self, x):
        y = self.linear(x)
        handle = None
        if self.group is not None and dist.is_initialized() and dist.get_world_size(group=self.group) > 1:
            handle = dist.all_reduce(y, group=self.group, async_op=True)
        return y, handle


class _AsyncAllreduce(torch.autograd.Function):
    """Forward: start the all-reduce and return immediately (handle stashed on the module). Backward: identity."""

    @staticmethod
    def forward(ctx, x, group, holder, key):
        if group is not None and dist.is_initialized() and dist.get_world_size(group=group) > 1:
            holder[key] = dist.all_reduce(x, group=group, async_op=True)
        return x

    @staticmethod
    def backward(ctx, g):
        return g, None, None, None


class DominoTransformerLayer(torch.nn.Module):
    """``attn(x) -> row-parallel out`` and ``mlp(x) -> row-parallel out`` are user modules returning PARTIAL sums; this
    layer owns the residual/norm plumbing and the two-micro-batch interleave."""

    def __init__(self, attn: torch.nn.Module, mlp: torch.nn.Module, hidden: int, tp_group=None, eps: float = 1e-5):
        super().__init__()
        self.attn, self.mlp, self.group = attn, mlp, tp_group
        self.ln1, self.ln2 = torch.nn.LayerNorm(hidden, eps=eps), torch.nn.LayerNorm(hidden, eps=eps)
        self._handles = {}

    def _wait(self, key):
        h = self._handles.pop(key, None)
        if h is not None:
            h.wait()

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x0, x1 = torch.chunk(x, 2, dim=0)
        a0 = _AsyncAllreduce.apply(self.attn(self.ln1(x0)), self.group, self._handles, "a0")
        a1 = _AsyncAllreduce.apply(self.attn(self.ln1(x1)), self.group, self._handles, "a1")   # overlaps a0's reduce
        self._wait("a0")
        h0 = x0 + a0
        m0 = _AsyncAllreduce.apply(self.mlp(self.ln2(h0)), self.group, self._handles, "m0")      # overlaps a1's reduce
        self._wait("a1")
        h1 = x1 + a1
        m1 = _AsyncAllreduce.apply(self.mlp(self.ln2(h1)), self.group, self._handles, "m1")      # overlaps m0's reduce
        self._wait("m0")
        y0 = h0 + m0
        self._wait("m1")
        y1 = h1 + m1
        return torch.cat([y0, y1], dim=0)
