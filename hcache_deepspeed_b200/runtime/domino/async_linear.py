"""Tensor-parallel linears for Domino (``deepspeed/runtime/domino/async_linear.py:18-140``).

``DominoAsyncColumnParallelLinear``: the forward needs no communication (the weight is split over output features);
the BACKWARD has to all-reduce the input gradient over the tensor-parallel group — that all-reduce is launched
asynchronously and its handle is parked in ``handle_dic[h_id]`` so the caller waits for it only where the gradient is
consumed, with the other micro-batch's backward compute in between. ``RowParallelLinearNoComm`` returns the partial
sum and leaves the (asynchronous) all-reduce to the caller."""
from __future__ import annotations

import math
from typing import Callable, Optional

import torch
import torch.nn.functional as F

from ... import comm as dist


def _group_size(group) -> int:
    return dist.get_world_size(group=group) if (group is not None and dist.is_initialized()) else 1


class DominoAsyncColumnParallelLinearImpl(torch.autograd.Function):

    @staticmethod
    def forward(ctx, inp, weight, bias, handle_dic, h_id, group):
        ctx.save_for_backward(inp, weight)
        ctx.has_bias, ctx.handle_dic, ctx.h_id, ctx.group = bias is not None, handle_dic, h_id, group
        return F.linear(inp, weight, bias)

    @staticmethod
    def backward(ctx, grad_output):
        inp, weight = ctx.saved_tensors
        grad_input = grad_output.matmul(weight)
        if _group_size(ctx.group) > 1:
            ctx.handle_dic[ctx.h_id] = dist.all_reduce(grad_input, group=ctx.group, async_op=True)
        go = grad_output.reshape(-1, grad_output.shape[-1])
        grad_weight = go.t().matmul(inp.reshape(-1, inp.shape[-1]))       # overlaps the all-reduce above
        grad_bias = go.sum(0) if ctx.has_bias else None
        return grad_input, grad_weight, grad_bias, None, None, None


class DominoAsyncColumnParallelLinear(torch.nn.Module):

    def __init__(self, input_size: int, output_size: int, _tp_group=None, config=None,
                 init_method: Optional[Callable] = None, bias: bool = True, skip_bias_add: bool = False, dtype=None):
        super().__init__()
        self.tp_group = _tp_group
        tp = _group_size(_tp_group)
        assert output_size % tp == 0, "output_size must divide by the tensor-parallel size"
        self.output_size_per_partition = output_size // tp
        self.skip_bias_add = skip_bias_add
        dtype = dtype or getattr(config, "params_dtype", None)
        self.weight = torch.nn.Parameter(torch.empty(self.output_size_per_partition, input_size, dtype=dtype))
        (init_method or (lambda w: torch.nn.init.kaiming_uniform_(w, a=math.sqrt(5))))(self.weight)
        self.bias = torch.nn.Parameter(torch.zeros(self.output_size_per_partition, dtype=dtype)) if bias else None

    def forward(self, input_: torch.Tensor, handle_dic: dict, h_id):
        bias = self.bias if not self.skip_bias_add else None
        out = DominoAsyncColumnParallelLinearImpl.apply(input_, self.weight, bias, handle_dic, h_id, self.tp_group)
        return out, (self.bias if self.skip_bias_add else None)


class RowParallelLinearNoComm(torch.nn.Module):

    def __init__(self, input_size: int, output_size: int, config=None, init_method: Optional[Callable] = None,
                 bias: bool = True, stride: int = 1, skip_bias_add: bool = False, _tp_group=None, dtype=None):
        super().__init__()
        tp = _group_size(_tp_group)
        assert input_size % tp == 0, "input_size must divide by the tensor-parallel size"
        self.input_size_per_partition = input_size // tp
        self.skip_bias_add = skip_bias_add
        dtype = dtype or getattr(config, "params_dtype", None)
        self.weight = torch.nn.Parameter(torch.empty(output_size, self.input_size_per_partition, dtype=dtype))
        (init_method or (lambda w: torch.nn.init.kaiming_uniform_(w, a=math.sqrt(5))))(self.weight)
        self.bias = torch.nn.Parameter(torch.zeros(output_size, dtype=dtype)) if bias else None

    def forward(self, input_: torch.Tensor):
        """Partial sum of this rank's input slice — NOT reduced; the bias is returned for the caller to add after the
        reduction (adding it here would count it once per rank)."""
        return F.linear(input_, self.weight), self.bias
