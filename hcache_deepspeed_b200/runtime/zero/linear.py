"""Memory-lean linear for ZeRO-3: the autograd node saves the input and the weight PARAMETER OBJECT (not a copy of the
gathered weight), so the full weight can be released after the forward and re-gathered for the backward.
Reference: ``deepspeed/runtime/zero/linear.py:40-140`` (``LinearFunctionForZeroStage3``, ``zero3_linear_wrap``,
``LinearModuleForZeroStage3``)."""
from __future__ import annotations

import math
from typing import Optional

import torch
from torch import nn


class LinearFunctionForZeroStage3(torch.autograd.Function):

    @staticmethod
    def forward(ctx, input: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None):
        ctx.save_for_backward(input, weight, bias)
        out = input.matmul(weight.t())
        if bias is not None:
            out = out + bias
        return out

    @staticmethod
    def backward(ctx, grad_output: torch.Tensor):
        input, weight, bias = ctx.saved_tensors
        grad_input = grad_weight = grad_bias = None
        if ctx.needs_input_grad[0]:
            grad_input = grad_output.matmul(weight)
        if ctx.needs_input_grad[1]:
            go = grad_output.reshape(-1, grad_output.shape[-1])
            grad_weight = go.t().matmul(input.reshape(-1, input.shape[-1]))
        if bias is not None and ctx.needs_input_grad[2]:
            grad_bias = grad_output.reshape(-1, grad_output.shape[-1]).sum(0)
        return grad_input, grad_weight, grad_bias


def zero3_linear_wrap(input: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    return LinearFunctionForZeroStage3.apply(input, weight, bias)


class LinearModuleForZeroStage3(nn.Module):
    """Drop-in ``nn.Linear`` whose forward goes through :class:`LinearFunctionForZeroStage3`."""

    def __init__(self, in_features: int, out_features: int, bias: bool = True) -> None:
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        self.weight = nn.Parameter(torch.empty(out_features, in_features))
        self.bias = nn.Parameter(torch.empty(out_features)) if bias else None
        self.reset_parameters()

    def reset_parameters(self) -> None:
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        if self.bias is not None:
            bound = 1 / math.sqrt(self.in_features)
            nn.init.uniform_(self.bias, -bound, bound)

    def forward(self, input: torch.Tensor) -> torch.Tensor:
        return zero3_linear_wrap(input, self.weight, self.bias)

    def extra_repr(self) -> str:
        return f"in_features={self.in_features}, out_features={self.out_features}, bias={self.bias is not None}"
