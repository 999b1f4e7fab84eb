"""Quantiser / binariser autograd functions under their reference names (``deepspeed/compression/utils.py:12-222``).
Signature: ``apply(input, num_bits, min_value=None, max_value=None, num_groups=1)``; gradients pass straight through."""
from __future__ import annotations

import torch

from .basic_layer import _STEBinary, _STEQuant, _STETernary, _TopKBinarizer


class TopKBinarizer(_TopKBinarizer):
    pass


class SymQuantizer(torch.autograd.Function):

    @staticmethod
    def forward(ctx, input, num_bits, min_value=None, max_value=None, num_groups=1):
        assert (min_value is None and max_value is None) or (min_value is not None and max_value is not None and num_groups == 1)
        if max_value is not None:      # static range (activation quantisation with a tracked range)
            qmax = 2 ** (num_bits - 1) - 1
            scale = max(abs(float(min_value)), abs(float(max_value))) / qmax
            return (torch.clamp(torch.round(input / scale), -qmax - 1, qmax) * scale).to(input.dtype)
        return _STEQuant.forward(ctx, input, num_bits, True, num_groups)

    @staticmethod
    def backward(ctx, grad_output):
        return grad_output, None, None, None, None


class AsymQuantizer(torch.autograd.Function):

    @staticmethod
    def forward(ctx, input, num_bits, min_value=None, max_value=None, num_groups=1):
        assert (min_value is None and max_value is None) or (min_value is not None and max_value is not None and num_groups == 1)
        if max_value is not None:
            scale = (float(max_value) - float(min_value)) / (2 ** num_bits - 1)
            zero = float(min_value)
            return (torch.clamp(torch.round((input - zero) / scale), 0, 2 ** num_bits - 1) * scale + zero).to(input.dtype)
        return _STEQuant.forward(ctx, input, num_bits, False, num_groups)

    @staticmethod
    def backward(ctx, grad_output):
        return grad_output, None, None, None, None


class TernaryQuantizer(torch.autograd.Function):

    @staticmethod
    def forward(ctx, input, num_bits, min_value=None, max_value=None, num_groups=1):
        assert min_value is None and max_value is None
        return _STETernary.forward(ctx, input, num_groups)

    @staticmethod
    def backward(ctx, grad_output):
        return grad_output, None, None, None, None


class BinaryQuantizer(torch.autograd.Function):

    @staticmethod
    def forward(ctx, input, num_bits, min_value=None, max_value=None, num_groups=1):
        assert min_value is None and max_value is None
        return _STEBinary.forward(ctx, input, num_groups)

    @staticmethod
    def backward(ctx, grad_output):
        return grad_output, None, None, None, None
