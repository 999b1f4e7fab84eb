"""``QuantizedParameter``: a parameter that stores minifloat codes (FP8/6/4/12) and dequantises on access;
``QuantizedLinear`` uses it for its weight. Reference: ``deepspeed/linear/quantization.py:18-150``."""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops.fp_quantizer import FP_Quantize
from .config import QuantizationConfig


class QuantizedParameter(nn.Parameter):
    """Holds packed codes in ``.data`` once quantised (on construction for CPU data is deferred until ``.to(device)``,
    matching the reference where quantisation happens on the accelerator)."""

    def __new__(cls, data: Optional[torch.Tensor] = None, requires_grad: bool = False,
                quantization_config: Optional[QuantizationConfig] = None, quantizer: Optional[FP_Quantize] = None):
        if requires_grad:
            raise ValueError("requires_grad=True is not supported with QuantizedParameter")
        if data is None:
            data = torch.empty(0)
        self = torch.Tensor._make_subclass(cls, data, requires_grad)
        self.quantization_config = QuantizationConfig() if quantization_config is None else quantization_config
        self.quantizer = quantizer if quantizer is not None else FP_Quantize(group_size=self.quantization_config.group_size)
        self._quantized = False
        self._ensure_quantized(self)
        return self

    def _ensure_quantized(self, tensor: torch.Tensor):
        if tensor.dtype == self.quantization_config.q_dtype or tensor.numel() == 0 or self._quantized:
            return
        n = tensor.numel()
        gs = self.quantization_config.group_size
        assert n % gs == 0, f"parameter of {n} elements is not a multiple of the quantisation group size {gs}"
        with torch.no_grad():
            tensor.data = self.quantizer.quantize(tensor.data.contiguous(), q_bits=self.quantization_config.q_bits,
                                                  q_mantisa_bits=self.quantization_config.mantissa_bits)
        self._quantized = True

    def dequantized(self) -> torch.Tensor:
        if self._quantized:
            with torch.no_grad():
                return self.quantizer.dequantize(self.data, q_bits=self.quantization_config.q_bits,
                                                 q_mantisa_bits=self.quantization_config.mantissa_bits)
        return self.data

    def __getstate__(self):
        return {**self.__dict__, "data": self.data, "requires_grad": self.requires_grad}

    def to(self, *args, **kwargs):
        t = QuantizedParameter.__new__(QuantizedParameter, None, False, self.quantization_config, self.quantizer)
        t.data = self.data.to(*args, **kwargs)
        t._quantized = self._quantized
        return t

    def cuda(self, device=None, non_blocking=False):
        return self.to(device="cuda" if device is None else device, non_blocking=non_blocking)


class QuantizedLinear(nn.Linear):
    """Linear whose weight is a :class:`QuantizedParameter` (dequantised per forward)."""

    def __init__(self, input_dim: int, output_dim: int, bias: bool = False,
                 quantization_config: Optional[QuantizationConfig] = None, dtype=torch.bfloat16):
        super().__init__(input_dim, output_dim, bias=bias, dtype=dtype)
        assert dtype == torch.bfloat16, "currently only supports bfloat16 dtype"
        self.weight = QuantizedParameter(self.weight.data, quantization_config=quantization_config)

    def forward(self, input: torch.Tensor) -> torch.Tensor:
        return F.linear(input, self.weight.dequantized(), self.bias)
