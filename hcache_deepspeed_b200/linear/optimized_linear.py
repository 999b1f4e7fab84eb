"""``OptimizedLinear``: a Linear that can be (a) LoRA-adapted with a frozen, optionally sharded and/or quantised base
weight, (b) quantised only, or (c) plain. Reference: ``deepspeed/linear/optimized_linear.py:19-220``."""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import comm as dist
from .config import LoRAConfig, QuantizationConfig
from .quantization import QuantizedLinear, QuantizedParameter


class OptimizedLinear(nn.Module):

    def __new__(cls, input_dim: int, output_dim: int, bias: bool = False, lora_config: Optional[LoRAConfig] = None,
                quantization_config: Optional[QuantizationConfig] = None, dtype=torch.bfloat16):
        if quantization_config is not None and not lora_config:
            return QuantizedLinear(input_dim, output_dim, bias=bias, quantization_config=quantization_config, dtype=dtype)
        if lora_config is not None:
            return LoRAOptimizedLinear(input_dim, output_dim, bias=bias, lora_config=lora_config,
                                       quantization_config=quantization_config, dtype=dtype)
        return nn.Linear(input_dim, output_dim, bias=bias, dtype=dtype)


class LoRAOptimizedLinear(nn.Module):

    def __init__(self, input_dim: int, output_dim: int, bias: bool = False, lora_config: LoRAConfig = None,
                 quantization_config: Optional[QuantizationConfig] = None, device=None, dtype=torch.bfloat16):
        super().__init__()
        self.input_dim, self.output_dim, self.bias = input_dim, output_dim, bias
        self.lora_config, self.quantization_config = lora_config, quantization_config
        self.dtype = dtype
        assert not bias, "bias=True is not supported by LoRAOptimizedLinear"
        self.zero_shards = max(1, lora_config.base_weight_sharding)
        self.sharded_weight_size = int(float(input_dim) // self.zero_shards)
        w = torch.empty(output_dim, self.sharded_weight_size, dtype=dtype, device=device)
        nn.init.xavier_uniform_(w)
        if quantization_config is not None:
            self.weight = QuantizedParameter(w, quantization_config=quantization_config)
        else:
            self.weight = nn.Parameter(w, requires_grad=False)
        self.lora_scaling_factor = lora_config.lora_alpha / lora_config.lora_r
        self.lora_weight_1 = nn.Linear(input_dim, lora_config.lora_r, bias=False, dtype=dtype, device=device)
        self.lora_weight_2 = nn.Linear(lora_config.lora_r, output_dim, bias=False, dtype=dtype, device=device)
        nn.init.kaiming_uniform_(self.lora_weight_1.weight, a=math.sqrt(5))
        nn.init.zeros_(self.lora_weight_2.weight)

    def full_weight(self) -> torch.Tensor:
        local = self.weight.dequantized() if isinstance(self.weight, QuantizedParameter) else self.weight
        if self.zero_shards == 1 or not dist.is_initialized():
            return local
        parts = [torch.empty_like(local) for _ in range(self.zero_shards)]
        dist.all_gather(parts, local.contiguous())
        return torch.cat(parts, dim=1)

    def forward(self, input_tensor: torch.Tensor) -> torch.Tensor:
        with torch.no_grad():
            base = F.linear(input_tensor, self.full_weight())
        lora = self.lora_weight_2(self.lora_weight_1(input_tensor))
        return base + self.lora_scaling_factor * lora
