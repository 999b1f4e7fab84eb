"""Weight-quantised Linear / Embedding: weights are stored as int4 / int8 codes with per-group (scale, offset) and
dequantised on the fly in forward. Reference: ``deepspeed/inference/quantization/layers.py:24-115``."""
from __future__ import annotations

from typing import Dict

import torch
import torch.nn.functional as F

from .utils import DeQuantizer, Quantizer


class QuantizedLinear(torch.nn.Linear):

    def __init__(self, config: Dict, pre_quant_layer: torch.nn.Linear) -> None:
        super().__init__(in_features=pre_quant_layer.in_features, out_features=pre_quant_layer.out_features,
                         bias=pre_quant_layer.bias is not None, device=pre_quant_layer.weight.device,
                         dtype=pre_quant_layer.weight.dtype)
        self.config = config
        self.bias = pre_quant_layer.bias
        self.weight_shape, self.weight_dtype = tuple(pre_quant_layer.weight.shape), pre_quant_layer.weight.dtype
        codes, scale, offset = Quantizer(config).quantize(pre_quant_layer.weight.data)
        self.weight = torch.nn.Parameter(codes, requires_grad=False)
        self.weight.dequantizer = DeQuantizer(config, self.weight_dtype)
        self.register_buffer("quant_scale", scale)
        self.register_buffer("quant_min", offset)

    def dequantized_weight(self) -> torch.Tensor:
        return self.weight.dequantizer.dequantize(self.weight.data, self.quant_scale, self.quant_min, self.weight_shape)

    def forward(self, input: torch.Tensor) -> torch.Tensor:
        return F.linear(input, self.dequantized_weight(), self.bias)


class QuantizedEmbedding(torch.nn.Embedding):

    def __init__(self, config: Dict, pre_quant_layer: torch.nn.Embedding) -> None:
        super().__init__(num_embeddings=pre_quant_layer.num_embeddings, embedding_dim=pre_quant_layer.embedding_dim,
                         padding_idx=pre_quant_layer.padding_idx, max_norm=pre_quant_layer.max_norm,
                         norm_type=pre_quant_layer.norm_type, scale_grad_by_freq=pre_quant_layer.scale_grad_by_freq,
                         sparse=pre_quant_layer.sparse, _weight=pre_quant_layer.weight, device=pre_quant_layer.weight.device,
                         dtype=pre_quant_layer.weight.dtype)
        self.config = config
        self.weight_shape, self.weight_dtype = tuple(pre_quant_layer.weight.shape), pre_quant_layer.weight.dtype
        codes, scale, offset = Quantizer(config).quantize(pre_quant_layer.weight.data)
        self.weight = torch.nn.Parameter(codes, requires_grad=False)
        self.weight.dequantizer = DeQuantizer(config, self.weight_dtype)
        self.register_buffer("quant_scale", scale)
        self.register_buffer("quant_min", offset)

    def forward(self, input: torch.Tensor) -> torch.Tensor:
        w = self.weight.dequantizer.dequantize(self.weight.data, self.quant_scale, self.quant_min, self.weight_shape)
        return F.embedding(input, w, self.padding_idx, self.max_norm, self.norm_type, self.scale_grad_by_freq, self.sparse)


QUANTIZATION_LAYER_MAPPINGS = {torch.nn.Linear: QuantizedLinear, torch.nn.Embedding: QuantizedEmbedding}
