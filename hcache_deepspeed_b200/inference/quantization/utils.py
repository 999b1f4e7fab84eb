"""Group-wise asymmetric weight quantisation helpers for ZeroQuant-style inference (4 / 8 bit).
Reference: ``deepspeed/inference/quantization/utils.py:23-190``."""
from __future__ import annotations

from typing import Dict, Tuple

import torch

from ...ops.quantizer import dequantize_int, quantize_int


class Quantizer:

    def __init__(self, config: Dict):
        self.config = config
        assert config["num_bits"] in (4, 8), "Only INT4 and INT8 quantization is supported."
        assert config["symmetric"] is False, "Only asymmetric quantization is supported at this moment."

    def quantize(self, tensor: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        """Returns (codes, scale [groups], offset [groups]) with groups along ``group_dim`` of size ``group_size``."""
        gs = self.config["group_size"]
        dim = self.config["group_dim"]
        assert tensor.shape[dim] % gs == 0, f"Tensor shape: {tensor.shape} quantization config {self.config}"
        t = tensor.transpose(dim, -1).contiguous() if dim not in (-1, tensor.dim() - 1) else tensor.contiguous()
        groups = t.numel() // gs
        codes, params = quantize_int(t, groups, self.config["num_bits"], symmetric=False)
        return codes, params[:, 0].contiguous(), params[:, 1].contiguous()


class DeQuantizer:

    def __init__(self, config: Dict, dtype: torch.dtype):
        self.config, self.dtype = config, dtype

    def dequantize(self, codes: torch.Tensor, scale: torch.Tensor, offset: torch.Tensor, shape) -> torch.Tensor:
        params = torch.stack([scale, offset], dim=1).contiguous()
        dim = self.config["group_dim"]
        nd = len(shape)
        if dim in (-1, nd - 1):
            return dequantize_int(codes, params, self.config["num_bits"], self.dtype, shape)
        tshape = list(shape)
        tshape[dim], tshape[-1] = tshape[-1], tshape[dim]
        return dequantize_int(codes, params, self.config["num_bits"], self.dtype, tshape).transpose(dim, -1).contiguous()


def get_AsyncPartitionedParameterSwapper(model: torch.nn.Module):
    for p in model.parameters():
        if hasattr(p, "nvme_swapper") and p.nvme_swapper is not None:
            return p.nvme_swapper
    return None


def recursive_setattr(model, module_name, module):
    parts = module_name.split(".")
    obj = model
    for p in parts[:-1]:
        obj = getattr(obj, p)
    setattr(obj, parts[-1], module)
