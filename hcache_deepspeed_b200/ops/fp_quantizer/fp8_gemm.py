"""Weight-only FP8 matmul (``deepspeed/ops/fp_quantizer/fp8_gemm.py``; the reference's fast path is a Triton kernel in
``fp8_gemm_triton.py``): bf16 activations x group-quantised FP8 weights. The weights are de-quantised group-wise and go
through the regular bf16 GEMM; the W8A8 tcgen05 path with per-channel scales is ``ops/fp8.py`` / ``hds_gemm_fp8``."""
from __future__ import annotations

import torch


def matmul_fp8(inp: torch.Tensor, weight: torch.Tensor, scale: torch.Tensor, quantization_group_size: int, quantizer) -> torch.Tensor:
    return matmul_fp8_fallback(inp, weight, scale, quantization_group_size, quantizer)


def matmul_fp8_fallback(inp: torch.Tensor, weight: torch.Tensor, scale: torch.Tensor, quantization_group_size: int, quantizer) -> torch.Tensor:
    w = quantizer.dequantize(weight, scale=scale)
    return torch.matmul(inp, w.to(inp.dtype))
