"""FP8 (E4M3) weights for the ragged executor (``quantization_mode = "wf8af8"``): a weight is stored as its fp8 codes
plus one fp32 scale per output channel; the op back ends quantise activations per token and run the
``tcgen05.mma.kind::f8f6f4`` GEMM (``ops/csrc/gemm_fp8.cu``, ``gemm_2sm.cu``). Takes the place of the reference's
weight-only FP6 linear (``modules/implementations/linear/quantized_linear.py:24-205``)."""
from __future__ import annotations

from typing import Tuple

import torch

E4M3_MAX = 448.0


def quantize_rows_reference(x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """torch restatement of ``hds_quantize_fp8_rows``: (uint8 codes, fp32 per-row scales)."""
    amax = x.float().abs().amax(dim=1)
    scale = torch.where(amax > 0, amax / E4M3_MAX, torch.ones_like(amax))
    q = (x.float() / scale[:, None]).to(torch.float8_e4m3fn).view(torch.uint8)
    return q, scale


class Fp8Weight:
    """``[N, K]`` weight as E4M3 codes (``q``: uint8) and per-output-channel scales (``scale``: fp32 ``[N]``). Row slices
    (``w[r0:r1]``) are views, like the bf16 tensors they replace (the K/V rows of a fused QKV weight)."""

    __slots__ = ("q", "scale")

    def __init__(self, q: torch.Tensor, scale: torch.Tensor):
        assert q.dtype == torch.uint8 and q.dim() == 2 and scale.dtype == torch.float32 and scale.shape == (q.shape[0], )
        self.q, self.scale = q, scale

    @classmethod
    def from_tensor(cls, w: torch.Tensor) -> "Fp8Weight":
        if w.is_cuda:
            from . import kernels
            q, s = kernels.quantize_fp8_rows(w.contiguous())
        else:
            q, s = quantize_rows_reference(w)
        return cls(q, s)

    @property
    def shape(self):
        return self.q.shape

    @property
    def device(self):
        return self.q.device

    def __getitem__(self, rows) -> "Fp8Weight":
        assert isinstance(rows, slice), "Fp8Weight supports row slices only"
        return Fp8Weight(self.q[rows], self.scale[rows])

    def dequantize(self, dtype=torch.float32) -> torch.Tensor:
        return (self.q.view(torch.float8_e4m3fn).float() * self.scale[:, None]).to(dtype)

    def nbytes(self) -> int:
        return self.q.numel() + 4 * self.scale.numel()
