"""Grouped / mixed-input GEMM wrappers (``deepspeed/inference/v2/kernels/cutlass_ops/``: ``moe_gemm``, ``mixed_gemm``).
The reference instantiates CUTLASS templates; here the grouped GEMM is the framework's own tcgen05 kernel
(``hds_moe_gemm``), and the weight-only path de-quantises into the same bf16 GEMM."""
from __future__ import annotations

from typing import Optional

import torch

from ....ops import reference as ref
from .core_ops import _act_id
from .ds_kernel import DSKernelBase


class MoEGEMM(DSKernelBase):
    """``ordered_output[rows of expert e] = act(ordered_input[rows of e] @ W[e].T + b[e])`` for all experts in one launch.
    ``weights``: ``[E, N, K]``; ``expert_cumsum`` int32 [E+1]: 128-row padded range starts (from :class:`MoEScatter`)."""

    def __init__(self, fp_dtype: torch.dtype, act_fn="identity", gated: bool = False):
        self._check_dtype(fp_dtype)
        self.act, self.gated = _act_id(act_fn), gated

    def __call__(self, ordered_output, ordered_input, weights, expert_cumsum, biases: Optional[torch.Tensor] = None):
        E, N, K = weights.shape
        if ordered_input.is_cuda:
            if biases is not None:
                raise NotImplementedError("the grouped tcgen05 GEMM has no per-expert bias epilogue")
            from ....ops import kernels
            kernels.moe_gemm(ordered_input, weights.reshape(E * N, K), expert_cumsum, E, ordered_output, act=self.act, gated=self.gated)
            return ordered_output
        bounds = expert_cumsum.tolist()
        for e in range(E):
            lo, hi = bounds[e], bounds[e + 1]
            if hi > lo:
                ordered_output[lo:hi] = ref.gemm(ordered_input[lo:hi], weights[e], bias=None if biases is None else biases[e],
                                                 act=self.act, gated=self.gated, out_dtype=ordered_output.dtype)
        return ordered_output


class MixedGEMM(DSKernelBase):
    """bf16/fp16 activations x int8 / int4 (packed two per byte) weights with per-output-channel scales."""

    def __init__(self, fp_dtype: torch.dtype, act_fn="identity", num_bits: int = 8):
        self._check_dtype(fp_dtype)
        if num_bits not in (4, 8):
            raise ValueError("num_bits must be 4 or 8")
        self.act, self.num_bits = _act_id(act_fn), num_bits

    def dequantize(self, weights: torch.Tensor, scales: torch.Tensor, dtype) -> torch.Tensor:
        if self.num_bits == 4:
            lo = (weights & 0xF).to(torch.int8)
            hi = ((weights >> 4) & 0xF).to(torch.int8)
            w = torch.stack([lo, hi], dim=-1).reshape(weights.shape[0], -1) - 8
        else:
            w = weights.to(torch.int8)
        return (w.float() * scales.float().reshape(-1, 1)).to(dtype)

    def __call__(self, output, hidden_states, weights, scales, biases: Optional[torch.Tensor] = None):
        w = self.dequantize(weights, scales, hidden_states.dtype)
        x2, o2 = hidden_states.reshape(-1, hidden_states.shape[-1]), output.view(-1, output.shape[-1])
        self._ops(x2).gemm(x2, w, o2, bias=biases, act=self.act)
        return output
