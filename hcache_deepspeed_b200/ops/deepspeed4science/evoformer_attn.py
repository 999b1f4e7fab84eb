"""Evoformer attention (AlphaFold-style MSA row / column / triangle attention): softmax(QK^T / sqrt(d) + bias1 + bias2) V
with up to two broadcast biases, forward and backward. The reference ships a CUTLASS-2 kernel pair
(``csrc/deepspeed4science/evoformer_attn``, sm_70-80); here the op is expressed with ``scaled_dot_product_attention``
on the summed bias, which dispatches to the memory-efficient kernel and gives bias gradients through autograd.
Reference: ``deepspeed/ops/deepspeed4science/evoformer_attn.py:15-100``."""
from __future__ import annotations

from typing import List

import torch


def DS4Sci_EvoformerAttention(Q: torch.Tensor, K: torch.Tensor, V: torch.Tensor, biases: List[torch.Tensor]) -> torch.Tensor:
    """Q/K/V: [batch, n_seq, n_res, heads, dim]; biases: up to two of [batch, n_seq, 1, 1, n_res] (mask bias) and
    [batch, 1, heads, n_res, n_res] (pair bias). Returns [batch, n_seq, n_res, heads, dim]."""
    assert len(biases) <= 2
    if len(biases) == 0:
        biases = [None]
    if len(biases) == 1:
        biases = list(biases) + [None]
    b1, b2 = biases
    B, N, S, H, D = Q.shape
    if b1 is not None:
        assert b1.shape == (B, N, 1, 1, S), "bias1 shape is incorrect"
    if b2 is not None:
        assert b2.shape == (B, 1, H, S, S), "bias2 shape is incorrect"
    q, k, v = (t.permute(0, 1, 3, 2, 4) for t in (Q, K, V))  # [B, N, H, S, D]
    bias = None
    if b1 is not None:
        bias = b1.to(q.dtype)
    if b2 is not None:
        bias = b2.to(q.dtype) if bias is None else bias + b2.to(q.dtype)
    if bias is not None:
        bias = bias.expand(B, N, H, S, S)
    out = torch.nn.functional.scaled_dot_product_attention(q, k, v, attn_mask=bias)
    return out.permute(0, 1, 3, 2, 4).contiguous()
