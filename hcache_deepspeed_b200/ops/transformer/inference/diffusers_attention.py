"""Fused attention of diffusers UNet / VAE transformer blocks (``ops/transformer/inference/diffusers_attention.py:35-196``):
self-attention takes one fused QKV GEMM, cross-attention projects K/V from the text context; no mask, no cache."""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn.functional as F
from torch import nn


class DeepSpeedDiffusersAttention(nn.Module):
    layer_id = 0

    def __init__(self, config):
        super().__init__()
        self.config = config
        self.config.layer_id = DeepSpeedDiffusersAttention.layer_id
        DeepSpeedDiffusersAttention.layer_id += 1
        h, dt = config.hidden_size, getattr(config, "dtype", torch.float32)
        self.heads = config.heads
        mk = lambda *shape: nn.Parameter(torch.empty(*shape, dtype=dt), requires_grad=False)  # noqa: E731
        self.attn_qkvw = mk(h, 3 * h)          # self-attention: [in, q|k|v]
        self.attn_qw, self.attn_kw, self.attn_vw = mk(h, h), None, None      # cross-attention: k/v take the context width
        self.attn_qkvb = None
        self.attn_ow, self.attn_ob = mk(h, h), mk(h)
        self.do_out_bias = True
        self.norm_factor = math.sqrt(h // self.heads)

    def _heads(self, x: torch.Tensor) -> torch.Tensor:
        b, s, _ = x.shape
        return x.view(b, s, self.heads, -1).transpose(1, 2)

    def forward(self, input: torch.Tensor, context: Optional[torch.Tensor] = None, input_mask=None) -> torch.Tensor:
        if context is None:
            q, k, v = (input @ self.attn_qkvw).chunk(3, dim=-1)
        else:
            q, k, v = input @ self.attn_qw, context @ self.attn_kw, context @ self.attn_vw
        ctx = F.scaled_dot_product_attention(self._heads(q), self._heads(k), self._heads(v))
        ctx = ctx.transpose(1, 2).reshape(input.shape[0], input.shape[1], -1)
        out = ctx @ self.attn_ow
        return out + self.attn_ob if self.do_out_bias else out
