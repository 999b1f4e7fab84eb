"""Block-sparse self-attention: scores outside the layout's blocks are masked out. The reference multiplies only the
live blocks with Triton SDD/DSD kernels (``ops/sparse_attention/{matmul,softmax}.py``); here the layout becomes an
additive mask for ``scaled_dot_product_attention`` (same result; Triton is excluded by design and block-sparse
training is not on this framework's optimised path). Reference: ``sparse_self_attention.py:12-152``."""
from __future__ import annotations

import torch
from torch import nn

from .sparsity_config import SparsityConfig


class SparseSelfAttention(nn.Module):

    def __init__(self, sparsity_config=SparsityConfig(num_heads=4), key_padding_mask_mode="add", attn_mask_mode="mul",
                 max_seq_length=2048):
        super().__init__()
        self.sparsity_config = sparsity_config
        self.key_padding_mask_mode, self.attn_mask_mode = key_padding_mask_mode, attn_mask_mode
        self.register_buffer("master_layout", sparsity_config.make_layout(max_seq_length))
        self._cache = {}

    def get_layout(self, L: int) -> torch.Tensor:
        if L % self.sparsity_config.block != 0:
            raise ValueError(f"Sequence Length, {L}, needs to be dividable by Block size {self.sparsity_config.block}!")
        nb = L // self.sparsity_config.block
        return self.master_layout[..., :nb, :nb]

    def _block_mask(self, L: int, device) -> torch.Tensor:
        if L not in self._cache:
            b = self.sparsity_config.block
            lay = self.get_layout(L).bool()
            self._cache[L] = lay.repeat_interleave(b, dim=1).repeat_interleave(b, dim=2)
        return self._cache[L].to(device)

    def forward(self, query, key, value, rpe=None, key_padding_mask=None, attn_mask=None):
        """query/key/value: [batch, heads, seq, head_dim]."""
        assert query.dtype == key.dtype == value.dtype
        bsz, heads, L, d = query.shape
        if query.shape != key.shape or key.shape != value.shape:
            raise NotImplementedError("only self-attention is supported for now")
        bias = torch.zeros(1, heads, L, L, dtype=query.dtype, device=query.device)
        bias = bias.masked_fill(~self._block_mask(L, query.device).unsqueeze(0), float("-inf"))
        if rpe is not None:
            bias = bias + rpe
        if key_padding_mask is not None:
            kpm = key_padding_mask.to(query.dtype)
            kpm = kpm if self.key_padding_mask_mode == "add" else (1.0 - kpm) * torch.finfo(query.dtype).min
            bias = bias + kpm.reshape(bsz, 1, 1, L)
        if attn_mask is not None:
            am = attn_mask.to(query.dtype)
            am = am if self.attn_mask_mode == "add" else torch.where(am > 0, torch.zeros_like(am), torch.full_like(am, float("-inf")))
            bias = bias + am.reshape(1, 1, L, L)
        return torch.nn.functional.scaled_dot_product_attention(query, key, value, attn_mask=bias)
