"""Softmax over block-sparse scores (reference ``ops/sparse_attention/softmax.py``: Triton kernel with fused scale,
relative position embedding, key-padding and attention masks). The packed blocks of each (head, block-row) are
normalised together; masked-out positions contribute nothing."""
from __future__ import annotations

from typing import Optional

import torch


class Softmax:

    def __init__(self, layout: torch.Tensor, block: int, bench: bool = False):
        self.layout, self.block = layout, block
        nz = layout.nonzero(as_tuple=False)
        self.h, self.i, self.j = nz[:, 0], nz[:, 1], nz[:, 2]
        self.spdims = layout.shape
        # segment id of every packed block = its (head, block-row)
        self.segment = self.h * layout.shape[1] + self.i
        self.n_segments = layout.shape[0] * layout.shape[1]

    def __call__(self, x: torch.Tensor, scale: float = 1.0, rpe: Optional[torch.Tensor] = None,
                 key_padding_mask: Optional[torch.Tensor] = None, attn_mask: Optional[torch.Tensor] = None,
                 key_padding_mask_mode: str = "add", attn_mask_mode: str = "add") -> torch.Tensor:
        """``x``: ``[batch, nnz, block, block]`` -> same shape, rows normalised across their block-row."""
        b = self.block
        dev = x.device
        h, i, j = self.h.to(dev), self.i.to(dev), self.j.to(dev)
        s = x.float() * scale

        if rpe is not None:
            r = rpe.float()
            r = r.expand(x.shape[0], self.spdims[0], r.shape[-2], r.shape[-1]) if r.dim() == 4 else r.expand(self.spdims[0], -1, -1)[None]
            rb = r.unflatten(-2, (-1, b)).unflatten(-1, (-1, b))            # [B,H,nbr,b,nbc,b]
            s = s + rb[:, h, i, :, j, :].movedim(0, 1) if rb.shape[0] == x.shape[0] else s + rb[0, h, i, :, j, :][None]
        if key_padding_mask is not None:
            k = key_padding_mask.float()
            if key_padding_mask_mode == "mul":
                k = torch.where(k > 0, torch.zeros_like(k), torch.full_like(k, float("-inf")))
            s = s + k.unflatten(-1, (-1, b))[:, j, :][:, :, None, :]
        if attn_mask is not None:
            a = attn_mask.float()
            if attn_mask_mode == "mul":
                a = torch.where(a > 0, torch.zeros_like(a), torch.full_like(a, float("-inf")))
            ab = a.unflatten(-2, (-1, b)).unflatten(-1, (-1, b))            # [nbr,b,nbc,b]
            s = s + ab[i, :, j, :][None]
        seg = self.segment.to(dev)
        row_max = torch.full((x.shape[0], self.n_segments, b), float("-inf"), device=dev)
        row_max = row_max.scatter_reduce(1, seg[None, :, None].expand(x.shape[0], -1, b), s.amax(-1), reduce="amax")
        m = row_max[:, seg]                                                  # [B,nnz,b]
        m = torch.where(torch.isinf(m), torch.zeros_like(m), m)
        e = torch.exp(s - m[..., None])
        den = torch.zeros(x.shape[0], self.n_segments, b, device=dev).index_add_(1, seg, e.sum(-1))
        return (e / den[:, seg][..., None].clamp_min(1e-30)).to(x.dtype)
