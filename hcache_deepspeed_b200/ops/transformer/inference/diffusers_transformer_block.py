"""Fused replacement of a diffusers ``BasicTransformerBlock``
(``ops/transformer/inference/diffusers_transformer_block.py:18-110``): LN -> self-attn -> LN -> cross-attn -> LN -> GEGLU
feed-forward, each with its residual; the weights are taken over from the module it replaces."""
from __future__ import annotations

import torch
import torch.nn.functional as F
from torch import nn

from .diffusers_2d_transformer import Diffusers2DTransformerConfig
from .diffusers_attention import DeepSpeedDiffusersAttention


class _AttnCfg:
    def __init__(self, hidden_size, heads, dtype):
        self.hidden_size, self.heads, self.dtype = hidden_size, heads, dtype


def _fused_attention(attn: nn.Module, cross: bool) -> DeepSpeedDiffusersAttention:
    w = attn.to_q.weight
    out = DeepSpeedDiffusersAttention(_AttnCfg(w.shape[0], attn.heads, w.dtype))
    if cross:
        out.attn_qw.data = attn.to_q.weight.detach().t().contiguous()
        out.attn_kw = nn.Parameter(attn.to_k.weight.detach().t().contiguous(), requires_grad=False)
        out.attn_vw = nn.Parameter(attn.to_v.weight.detach().t().contiguous(), requires_grad=False)
    else:
        out.attn_qkvw.data = torch.cat([attn.to_q.weight, attn.to_k.weight, attn.to_v.weight], dim=0).detach().t().contiguous()
    proj = attn.to_out[0]
    out.attn_ow.data = proj.weight.detach().t().contiguous()
    out.do_out_bias = proj.bias is not None
    if proj.bias is not None:
        out.attn_ob.data = proj.bias.detach().clone()
    return out


class DeepSpeedDiffusersTransformerBlock(nn.Module):

    def __init__(self, equivalent_module: nn.Module, config: Diffusers2DTransformerConfig):
        super().__init__()
        self.config = config
        m = equivalent_module
        p = lambda t: nn.Parameter(t.detach().clone(), requires_grad=False)  # noqa: E731
        self.norm1_g, self.norm1_b, self.norm1_eps = p(m.norm1.weight), p(m.norm1.bias), m.norm1.eps
        self.norm2_g, self.norm2_b, self.norm2_eps = p(m.norm2.weight), p(m.norm2.bias), m.norm2.eps
        self.norm3_g, self.norm3_b, self.norm3_eps = p(m.norm3.weight), p(m.norm3.bias), m.norm3.eps
        self.attn_1 = _fused_attention(m.attn1, cross=False)
        self.attn_2 = _fused_attention(m.attn2, cross=True)
        ff1, ff2 = m.ff.net[0].proj, m.ff.net[2]
        self.ff1_w, self.ff1_b = p(ff1.weight.t().contiguous()), p(ff1.bias)
        self.ff2_w, self.ff2_b = p(ff2.weight.t().contiguous()), p(ff2.bias)

    def forward(self, hidden_states, context=None, timestep=None, **kwargs):
        context = kwargs.get("encoder_hidden_states", context)
        h = hidden_states
        n = F.layer_norm(h, h.shape[-1:], self.norm1_g, self.norm1_b, self.norm1_eps)
        h = h + self.attn_1(n)
        n = F.layer_norm(h, h.shape[-1:], self.norm2_g, self.norm2_b, self.norm2_eps)
        h = h + self.attn_2(n, context=context)
        n = F.layer_norm(h, h.shape[-1:], self.norm3_g, self.norm3_b, self.norm3_eps)
        a, gate = (n @ self.ff1_w + self.ff1_b).chunk(2, dim=-1)
        return h + (a * F.gelu(gate)) @ self.ff2_w + self.ff2_b
