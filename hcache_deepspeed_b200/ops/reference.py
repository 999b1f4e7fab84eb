"""Plain-PyTorch restatements of every device op (fp32 math, device-agnostic).

Two uses: (1) the numerics oracle the GPU kernel tests compare against, (2) the CPU execution
backend of the engines (``init_inference`` on CPU, CPU-only CI). They are never used on a CUDA
device by the framework itself.

The math follows the reference kernels cited in ``include/hds_b200.h``; differences that matter for
tolerances: normalisation statistics and both multiplies are fp32 here (the reference truncates the
reciprocal to the I/O dtype, ``rms_norm_cuda.cu:57,73-74``), RoPE angles are ``pos * inv_freq`` in fp32.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn.functional as F


def apply_act(x: torch.Tensor, act: int) -> torch.Tensor:
    if act == 1:
        return F.gelu(x)
    if act == 2:
        return F.relu(x)
    if act == 3:
        return F.silu(x)
    if act == 4:
        return F.gelu(x, approximate="tanh")
    return x


def gemm(a: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None, act: int = 0, gated: bool = False,
         out_dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    y = a.float() @ w.float().t()
    if bias is not None:
        y = y + bias.float()
    if gated:
        y = apply_act(y[:, 0::2], act) * y[:, 1::2]
    else:
        y = apply_act(y, act)
    return y.to(out_dtype or a.dtype)


def norm(x: torch.Tensor, gamma: torch.Tensor, beta: Optional[torch.Tensor] = None, eps: float = 1e-5,
         layer_norm: bool = False) -> torch.Tensor:
    xf = x.float()
    if layer_norm:
        mean = xf.mean(-1, keepdim=True)
        var = ((xf - mean) ** 2).mean(-1, keepdim=True)
        y = (xf - mean) * torch.rsqrt(var + eps) * gamma.float()
        if beta is not None:
            y = y + beta.float()
    else:
        y = xf * torch.rsqrt((xf * xf).mean(-1, keepdim=True) + eps) * gamma.float()
    return y.to(x.dtype)


def rope_rotate_half(x: torch.Tensor, pos: torch.Tensor, inv_freq: torch.Tensor, rotary_dim: int) -> torch.Tensor:
    """x: [T, heads, d]; rotate-half RoPE over the first ``rotary_dim`` dims (HF Llama convention)."""
    xf = x.float()
    half = rotary_dim // 2
    ang = pos.float()[:, None] * inv_freq.float()[None, :half]  # [T, half]
    cos, sin = ang.cos()[:, None, :], ang.sin()[:, None, :]
    lo, hi = xf[..., :half], xf[..., half:rotary_dim]
    out = xf.clone()
    out[..., :half] = lo * cos - hi * sin
    out[..., half:rotary_dim] = hi * cos + lo * sin
    return out.to(x.dtype)


def kv_append(kv_cache_layer: torch.Tensor, k: torch.Tensor, v: torch.Tensor, tok_slot: torch.Tensor) -> None:
    """kv_cache_layer: [blocks, 2, kvh, bs, d]; k, v: [T, kvh, d]; tok_slot: [T] (negative = skip)."""
    bs = kv_cache_layer.shape[3]
    valid = tok_slot >= 0
    slots = tok_slot[valid].long()
    blk, off = slots // bs, slots % bs
    kv_cache_layer[blk, 0, :, off, :] = k[valid].to(kv_cache_layer.dtype)
    kv_cache_layer[blk, 1, :, off, :] = v[valid].to(kv_cache_layer.dtype)


def gather_kv(kv_cache_layer: torch.Tensor, block_table_row: torch.Tensor, kv_len: int):
    """Dense K, V ([kv_len, kvh, d]) of one sequence out of the paged cache."""
    bs = kv_cache_layer.shape[3]
    nb = (kv_len + bs - 1) // bs
    blocks = kv_cache_layer[block_table_row[:nb].long()]  # [nb, 2, kvh, bs, d]
    k = blocks[:, 0].permute(0, 2, 1, 3).reshape(nb * bs, blocks.shape[2], blocks.shape[4])[:kv_len]
    v = blocks[:, 1].permute(0, 2, 1, 3).reshape(nb * bs, blocks.shape[2], blocks.shape[4])[:kv_len]
    return k, v


def paged_attention(q: torch.Tensor, kv_cache_layer: torch.Tensor, block_tables: torch.Tensor, items: torch.Tensor,
                    n_q_heads: int, n_kv_heads: int, scale: float, out: Optional[torch.Tensor] = None,
                    alibi: Optional[torch.Tensor] = None, window: int = 0) -> torch.Tensor:
    """q: [T, n_q_heads, d]; items: int32 [n, 4] rows (seq, q_row0, n_q, pos0). Causal over absolute positions.
    ``alibi``: [n_q_heads] slopes, score += slope * (key position - query position). ``window`` > 0: a query at position p
    sees the keys (p - window, p] only."""
    T, _, d = q.shape
    if out is None:
        out = torch.zeros_like(q)
    G = n_q_heads // n_kv_heads
    for seq, q_row0, n_q, pos0 in items.tolist():
        kv_len = pos0 + n_q
        k, v = gather_kv(kv_cache_layer, block_tables[seq], kv_len)
        k = k.float().repeat_interleave(G, dim=1)  # [kv_len, n_q_heads, d]
        v = v.float().repeat_interleave(G, dim=1)
        qq = q[q_row0:q_row0 + n_q].float()  # [n_q, H, d]
        s = torch.einsum("qhd,khd->hqk", qq, k) * scale
        kv_pos = torch.arange(kv_len, device=q.device)[None, None, :]
        q_pos = (pos0 + torch.arange(n_q, device=q.device))[None, :, None]
        if alibi is not None:
            s = s + alibi.float().to(s.device)[:, None, None] * (kv_pos - q_pos).float()
        s = s.masked_fill(kv_pos > q_pos, float("-inf"))
        if window:
            s = s.masked_fill(kv_pos <= q_pos - window, float("-inf"))
        p = torch.softmax(s, dim=-1)
        out[q_row0:q_row0 + n_q] = torch.einsum("hqk,khd->qhd", p, v).to(out.dtype)
    return out


def embed(ids: torch.Tensor, weight: torch.Tensor, pos_weight: Optional[torch.Tensor] = None,
          tok_pos: Optional[torch.Tensor] = None, pos_offset: int = 0) -> torch.Tensor:
    out = weight[ids.long().clamp_min(0)].float()
    out = torch.where((ids >= 0)[:, None], out, torch.zeros_like(out))
    if pos_weight is not None:
        p = (tok_pos.long() + pos_offset).clamp(0, pos_weight.shape[0] - 1)
        out = (out.to(weight.dtype).float() + pos_weight[p].float())
    return out.to(weight.dtype)


def default_inv_freq(rotary_dim: int, theta: float = 10000.0, device=None) -> torch.Tensor:
    return 1.0 / (theta ** (torch.arange(0, rotary_dim, 2, dtype=torch.float32, device=device) / rotary_dim))


def alibi_slopes(n_heads: int) -> torch.Tensor:
    """ALiBi head slopes (Press et al.): a geometric sequence for the largest power of two <= n_heads, interleaved with
    the odd terms of the next power's sequence for the remaining heads. fp32 [n_heads]."""
    base = 2 ** math.floor(math.log2(n_heads))
    first = [2.0 ** (-8.0 * (i + 1) / base) for i in range(base)]
    if base < n_heads:
        first += [2.0 ** (-4.0 * (2 * i + 1) / base) for i in range(n_heads - base)]
    return torch.tensor(first, dtype=torch.float32)


def attention_scale(head_dim: int) -> float:
    return 1.0 / math.sqrt(head_dim)
