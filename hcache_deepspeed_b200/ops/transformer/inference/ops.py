"""Operation set of the v1 fused inference layer — the functions the reference exposes through its pybind module
``transformer_inference`` (``csrc/transformer/inference/csrc/pt_binding.cpp:1969-2042``) and wraps in
``ops/transformer/inference/op_binding/*.py``: norms, bias/activation epilogues, residual adds, rotary embedding,
QKV / vector / MLP GEMMs and ``softmax_context`` (attention over a contiguous per-layer KV cache).

Dense ``[batch, seq, hidden]`` activations with a contiguous KV cache are the LEGACY data path of the reference; the
serving path of this framework is the ragged executor with paged KV. These functions are therefore expressed with
torch ops (cuBLAS / SDPA underneath on a GPU) and exist for API parity and for models that are not lifted onto the
ragged executor.
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch
import torch.nn.functional as F


class InferenceOps:
    """Stateless helpers + a per-engine contiguous KV workspace (``inference_context.h:98-181``)."""

    def __init__(self):
        self.kv_cache = {}     # layer_id -> (k [b, h, max, d], v)
        self.seq_len = {}      # layer_id -> tokens currently cached

    # ---- workspace ---------------------------------------------------------------------------------------
    def allocate_workspace(self, layer_id: int, batch: int, heads: int, head_dim: int, max_tokens: int, dtype, device):
        if layer_id not in self.kv_cache or self.kv_cache[layer_id][0].shape[:1] != (batch, ) or \
                self.kv_cache[layer_id][0].shape[2] < max_tokens:
            k = torch.zeros(batch, heads, max_tokens, head_dim, dtype=dtype, device=device)
            self.kv_cache[layer_id] = (k, torch.zeros_like(k))
            self.seq_len[layer_id] = 0

    def reset_cache(self):
        for k in self.seq_len:
            self.seq_len[k] = 0

    def release_workspace(self):
        self.kv_cache.clear()
        self.seq_len.clear()

    # ---- norms ---------------------------------------------------------------------------------------------
    @staticmethod
    def layer_norm(x, gamma, beta, eps):
        return F.layer_norm(x, x.shape[-1:], gamma, beta, eps)

    @staticmethod
    def layer_norm_residual(x, bias, residual, gamma, beta, eps):
        return F.layer_norm(x + (bias if bias is not None else 0) + residual, x.shape[-1:], gamma, beta, eps)

    @staticmethod
    def layer_norm_residual_store_pre_ln_res(x, bias, residual, gamma, beta, eps):
        pre = x + (bias if bias is not None else 0) + residual
        return F.layer_norm(pre, x.shape[-1:], gamma, beta, eps), pre

    @staticmethod
    def rms_norm(x, gamma, eps):
        v = x.float()
        return (v * torch.rsqrt(v.pow(2).mean(-1, keepdim=True) + eps)).to(x.dtype) * gamma

    @staticmethod
    def pre_rms_norm(x, residual, gamma, eps):
        res = x + residual
        return InferenceOps.rms_norm(res, gamma, eps), res

    # ---- pointwise epilogues -------------------------------------------------------------------------------
    @staticmethod
    def bias_add(x, bias):
        return x + bias

    @staticmethod
    def bias_gelu(x, bias):
        return F.gelu(x + bias)

    @staticmethod
    def bias_relu(x, bias):
        return F.relu(x + bias)

    @staticmethod
    def bias_residual(x, residual, bias):
        return x + residual + bias

    @staticmethod
    def gated_activation(x, bias, act: str = "gelu"):
        """GEGLU / SiGLU over [.., 2F] with the halves (activation half first) split along the last dim."""
        if bias is not None:
            x = x + bias
        a, g = x.chunk(2, dim=-1)
        fn = {"gelu": F.gelu, "silu": F.silu, "relu": F.relu}[act]
        return a * fn(g)

    @staticmethod
    def residual_add_bias(hidden_state, residual, attention_output, attention_bias, final_bias, mp_size: int,
                          mlp_after_attn: bool, add_bias: bool, pre_layer_norm: bool):
        """The residual wiring of ``residual_add_bias`` (gelu.cu): sequential (GPT-2 style) or parallel (GPT-J style)."""
        if mlp_after_attn:
            out = hidden_state + residual
            if final_bias is not None:
                out = out + final_bias
            if pre_layer_norm is False and add_bias and attention_bias is not None:
                out = out + attention_bias
            return out
        tmp = residual + attention_output
        if add_bias and attention_bias is not None:
            tmp = tmp + attention_bias
        if final_bias is not None:
            tmp = tmp + final_bias
        return hidden_state + tmp / mp_size if mp_size > 1 else hidden_state + tmp

    @staticmethod
    def moe_res_matmul(residual, coef, output):
        c1, c2 = coef.chunk(2, dim=-1)
        return residual * c1 + output * c2

    # ---- rotary ---------------------------------------------------------------------------------------------
    @staticmethod
    def apply_rotary_pos_emb(q, k, rotary_dim: int, offset: int, heads: int, rotate_half: bool, theta: float = 10000.0):
        """q, k: [b, h, s, d]; positions start at ``offset``; rotate-half (NeoX/Llama) or rotate-every-two (GPT-J)."""
        s, d = q.shape[2], q.shape[3]
        rd = rotary_dim if rotary_dim > 0 else d
        inv = 1.0 / (theta ** (torch.arange(0, rd, 2, device=q.device, dtype=torch.float32) / rd))
        ang = torch.outer(torch.arange(offset, offset + s, device=q.device, dtype=torch.float32), inv)  # [s, rd/2]
        cos, sin = ang.cos()[None, None], ang.sin()[None, None]

        def rot(x):
            xr, rest = x[..., :rd].float(), x[..., rd:]
            if rotate_half:
                a, b = xr[..., :rd // 2], xr[..., rd // 2:]
                out = torch.cat([a * cos - b * sin, b * cos + a * sin], dim=-1)
            else:
                a, b = xr[..., 0::2], xr[..., 1::2]
                out = torch.stack([a * cos - b * sin, b * cos + a * sin], dim=-1).flatten(-2)
            return torch.cat([out.to(x.dtype), rest], dim=-1)

        return rot(q), rot(k)

    # ---- GEMMs ---------------------------------------------------------------------------------------------
    @staticmethod
    def qkv_gemm(x, weight, q_scale, bias, gamma, beta, eps, add_bias: bool, pre_layer_norm: bool = True,
                 rms_norm: bool = False, transposed_mode: bool = False):
        """(optional pre-norm) -> ``x @ W`` (+bias). Returns (qkv, normed input). ``weight``: [in, out] unless transposed."""
        inp = x
        if pre_layer_norm:
            inp = InferenceOps.rms_norm(x, gamma, eps) if rms_norm else F.layer_norm(x, x.shape[-1:], gamma, beta, eps)
        w = weight if transposed_mode else weight.t()
        out = F.linear(inp, w, bias if add_bias else None)
        return out, inp

    @staticmethod
    def vector_matmul(x, weight, async_op: bool = False, q_scale=None, q_int8: bool = False, transposed_mode: bool = False):
        return F.linear(x, weight if transposed_mode else weight.t())

    @staticmethod
    def mlp_gemm(x, residual, input_bias, weight_interm, weight_out, bias, gamma, beta, eps, pre_layer_norm: bool,
                 mlp_after_attn: bool, act: str = "gelu", transposed_mode: bool = False, rms_norm: bool = False,
                 skip_norm: bool = False):
        """Norm + MLP-1 (+bias, activation) + MLP-2. Sequential wiring (``mlp_after_attn``): the norm input is
        ``x + residual + input_bias`` (attention output + layer input). Parallel wiring: ``x`` is normed as given
        (``skip_norm``: ``x`` is already the normed layer input, GPT-J). Returns (output, tensor the MLP output is added
        to: the un-normed sum for pre-LN blocks, the normed one for post-LN blocks)."""
        res_add = x + residual + (input_bias if input_bias is not None else 0) if mlp_after_attn else x
        if skip_norm:
            normed = res_add
        elif rms_norm:
            normed = InferenceOps.rms_norm(res_add, gamma, eps)
        else:
            normed = F.layer_norm(res_add, x.shape[-1:], gamma, beta, eps)
        w1 = weight_interm if transposed_mode else weight_interm.t()
        w2 = weight_out if transposed_mode else weight_out.t()
        h = F.linear(normed, w1, bias)
        if act in ("gated_gelu", "gated_silu"):
            h = InferenceOps.gated_activation(h, None, act.split("_")[1])
        else:
            h = {"gelu": F.gelu, "relu": F.relu, "silu": F.silu,
                 "gelu_new": lambda t: F.gelu(t, approximate="tanh")}.get(act, F.gelu)(h)
        return F.linear(h, w2), (res_add if pre_layer_norm or not mlp_after_attn else normed)

    # ---- attention ---------------------------------------------------------------------------------------------
    def softmax_context(self, qkv: torch.Tensor, attn_mask: Optional[torch.Tensor], rotary_dim: int, rotate_half: bool,
                        rotate_every_two: bool, heads: int, num_kv: int, norm_factor: float, triangular_masking: bool,
                        local_attention: bool, window_size: int, no_masking: bool, layer_id: int, num_layers: int,
                        alibi: Optional[torch.Tensor], rope_theta: float, is_prompt: bool,
                        max_out_tokens: int = 1024) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        """qkv: [b, s, (heads + 2 * num_kv) * d]. Appends K/V to the layer's contiguous cache (reset when ``is_prompt``),
        returns (context [b, s, heads * d], k_cache view, v_cache view) — ``ds_softmax_context`` (pt_binding.cpp:440-567)."""
        b, s, width = qkv.shape
        d = width // (heads + 2 * num_kv)
        q, k, v = qkv.split([heads * d, num_kv * d, num_kv * d], dim=-1)
        q = q.view(b, s, heads, d).transpose(1, 2)
        k = k.view(b, s, num_kv, d).transpose(1, 2)
        v = v.view(b, s, num_kv, d).transpose(1, 2)
        self.allocate_workspace(layer_id, b, num_kv, d, max(max_out_tokens, s), qkv.dtype, qkv.device)
        if is_prompt:
            self.seq_len[layer_id] = 0
        past = self.seq_len[layer_id]
        if rotary_dim > 0 or rotate_half or rotate_every_two:
            q, k = self.apply_rotary_pos_emb(q, k, rotary_dim, past, heads, rotate_half, rope_theta)
        kc, vc = self.kv_cache[layer_id]
        if past + s > kc.shape[2]:
            raise RuntimeError(f"KV cache overflow: {past + s} tokens > max_out_tokens {kc.shape[2]}")
        kc[:, :, past:past + s] = k
        vc[:, :, past:past + s] = v
        total = past + s
        self.seq_len[layer_id] = total
        kk, vv = kc[:, :, :total], vc[:, :, :total]
        if num_kv != heads:
            kk, vv = kk.repeat_interleave(heads // num_kv, dim=1), vv.repeat_interleave(heads // num_kv, dim=1)
        scores = torch.matmul(q.float(), kk.float().transpose(-1, -2)) * (norm_factor * norm_factor)
        if alibi is not None:
            scores = scores + alibi.view(b, heads, 1, -1)[..., :total].float()
        if attn_mask is not None and not no_masking:
            scores = scores + attn_mask.to(scores.dtype)[..., -s:, :total] if attn_mask.dim() == 4 else \
                scores + attn_mask.to(scores.dtype).view(b, 1, 1, -1)[..., :total]
        if triangular_masking and s > 1:
            i = torch.arange(s, device=qkv.device).view(s, 1) + past
            j = torch.arange(total, device=qkv.device).view(1, total)
            masked = j > i
            if local_attention:
                masked = masked | (j < i - window_size + 1)
            scores = scores.masked_fill(masked, float("-inf"))
        ctx = torch.matmul(torch.softmax(scores, dim=-1), vv.float()).to(qkv.dtype)
        return ctx.transpose(1, 2).reshape(b, s, heads * d), kk, vv
