"""``DeepSpeedMoEInference``: the v1 fused decoder layer whose MLP is a mixture of experts — attention (shared with
:class:`DeepSpeedTransformerInference`), post-attention norm, top-k gate, this rank's slice of the experts, optional
residual MLP blended by a learned 2-way coefficient (PR-MoE), expert-parallel combination.

Reference: ``deepspeed/ops/transformer/inference/moe_inference.py:22-365`` (config :22-95, ``DeepSpeedMoEMLP`` :131-156,
layer :159-365: capacity-based einsum dispatch, ``all_to_all`` before and after the experts, ``all_gather`` of tokens
across the expert-model-parallel group). Here routing is capacity-free (inference never drops tokens): every rank of
the expert-parallel group sees the (replicated) tokens, evaluates only the tokens routed to ITS experts and the partial
outputs are summed with one all-reduce — same result, one collective instead of two all-to-alls.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F
from torch import nn

from .... import comm as dist
from .config import DeepSpeedInferenceConfig
from .ops import InferenceOps


class DeepSpeedMoEInferenceConfig(DeepSpeedInferenceConfig):

    def __init__(self, hidden_size=-1, intermediate_size=-1, heads=-1, num_hidden_layers=-1, layer_norm_eps=1e-12, local_rank=-1,
                 mp_size=1, fp16=False, bf16=False, q_int8=False, pre_layer_norm=True, stochastic_mode=False, scale_attention=True,
                 triangular_masking=True, local_attention=False, window_size=256, return_tuple=True, moe_experts=1,
                 global_experts=1, k=1, capacity_factor=1., eval_capacity_factor=1., min_capacity=1, noisy_gate_policy=None,
                 drop_tokens=True, use_rts=False, mlp_type="standard", scale_attn_by_inverse_layer_idx=False, dtype=None):
        dt = dtype if dtype is not None else torch.float16 if fp16 else torch.bfloat16 if bf16 else torch.float32
        super().__init__(hidden_size=hidden_size, intermediate_size=intermediate_size, heads=heads,
                         num_hidden_layers=num_hidden_layers, layer_norm_eps=layer_norm_eps, local_rank=local_rank,
                         mp_size=mp_size, dtype=dt, pre_layer_norm=pre_layer_norm, stochastic_mode=stochastic_mode,
                         scale_attention=scale_attention, triangular_masking=triangular_masking,
                         local_attention=local_attention, window_size=window_size, return_tuple=return_tuple,
                         scale_attn_by_inverse_layer_idx=scale_attn_by_inverse_layer_idx)
        self.moe_experts, self.global_experts, self.k = moe_experts, max(global_experts, moe_experts), k
        self.capacity_factor, self.eval_capacity_factor, self.min_capacity = capacity_factor, eval_capacity_factor, min_capacity
        self.noisy_gate_policy, self.drop_tokens, self.use_rts, self.mlp_type = noisy_gate_policy, drop_tokens, use_rts, mlp_type


class DeepSpeedMoEMLP(nn.Module):
    """One expert: ``gelu(x W1 + b1) W2 + b2`` with ``[in, out]`` weights, optionally sliced over ``mp_group``."""

    def __init__(self, config: DeepSpeedInferenceConfig, mp_group=None):
        super().__init__()
        self.config, self.mp_group = config, mp_group
        shard = dist.get_world_size(group=mp_group) if mp_group is not None and dist.is_initialized() else 1
        h, inter, dt = config.hidden_size, config.intermediate_size // shard, config.dtype
        self.inter_w = nn.Parameter(torch.empty(h, inter, dtype=dt), requires_grad=False)
        self.inter_b = nn.Parameter(torch.zeros(inter, dtype=dt), requires_grad=False)
        self.output_w = nn.Parameter(torch.empty(inter, h, dtype=dt), requires_grad=False)
        self.output_b = nn.Parameter(torch.zeros(h, dtype=dt), requires_grad=False)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        out = F.linear(F.gelu(F.linear(x, self.inter_w.t(), self.inter_b)), self.output_w.t())
        if self.mp_group is not None and dist.is_initialized() and dist.get_world_size(group=self.mp_group) > 1:
            dist.all_reduce(out, group=self.mp_group)
        return out + self.output_b


class DeepSpeedMoEInference(nn.Module):
    layer_id = 0

    def __init__(self, config: DeepSpeedMoEInferenceConfig, mp_group=None, ep_group=None, expert_mp_group=None,
                 quantize_scales=None, quantize_groups=1, merge_count=1, mlp_extra_grouping=False):
        super().__init__()
        from ....model_implementations.transformers.ds_transformer import DeepSpeedSelfAttention
        self.config = config
        config.layer_id = DeepSpeedMoEInference.layer_id
        DeepSpeedMoEInference.layer_id += 1
        self.ops = InferenceOps()
        self.attention = DeepSpeedSelfAttention(config, self.ops, mp_group)
        h, dt = config.hidden_size, config.dtype
        self.attn_nw = nn.Parameter(torch.ones(h, dtype=dt), requires_grad=False)
        self.attn_nb = nn.Parameter(torch.zeros(h, dtype=dt), requires_grad=False)
        self.norm_w = nn.Parameter(torch.ones(h, dtype=dt), requires_grad=False)
        self.norm_b = nn.Parameter(torch.zeros(h, dtype=dt), requires_grad=False)
        if config.mlp_type == "residual":
            self.res_mlp = DeepSpeedMoEMLP(config, mp_group)
            self.res_coef = nn.Parameter(torch.zeros(h, 2, dtype=dt), requires_grad=False)
        self.mlp = nn.ModuleList(DeepSpeedMoEMLP(config, expert_mp_group) for _ in range(config.moe_experts))
        self.gate_w = nn.Parameter(torch.empty(config.global_experts, h, dtype=torch.float32), requires_grad=False)
        self.mp_group, self.ep_group, self.expert_mp_group = mp_group, ep_group, expert_mp_group

    # ---- expert-parallel bookkeeping ---------------------------------------------------------------------------------
    def _ep(self):
        if self.ep_group is not None and dist.is_initialized():
            return dist.get_rank(group=self.ep_group), dist.get_world_size(group=self.ep_group)
        return 0, 1

    def moe_forward(self, x: torch.Tensor) -> torch.Tensor:
        """x: [tokens, hidden] (replicated over the expert-parallel group) -> combined expert outputs."""
        c = self.config
        logits = F.linear(x.float(), self.gate_w)
        probs = torch.softmax(logits, dim=-1)
        top_p, top_e = probs.topk(c.k, dim=-1)
        if c.k > 1:
            top_p = top_p / top_p.sum(dim=-1, keepdim=True)
        rank, size = self._ep()
        first = rank * c.moe_experts                       # this rank owns global experts [first, first + moe_experts)
        out = torch.zeros_like(x, dtype=torch.float32)
        for local, expert in enumerate(self.mlp):
            hit = top_e == first + local                   # [tokens, k]
            rows = hit.any(dim=-1).nonzero(as_tuple=True)[0]
            if rows.numel() == 0:
                continue
            w = (top_p * hit).sum(dim=-1)[rows]
            out.index_add_(0, rows, expert(x[rows]).float() * w[:, None])
        if size > 1:
            dist.all_reduce(out, group=self.ep_group)
        return out.to(x.dtype)

    def forward(self, input, input_mask=None, attention_mask=None, head_mask=None, layer_past=None, get_key_value=False,
                get_present=False, encoder_output=None, enc_dec_attn_mask=None, encoder_hidden_states=None,
                encoder_attention_mask=None, use_cache=False, output_attentions=False, alibi=None, **kwargs):
        c = self.config
        get_present = get_present or get_key_value or use_cache
        mask = input_mask if input_mask is not None else attention_mask
        attn_out, k, v, _, _ = self.attention(input, mask, self.norm_w, self.norm_b, alibi,
                                              is_prompt=kwargs.get("is_prompt", layer_past is None or input.shape[1] > 1))
        residual = input + attn_out + self.attention.attn_ob
        normed = self.ops.layer_norm(residual, self.attn_nw, self.attn_nb, c.layer_norm_eps)
        flat = normed.reshape(-1, c.hidden_size)
        moe = self.moe_forward(flat)
        if c.mlp_type == "residual":
            # PR-MoE: a dense MLP runs beside the experts; a learned per-token softmax mixes the two
            dense = self.res_mlp(flat)
            coef = torch.softmax(F.linear(flat, self.res_coef.t()), dim=-1)
            moe = dense * coef[:, 0:1] + moe * coef[:, 1:2]
        out = residual + moe.view_as(residual)
        if not c.pre_layer_norm:
            out = self.ops.layer_norm(out, self.norm_w, self.norm_b, c.layer_norm_eps)
        if get_present:
            return out, (k, v)
        return (out, ) if c.return_tuple else out
