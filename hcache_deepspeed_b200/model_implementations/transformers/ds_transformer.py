"""``DeepSpeedTransformerInference`` — the v1 fused decoder layer: attention (QKV GEMM -> rotary -> softmax over a
contiguous KV cache -> output GEMM -> TP all-reduce) and MLP (norm -> GEMM -> activation -> GEMM -> TP all-reduce),
for dense ``[batch, seq, hidden]`` inputs. Variants of the reference (``ds_bert/bloom/gpt/llama2/megatron_gpt/opt``)
differ only in config flags (pre/post norm, masking, rotary style, gated activation), so they are aliases here.
Reference: ``deepspeed/model_implementations/transformers/ds_transformer.py:20-191``,
``ops/transformer/inference/ds_attention.py:18-290``, ``ds_mlp.py:13-125``.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
from torch import nn

from ... import comm as dist
from ...ops.transformer.inference import DeepSpeedInferenceConfig, InferenceOps


class DeepSpeedSelfAttention(nn.Module):

    def __init__(self, config: DeepSpeedInferenceConfig, ops: InferenceOps, mp_group=None):
        super().__init__()
        c = self.config = config
        self.ops, self.mp_group = ops, mp_group
        h = c.hidden_size
        self.heads_local = c.heads // c.mp_size
        self.kv_local = max(1, c.num_kv // c.mp_size)
        self.head_dim = h // c.heads
        qkv_out = (self.heads_local + 2 * self.kv_local) * self.head_dim
        dt = c.dtype
        self.attn_qkvw = nn.Parameter(torch.empty(h, qkv_out, dtype=dt), requires_grad=False)
        self.attn_qkvb = nn.Parameter(torch.zeros(qkv_out, dtype=dt), requires_grad=False)
        self.attn_ow = nn.Parameter(torch.empty(self.heads_local * self.head_dim, h, dtype=dt), requires_grad=False)
        self.attn_ob = nn.Parameter(torch.zeros(h, dtype=dt), requires_grad=False)
        self.norm_factor = math.sqrt(math.sqrt(self.head_dim)) ** -1 if c.scale_attention else 1.0
        if c.scale_attn_by_inverse_layer_idx:
            # scores are multiplied by norm_factor ** 2: an extra 1 / (layer index + 1), index counted within the model
            self.norm_factor *= (getattr(c, "layer_index", c.layer_id) + 1) ** -0.5

    def forward(self, x, input_mask=None, norm_w=None, norm_b=None, alibi=None, is_prompt: Optional[bool] = None):
        c = self.config
        qkv, normed = self.ops.qkv_gemm(x, self.attn_qkvw, None, self.attn_qkvb, norm_w, norm_b, c.layer_norm_eps,
                                        add_bias=True, pre_layer_norm=c.pre_layer_norm, rms_norm=c.norm_type == "rms_norm")
        if is_prompt is None:
            is_prompt = x.shape[1] > 1
        ctx, k, v = self.ops.softmax_context(qkv, input_mask, c.rotary_dim, c.rotate_half, c.rotate_every_two and c.rotary_dim > 0,
                                             self.heads_local, self.kv_local, self.norm_factor, c.triangular_masking,
                                             c.local_attention, c.window_size, input_mask is None, c.layer_id,
                                             c.num_hidden_layers, alibi, c.rope_theta, is_prompt, c.max_out_tokens)
        out = self.ops.vector_matmul(ctx, self.attn_ow)
        if c.mp_size > 1 and dist.is_initialized() and self.mp_group is not None:
            dist.all_reduce(out, group=self.mp_group)
        return out, k, v, ctx, normed


class DeepSpeedMLP(nn.Module):

    def __init__(self, config: DeepSpeedInferenceConfig, ops: InferenceOps, mp_group=None):
        super().__init__()
        c = self.config = config
        self.ops, self.mp_group = ops, mp_group
        h, inter = c.hidden_size, c.intermediate_size // c.mp_size
        gated = c.mlp_act_func_type in ("gated_gelu", "gated_silu")
        dt = c.dtype
        self.attn_nw = nn.Parameter(torch.ones(h, dtype=dt), requires_grad=False)
        self.attn_nb = nn.Parameter(torch.zeros(h, dtype=dt), requires_grad=False) if c.norm_type != "rms_norm" else None
        self.inter_w = nn.Parameter(torch.empty(h, inter * (2 if gated else 1), dtype=dt), requires_grad=False)
        self.inter_b = nn.Parameter(torch.zeros(inter * (2 if gated else 1), dtype=dt), requires_grad=False)
        self.output_w = nn.Parameter(torch.empty(inter, h, dtype=dt), requires_grad=False)
        self.output_b = nn.Parameter(torch.zeros(h, dtype=dt), requires_grad=False)

    def forward(self, attn_out, residual, attn_bias, skip_norm: bool = False):
        c = self.config
        out, res_add = self.ops.mlp_gemm(attn_out, residual, attn_bias, self.inter_w, self.output_w, self.inter_b, self.attn_nw,
                                         self.attn_nb, c.layer_norm_eps, c.pre_layer_norm, c.mlp_after_attn,
                                         act=c.mlp_act_func_type, rms_norm=c.norm_type == "rms_norm", skip_norm=skip_norm)
        if c.mp_size > 1 and dist.is_initialized() and self.mp_group is not None:
            dist.all_reduce(out, group=self.mp_group)
        return out, res_add


class DeepSpeedTransformerInference(nn.Module):
    layer_id = 0
    workspace: Optional[InferenceOps] = None

    def __init__(self, config: DeepSpeedInferenceConfig, mp_group=None, quantize_scales=None, quantize_groups=1,
                 merge_count=1, mlp_extra_grouping=False):
        super().__init__()
        self.config = config
        self.config.layer_id = DeepSpeedTransformerInference.layer_id
        DeepSpeedTransformerInference.layer_id += 1
        if DeepSpeedTransformerInference.workspace is None:
            DeepSpeedTransformerInference.workspace = InferenceOps()
        ops = DeepSpeedTransformerInference.workspace
        self.attention = DeepSpeedSelfAttention(self.config, ops, mp_group)
        self.mlp = DeepSpeedMLP(self.config, ops, mp_group)
        dt = config.dtype
        self.norm_w = nn.Parameter(torch.ones(config.hidden_size, dtype=dt), requires_grad=False)
        self.norm_b = nn.Parameter(torch.zeros(config.hidden_size, dtype=dt), requires_grad=False) \
            if config.norm_type != "rms_norm" else None

    @classmethod
    def reset_cache(cls):
        if cls.workspace is not None:
            cls.workspace.reset_cache()

    @classmethod
    def release(cls):
        cls.layer_id = 0
        if cls.workspace is not None:
            cls.workspace.release_workspace()
        cls.workspace = None

    def forward(self, input=None, input_mask=None, attention_mask=None, attn_mask=None, head_mask=None, layer_past=None,
                get_key_value=False, get_present=False, encoder_output=None, enc_dec_attn_mask=None, x=None,
                encoder_hidden_states=None, encoder_attention_mask=None, use_cache=False, alibi=None,
                output_attentions=False, layer_head_mask=None, past_key_value=None, **kwargs):
        if x is not None:
            input = x
        if "hidden_states" in kwargs:
            input = kwargs["hidden_states"]
        mask = input_mask if input_mask is not None else (attention_mask if attention_mask is not None else attn_mask)
        get_present = get_present or get_key_value or use_cache
        c = self.config
        attn_out, k, v, ctx, normed = self.attention(input, mask, self.norm_w, self.norm_b, alibi,
                                                     is_prompt=kwargs.get("is_prompt", (layer_past is None and past_key_value is None)
                                                                          or input.shape[1] > 1))
        ob, fb = self.attention.attn_ob, self.mlp.output_b
        if c.mlp_after_attn:
            # sequential: h = x + attn(x); out = h + mlp(norm(h))   (post-LN: norms sit after each residual add instead)
            mlp_out, res_add = self.mlp(attn_out, input, ob)
            out = mlp_out + res_add + fb
            if not c.pre_layer_norm:
                out = self.mlp.ops.layer_norm(out, self.norm_w, self.norm_b, c.layer_norm_eps)
        else:
            # parallel: out = x + attn(norm1(x)) + mlp(norm2(x)); one shared norm when the MLP has none of its own (GPT-J)
            own_norm = getattr(self.mlp, "has_own_norm", True)
            mlp_out, _ = self.mlp(input if own_norm else normed, None, None, skip_norm=not own_norm)
            out = input + attn_out + ob + mlp_out + fb
        if get_present:
            return (out, (k, v))
        return (out, ) if c.return_tuple else out


DeepSpeedBERTInference = DeepSpeedBloomInference = DeepSpeedGPTInference = DeepSpeedTransformerInference
DeepSpeedLlama2Inference = DeepSpeedMegatronGPTInference = DeepSpeedOPTInference = DeepSpeedTransformerInference
