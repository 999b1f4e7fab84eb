"""BERT-style fused transformer layer for TRAINING (``DeepSpeedTransformerLayer``): self-attention + MLP with pre- or
post-LayerNorm, dropout, GELU, optional stochastic mode. The reference implements it as one hand-fused CUDA op
(``csrc/transformer/*.cu``, 6.6k lines of Volta-era SIMT kernels); on Blackwell the attention goes through
``scaled_dot_product_attention`` (flash kernels) and the rest through cuBLAS + elementwise fusions — this layer is not
on any path the framework optimises, it exists so models written against the API keep working.
Reference: ``deepspeed/ops/transformer/transformer.py:18-420``."""
from __future__ import annotations

import copy
import json
import math

import torch
from torch import nn


class TransformerConfig:
    def __init__(self, batch_size, hidden_size, intermediate_size, heads, attn_dropout_ratio, hidden_dropout_ratio,
                 num_hidden_layers, initializer_range):
        self.layer_id = -1
        self.batch_size, self.hidden_size, self.intermediate_size, self.heads = batch_size, hidden_size, intermediate_size, heads
        self.attn_dropout_ratio, self.hidden_dropout_ratio = attn_dropout_ratio, hidden_dropout_ratio
        self.num_hidden_layers, self.initializer_range = num_hidden_layers, initializer_range


class DeepSpeedTransformerConfig(TransformerConfig):

    def __init__(self, batch_size=-1, hidden_size=-1, intermediate_size=-1, heads=-1, attn_dropout_ratio=-1,
                 hidden_dropout_ratio=-1, num_hidden_layers=-1, initializer_range=-1, layer_norm_eps=1e-12, local_rank=-1,
                 seed=-1, fp16=False, pre_layer_norm=True, normalize_invertible=False, gelu_checkpoint=False,
                 adjust_init_range=True, attn_dropout_checkpoint=False, stochastic_mode=False, return_tuple=False,
                 training=True):
        super().__init__(batch_size, hidden_size, intermediate_size if intermediate_size > 0 else 4 * hidden_size, heads,
                         attn_dropout_ratio, hidden_dropout_ratio, num_hidden_layers, initializer_range)
        self.fp16, self.pre_layer_norm, self.local_rank, self.seed = fp16, pre_layer_norm, local_rank, seed
        self.normalize_invertible, self.gelu_checkpoint, self.adjust_init_range = normalize_invertible, gelu_checkpoint, adjust_init_range
        self.test_gemm = False
        self.layer_norm_eps = layer_norm_eps
        self.training, self.is_grad_enabled = training, True
        self.attn_dropout_checkpoint, self.stochastic_mode, self.return_tuple = attn_dropout_checkpoint, stochastic_mode, return_tuple

    @classmethod
    def from_dict(cls, json_object):
        c = cls()
        for k, v in json_object.items():
            c.__dict__[k] = v
        return c

    @classmethod
    def from_json_file(cls, json_file):
        with open(json_file, "r", encoding="utf-16") as r:
            return cls.from_dict(json.loads(r.read()))


class DeepSpeedTransformerLayer(nn.Module):
    layer_id = 0

    def __init__(self, config: DeepSpeedTransformerConfig, initial_weights=None, initial_biases=None):
        super().__init__()
        self.config = copy.copy(config)
        self.config.layer_id = DeepSpeedTransformerLayer.layer_id
        DeepSpeedTransformerLayer.layer_id += 1
        h, i = config.hidden_size, config.intermediate_size
        dt = torch.half if config.fp16 else torch.float32
        self.attn_qkvw = nn.Parameter(torch.empty(3 * h, h, dtype=dt))
        self.attn_qkvb = nn.Parameter(torch.zeros(3 * h, dtype=dt))
        self.attn_ow = nn.Parameter(torch.empty(h, h, dtype=dt))
        self.attn_ob = nn.Parameter(torch.zeros(h, dtype=dt))
        self.attn_nw, self.attn_nb = nn.Parameter(torch.ones(h, dtype=dt)), nn.Parameter(torch.zeros(h, dtype=dt))
        self.inter_w, self.inter_b = nn.Parameter(torch.empty(i, h, dtype=dt)), nn.Parameter(torch.zeros(i, dtype=dt))
        self.output_w, self.output_b = nn.Parameter(torch.empty(h, i, dtype=dt)), nn.Parameter(torch.zeros(h, dtype=dt))
        self.norm_w, self.norm_b = nn.Parameter(torch.ones(h, dtype=dt)), nn.Parameter(torch.zeros(h, dtype=dt))
        if initial_weights is None:
            self.init_transformer_weights(config.adjust_init_range)
        else:
            q, k, v = initial_weights[0].data, initial_weights[1].data, initial_weights[2].data
            self.attn_qkvw.data = torch.cat((q, k, v)).to(dt)
            self.attn_qkvb.data = torch.cat([b.data for b in initial_biases[:3]]).to(dt)
            for p, src in zip((self.attn_ow, self.attn_nw, self.inter_w, self.output_w, self.norm_w), initial_weights[3:8]):
                p.data = src.data.to(dt)
            for p, src in zip((self.attn_ob, self.attn_nb, self.inter_b, self.output_b, self.norm_b), initial_biases[3:8]):
                p.data = src.data.to(dt)

    def init_transformer_weights(self, adjust_init_range: bool = False):
        std = self.config.initializer_range if self.config.initializer_range > 0 else 0.02
        out_std = std
        if adjust_init_range and self.config.num_hidden_layers > 0 and self.config.local_rank <= 0:
            out_std = std / math.sqrt(2.0 * self.config.num_hidden_layers)
        for p in (self.attn_qkvw, self.inter_w):
            p.data.normal_(0.0, std)
        for p in (self.attn_ow, self.output_w):
            p.data.normal_(0.0, out_std)

    def forward(self, hidden_states, attention_mask=None, head_mask=None, layer_head_mask=None, encoder_hidden_states=None,
                encoder_attention_mask=None, past_key_value=None, output_attentions=False, grads=None):
        c = self.config
        F = torch.nn.functional
        x = hidden_states
        b, s, h = x.shape
        nh = c.heads
        p_attn = c.attn_dropout_ratio if (self.training and c.attn_dropout_ratio > 0) else 0.0
        p_hid = c.hidden_dropout_ratio if (self.training and c.hidden_dropout_ratio > 0) else 0.0

        def attn_block(inp):
            qkv = F.linear(inp, self.attn_qkvw, self.attn_qkvb).view(b, s, 3, nh, h // nh).permute(2, 0, 3, 1, 4)
            mask = attention_mask
            if mask is not None and mask.dtype != torch.bool:
                mask = mask.to(inp.dtype)
            ctx = F.scaled_dot_product_attention(qkv[0], qkv[1], qkv[2], attn_mask=mask, dropout_p=p_attn)
            return F.dropout(F.linear(ctx.transpose(1, 2).reshape(b, s, h), self.attn_ow, self.attn_ob), p_hid, self.training)

        def mlp_block(inp):
            inter = F.gelu(F.linear(inp, self.inter_w, self.inter_b))
            return F.dropout(F.linear(inter, self.output_w, self.output_b), p_hid, self.training)

        if c.pre_layer_norm:
            x = x + attn_block(F.layer_norm(x, (h, ), self.attn_nw, self.attn_nb, c.layer_norm_eps))
            x = x + mlp_block(F.layer_norm(x, (h, ), self.norm_w, self.norm_b, c.layer_norm_eps))
        else:
            x = F.layer_norm(x + attn_block(x), (h, ), self.attn_nw, self.attn_nb, c.layer_norm_eps)
            x = F.layer_norm(x + mlp_block(x), (h, ), self.norm_w, self.norm_b, c.layer_norm_eps)
        return (x, ) if c.return_tuple else x
