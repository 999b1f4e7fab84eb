"""Training-kernel injection (``deepspeed/module_inject/inject.py``): swap HF BERT layers for the fused training
transformer layer (:class:`hcache_deepspeed_b200.ops.transformer.DeepSpeedTransformerLayer`), copying the weights."""
from __future__ import annotations

import copy

import torch


def module_inject(layer_obj, model: torch.nn.Module, config, micro_batch_size: int, max_seq_length: int, seed: int,
                  preln: bool, fp16: bool = True) -> torch.nn.Module:
    from ..ops.transformer import DeepSpeedTransformerConfig, DeepSpeedTransformerLayer
    for name, child in model.named_children():
        if isinstance(child, layer_obj):
            cfg = DeepSpeedTransformerConfig(batch_size=micro_batch_size, hidden_size=config.hidden_size,
                                             heads=config.num_attention_heads, attn_dropout_ratio=config.attention_probs_dropout_prob,
                                             hidden_dropout_ratio=config.hidden_dropout_prob, num_hidden_layers=config.num_hidden_layers,
                                             initializer_range=config.initializer_range, seed=seed, fp16=fp16, pre_layer_norm=preln,
                                             intermediate_size=getattr(config, "intermediate_size", 4 * config.hidden_size))
            new = DeepSpeedTransformerLayer(cfg)
            att, out = child.attention, child.attention.output
            q, k, v = att.self.query, att.self.key, att.self.value
            with torch.no_grad():
                new.attn_qkvw.copy_(torch.cat([q.weight, k.weight, v.weight], 0))
                new.attn_qkvb.copy_(torch.cat([q.bias, k.bias, v.bias], 0))
                new.attn_ow.copy_(out.dense.weight)
                new.attn_ob.copy_(out.dense.bias)
                ln1 = child.PostAttentionLayerNorm if preln and hasattr(child, "PostAttentionLayerNorm") else out.LayerNorm
                new.attn_nw.copy_(ln1.weight)
                new.attn_nb.copy_(ln1.bias)
                inter = child.intermediate.dense_act if hasattr(child.intermediate, "dense_act") else child.intermediate.dense
                new.inter_w.copy_(inter.weight)
                new.inter_b.copy_(inter.bias)
                new.output_w.copy_(child.output.dense.weight)
                new.output_b.copy_(child.output.dense.bias)
                ln2 = child.PreAttentionLayerNorm if preln and hasattr(child, "PreAttentionLayerNorm") else child.output.LayerNorm
                new.norm_w.copy_(ln2.weight)
                new.norm_b.copy_(ln2.bias)
            setattr(model, name, copy.deepcopy(new))
        else:
            module_inject(layer_obj, child, config, micro_batch_size, max_seq_length, seed, preln, fp16)
    return model
