"""In-place int8 quantisation of the linear weights of transformer layers
(``deepspeed/module_inject/module_quantize.py:9-70``): weights become int8 with a per-tensor scale (returned), for
kernels that consume 8-bit weights; the walk is ``quantize_module(model, layer_class, fn)``."""
from __future__ import annotations

from typing import Callable, List

import torch


def _quantize_weight(weight: torch.Tensor):
    scale = weight.detach().abs().max().clamp_min(1e-8) / 127.0
    q = torch.clamp(torch.round(weight.detach() / scale), -128, 127).to(torch.int8)
    return q, float(scale)


def quantize_transformer_layer(orig_layer_impl, model: torch.nn.Module, megatron: bool = False, preln: bool = False) -> List[float]:
    """Quantise the attention / MLP projection weights of every ``orig_layer_impl`` in ``model``. Returns the scales in
    visiting order (one per quantised matrix)."""
    scales: List[float] = []

    def swap(lin: torch.nn.Module):
        if lin is None or not hasattr(lin, "weight"):
            return
        q, s = _quantize_weight(lin.weight)
        lin.weight = torch.nn.Parameter(q, requires_grad=False)
        scales.append(s)

    def megatron_layer(layer):
        swap(layer.attention.query_key_value)
        swap(layer.attention.dense)
        swap(layer.mlp.dense_h_to_4h)
        swap(layer.mlp.dense_4h_to_h)

    def bert_layer(layer):
        att = layer.attention
        for name in ("query", "key", "value"):
            swap(getattr(att.self, name))
        swap(att.output.dense)
        inter = getattr(layer, "intermediate", None)
        swap(inter.dense_act if (preln and hasattr(inter, "dense_act")) else inter.dense)
        swap(layer.output.dense)

    def fn(child):
        (megatron_layer if megatron else bert_layer)(child)
        return child

    quantize_module(model, orig_layer_impl, fn)
    return scales


def quantize_module(model: torch.nn.Module, orig_class, quantize_fn: Callable) -> torch.nn.Module:
    return _quantize_module(model, {orig_class: quantize_fn})


def _quantize_module(model: torch.nn.Module, policies: dict) -> torch.nn.Module:
    for name, child in model.named_children():
        fn = policies.get(child.__class__)
        if fn is not None:
            setattr(model, name, fn(child))
        else:
            _quantize_module(child, policies)
    return model
