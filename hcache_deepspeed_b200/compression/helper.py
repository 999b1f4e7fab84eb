"""Module surgery helpers of the compression library (``deepspeed/compression/helper.py:13-322``)."""
from __future__ import annotations

import re
from typing import List

import torch
from torch import nn

from .compress import _set, _to_compress_layer


def recursive_getattr(model: nn.Module, module_name: str):
    out = model
    for part in module_name.split("."):
        out = getattr(out, part)
    return out


def recursive_setattr(model: nn.Module, module_name: str, module: nn.Module) -> None:
    _set(model, module_name, module)


def module_replacement(model: nn.Module, module_name: str, compression_technique=None, mpu=None) -> nn.Module:
    """Swap one Linear / Conv2d / Embedding / BatchNorm for its compressible twin (weights carried over)."""
    old = recursive_getattr(model, module_name)
    new = _to_compress_layer(old)
    if new is None:
        return old
    recursive_setattr(model, module_name, new)
    return new


def is_module_compressible(module: nn.Module, mpu=None) -> bool:
    return isinstance(module, (nn.Linear, nn.Embedding, nn.Conv2d, nn.BatchNorm2d))


def get_module_name(group_name: str, model: nn.Module, key_word: str, exist_module_name: set, mpu=None, verbose: bool = True):
    """Names of compressible modules matching ``key_word`` (a regex) that no earlier group claimed."""
    found: List[str] = []
    for name, module in model.named_modules():
        if is_module_compressible(module, mpu) and re.search(key_word, name) is not None:
            if name in exist_module_name and verbose:
                raise ValueError(f"{name} is already added to compression, please check your config file for {group_name}.")
            if name not in exist_module_name:
                exist_module_name.add(name)
                found.append(name)
    return found, exist_module_name


def get_compress_methods(model: nn.Module, compress_methods: dict, mpu=None):
    """``[(module names, related names, {technique: params})]`` from the ``compression_training`` section."""
    out, seen = [], set()
    for method, content in compress_methods.items():
        if method == "layer_reduction" or not isinstance(content, dict):
            continue
        shared = content.get("shared_parameters", {})
        if not shared.get("enabled", False):
            continue
        for group_name, group in (content.get("different_groups") or {}).items():
            names: List[str] = []
            for kw in group.get("modules", []):
                got, seen = get_module_name(group_name, model, kw, seen, mpu, verbose=False)
                names += got
            related = [[n for n, _ in model.named_modules() if re.search(kw, n)] for kw in group.get("related_modules") or []]
            if names:
                out.append([names, related, {method: {**shared, **(group.get("params") or {})}}])
    return out


def convert_conv1d_to_linear(model: nn.Module, convert_type) -> nn.Module:
    """GPT-2 style ``Conv1D`` (weight ``[in, out]``) -> ``nn.Linear`` so that the compressible Linear applies."""
    for name, module in list(model.named_modules()):
        if isinstance(module, convert_type):
            w = module.weight.data
            lin = nn.Linear(w.shape[0], w.shape[1], bias=getattr(module, "bias", None) is not None)
            lin.weight.data = w.t().contiguous()
            if lin.bias is not None:
                lin.bias.data = module.bias.data.clone()
            recursive_setattr(model, name, lin)
    return model


def generate_pruners(config: dict, model: nn.Module):
    raise ImportError("snip_momentum pruning needs the `neural_compressor` package, which is not installed")


def register_on_step_begin(model: nn.Module):
    def hook(module, input):
        for p in getattr(module, "pruners", []):
            p.on_step_begin(0)
    return model.register_forward_pre_hook(hook)
