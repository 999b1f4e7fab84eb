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


# ---- the reference's remaining helper names (compression/helper.py:185-322) -----------------------------------------------
def compression_preparation(model: nn.Module, compression_technique_list, mpu=None) -> nn.Module:
    """Swap every compressible layer for its compressible twin, then switch on the techniques of
    ``compression_technique_list`` - entries ``(module name lists, related modules, {technique: parameters})`` as
    ``get_compress_methods`` returns them."""
    from .compress import _apply_technique
    for name, module in list(model.named_modules()):
        if name and is_module_compressible(module, mpu):
            module_replacement(model, name, mpu=mpu)
    for module_name_lists, _, technique in compression_technique_list:
        for names in module_name_lists:
            for name in names:
                layer = recursive_getattr(model, name)
                for tech, params in technique.items():
                    shared = {k: v for k, v in params.items() if k != "params"}
                    _apply_technique(layer, tech, shared, params.get("params", params))
    return model


def fix_compression(model: nn.Module, module_name: str, compression_technique: dict, mask=None, dim_reduction: bool = False):
    """Make one module's compression permanent (quantised weights written back, pruned rows / heads / channels removed
    when ``dim_reduction``); returns the mask the next layer needs, if any."""
    module = recursive_getattr(model, module_name)
    for tech, v in compression_technique.items():
        on = bool(v.get("enabled", False))
        if tech == "weight_quantization" and on and v.get("quantize_weight_in_forward", False):
            return module.fix_weight_quantization()
        if tech == "sparse_pruning" and on:
            return module.fix_sparse_pruning_helper()
        if tech == "row_pruning" and (on or mask is not None):
            return module.fix_row_col_pruning_helper(mask, dim_reduction=dim_reduction)
        if tech == "head_pruning" and (on or mask is not None):
            return module.fix_head_pruning_helper(mask, v.get("num_heads"), dim_reduction=dim_reduction)
        if tech == "channel_pruning" and (on or mask is not None):
            return module.fix_channel_pruning_helper(mask, dim_reduction=dim_reduction)
    return None


def rewrite_optimizer_step(opt):
    """Bracket ``opt.step`` with the ``on_before_optimizer_step`` / ``on_after_optimizer_step`` calls of the pruners hung
    on ``opt.pruners`` (the original step stays reachable as ``opt.orig_step``)."""
    import types

    def new_step(self, closure=None):
        for pruner in getattr(self, "pruners", []):
            pruner.on_before_optimizer_step()
        res = self.orig_step(closure) if closure is not None else self.orig_step()
        for pruner in getattr(self, "pruners", []):
            pruner.on_after_optimizer_step()
        return res

    opt.orig_step = opt.step
    opt.step = types.MethodType(new_step, opt)
    return opt
