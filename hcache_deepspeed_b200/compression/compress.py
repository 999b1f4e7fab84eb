"""``init_compression`` swaps the modules named by the ``compression_training`` config for their compression-aware
versions and records each technique's schedule; ``redundancy_clean`` bakes masks / quantisation into the weights (and
shrinks dimensions when ``dense_ratio`` pruning allows); layer reduction keeps a subset of teacher layers.
Reference: ``deepspeed/compression/compress.py:100-240``, ``helper.py``, ``config.py``."""
from __future__ import annotations

import json
import os
import re
from typing import Dict, List, Optional

import torch
from torch import nn

from ..utils.logging import logger
from .basic_layer import Conv2dLayer_Compress, Embedding_Compress, LinearLayer_Compress
from .scheduler import TECHNIQUES

WEIGHT_QUANTIZATION, ACTIVATION_QUANTIZATION = "weight_quantization", "activation_quantization"
SPARSE_PRUNING, ROW_PRUNING, HEAD_PRUNING, CHANNEL_PRUNING, LAYER_REDUCTION = ("sparse_pruning", "row_pruning", "head_pruning",
                                                                              "channel_pruning", "layer_reduction")


def check_deepspeed_config(config):
    """A config dict as is, a path to a JSON file loaded (duplicate keys are an error)."""
    if isinstance(config, dict):
        return config
    if isinstance(config, str) and os.path.exists(config):
        def no_dupes(pairs):
            out = {}
            for k, v in pairs:
                if k in out:
                    raise ValueError(f"Duplicate key in DeepSpeed config: {k}")
                out[k] = v
            return out
        with open(config) as f:
            return json.load(f, object_pairs_hook=no_dupes)
    raise ValueError(f"Expected a string path to an existing deepspeed config, or a dictionary. Received: {config}")


def get_compression_config(ds_config) -> Dict:
    if isinstance(ds_config, str):
        ds_config = json.load(open(ds_config))
    cfg = dict(ds_config.get("compression_training", {}))
    for t in TECHNIQUES:
        sec = dict(cfg.get(t, {}))
        shared = dict(sec.get("shared_parameters", {}))
        shared.setdefault("enabled", False)
        shared.setdefault("schedule_offset", 0)
        sec["shared_parameters"], sec["different_groups"] = shared, dict(sec.get("different_groups", {}))
        cfg[t] = sec
    cfg.setdefault(LAYER_REDUCTION, {"enabled": False})
    return cfg


def _matches(name: str, patterns: List[str]) -> bool:
    return any(p == "*" or re.search(p, name) for p in patterns)


def _set(model: nn.Module, name: str, new: nn.Module):
    parts = name.split(".")
    obj = model
    for p in parts[:-1]:
        obj = getattr(obj, p)
    setattr(obj, parts[-1], new)


def _to_compress_layer(m: nn.Module) -> Optional[nn.Module]:
    if isinstance(m, (LinearLayer_Compress, Embedding_Compress, Conv2dLayer_Compress)):
        return None
    if isinstance(m, nn.Linear):
        new = LinearLayer_Compress(m.in_features, m.out_features, bias=m.bias is not None)
    elif isinstance(m, nn.Embedding):
        new = Embedding_Compress(m.num_embeddings, m.embedding_dim, m.padding_idx)
    elif isinstance(m, nn.Conv2d):
        new = Conv2dLayer_Compress(m.in_channels, m.out_channels, m.kernel_size, m.stride, m.padding, m.dilation, m.groups,
                                   m.bias is not None, m.padding_mode)
    else:
        return None
    new.load_state_dict(m.state_dict(), strict=False)
    return new.to(next(m.parameters()).device).to(next(m.parameters()).dtype)


def _apply_technique(layer, tech: str, shared: Dict, params: Dict):
    if tech == WEIGHT_QUANTIZATION:
        layer.enable_weight_quantization(params.get("start_bits", 8), params.get("target_bits", 8),
                                         params.get("quantization_period", 1), shared.get("quantize_weight_in_forward", True),
                                         shared.get("quantization_type", "symmetric"), shared.get("quantize_groups", 1))
    elif tech == ACTIVATION_QUANTIZATION:
        layer.enable_activation_quantization(params.get("bits", 8), shared.get("quantization_type", "symmetric"),
                                             shared.get("range_calibration", "dynamic"))
    elif tech == SPARSE_PRUNING:
        layer.enable_sparse_pruning(params.get("dense_ratio", 0.5), shared.get("method", "l1"))
    elif tech == ROW_PRUNING:
        layer.enable_row_pruning(params.get("dense_ratio", 0.5), shared.get("method", "l1"))
    elif tech == HEAD_PRUNING:
        layer.enable_head_pruning(params.get("dense_ratio", 0.5), shared.get("method", "topk"), shared["num_heads"])
    elif tech == CHANNEL_PRUNING:
        layer.enable_channel_pruning(params.get("dense_ratio", 0.5), shared.get("method", "l1"))


def student_initialization(student_model, teacher_model, deepspeed_config):
    """Layer reduction: copy the chosen teacher layers (and the named other modules) into the smaller student."""
    cfg = get_compression_config(deepspeed_config)[LAYER_REDUCTION]
    prefix, teacher_layers = cfg["module_name_prefix"], cfg["teacher_layer"]
    t_sd, s_sd = teacher_model.state_dict(), student_model.state_dict()
    for s_idx, t_idx in enumerate(teacher_layers):
        for k in list(s_sd):
            if k.startswith(f"{prefix}.{s_idx}."):
                s_sd[k] = t_sd[k.replace(f"{prefix}.{s_idx}.", f"{prefix}.{t_idx}.", 1)].clone()
    for name in cfg.get("other_module_name", []):
        for k in list(s_sd):
            if k.startswith(name + ".") or k == name:
                s_sd[k] = t_sd[k].clone()
    student_model.load_state_dict(s_sd)
    return student_model


def init_compression(model: nn.Module, deepspeed_config, teacher_model=None, mpu=None) -> nn.Module:
    cfg = get_compression_config(deepspeed_config)
    inner = model.module if hasattr(model, "module") and isinstance(model.module, nn.Module) else model
    if cfg[LAYER_REDUCTION].get("enabled", False):
        assert teacher_model is not None, "Teacher model is required for layer reduction"
        student_initialization(inner, teacher_model, deepspeed_config)
    plan = []
    for tech in TECHNIQUES:
        sec = cfg[tech]
        if not sec["shared_parameters"]["enabled"]:
            continue
        for gname, g in sec["different_groups"].items():
            for name, m in list(inner.named_modules()):
                if not _matches(name, g.get("modules", ["*"])):
                    continue
                if not isinstance(m, (nn.Linear, nn.Embedding, nn.Conv2d)):
                    continue
                if tech in (SPARSE_PRUNING, ROW_PRUNING, HEAD_PRUNING, ACTIVATION_QUANTIZATION) and not isinstance(m, (nn.Linear, nn.Conv2d)):
                    continue
                if tech in (ROW_PRUNING, HEAD_PRUNING, ACTIVATION_QUANTIZATION) and not isinstance(m, nn.Linear):
                    continue
                new = _to_compress_layer(m)
                if new is not None:
                    _set(inner, name, new)
                    m = new
                _apply_technique(m, tech, sec["shared_parameters"], g.get("params", {}))
                plan.append((name, tech, gname, g.get("related_modules")))
    inner._compression_plan = plan
    inner._compression_config = cfg
    logger.info(f"init_compression: {len(plan)} (module, technique) pairs configured")
    return model


def redundancy_clean(model: nn.Module, deepspeed_config, mpu=None) -> nn.Module:
    """Make compression permanent: quantise weights to their target bits, apply masks, and — for row / head / channel
    pruning with ``related_modules`` — shrink the following layers' inputs to match."""
    cfg = get_compression_config(deepspeed_config)
    inner = model.module if hasattr(model, "module") and isinstance(model.module, nn.Module) else model
    mods = dict(inner.named_modules())
    for name, tech, gname, related in getattr(inner, "_compression_plan", []):
        m = mods.get(name)
        if m is None:
            continue
        dim_red = cfg[tech]["shared_parameters"].get("dense_ratio_reduce_dim", True) if tech in (ROW_PRUNING, HEAD_PRUNING, CHANNEL_PRUNING) else False
        if tech == WEIGHT_QUANTIZATION:
            m.fix_weight_quantization()
        elif tech == SPARSE_PRUNING and m.sparse_pruning_method:
            m.fix_sparse_pruning_helper()
        elif tech == ROW_PRUNING and m.row_pruning_method:
            mask = m.fix_row_col_pruning_helper(dim_reduction=dim_red)
            for rel_group in (related or []):
                for rname in rel_group:
                    for n2, m2 in mods.items():
                        if re.search(rname, n2) and isinstance(m2, LinearLayer_Compress) and n2.rpartition(".")[0] == name.rpartition(".")[0]:
                            m2.fix_row_col_pruning_helper(mask, dim_reduction=dim_red)
        elif tech == HEAD_PRUNING and m.head_pruning_method:
            m.fix_head_pruning_helper(dim_reduction=dim_red)
        elif tech == CHANNEL_PRUNING and getattr(m, "channel_pruning_method", None):
            m.fix_channel_pruning_helper(dim_reduction=dim_red)
    return model


from .helper import get_compress_methods, get_module_name  # noqa: E402,F401  (the reference exports them from here too)
