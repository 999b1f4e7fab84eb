"""Post-initialisation weight quantisation of a model according to ``weight_quantization.post_init_quant`` rules
(key = substring of the module name, value = {num_bits, group_size, group_dim, symmetric}).
Reference: ``deepspeed/inference/quantization/quantization.py:21-111``."""
from __future__ import annotations

from typing import Dict

import torch

from ...utils.logging import logger
from .layers import QUANTIZATION_LAYER_MAPPINGS
from .utils import recursive_setattr


def _init_group_wise_weight_quantization(model: torch.nn.Module, ds_config: Dict) -> torch.nn.Module:
    if "weight_quantization" not in ds_config or "post_init_quant" not in ds_config["weight_quantization"]:
        raise KeyError("Please provide quantization config")
    rules = ds_config["weight_quantization"]["post_init_quant"]
    matched_count = {k: 0 for k in rules}
    for name, module in list(model.named_modules()):
        if type(module) not in QUANTIZATION_LAYER_MAPPINGS:
            continue
        keys = [k for k in rules if k in name]
        if not keys:
            continue
        key = max(keys, key=len)  # the most specific rule wins
        cfg = dict(rules[key])
        cfg.setdefault("symmetric", False)
        cfg.setdefault("group_dim", 1)
        new = QUANTIZATION_LAYER_MAPPINGS[type(module)](cfg, module)
        recursive_setattr(model, name, new)
        matched_count[key] += 1
    for k, n in matched_count.items():
        logger.info(f"Quantization rule '{k}' matched {n} modules")
    return model
