"""Validated readers of the ``compression_training`` block (``deepspeed/compression/config.py``).

Every technique has ``shared_parameters`` (switch, schedule, method, ...) and ``different_groups`` (per group: the
parameters of the technique, the modules it applies to, and related modules whose shape follows). The reference spells the
defaults and checks out once per technique (22 functions); here one table holds them and the per-technique functions are
thin wrappers, so the rules can be read in one place:

* shared defaults and the allowed values of enumerated entries (``_SHARED``),
* which group parameters must be given (``_REQUIRED_GROUP_PARAMS``) and which get a default (``_GROUP_DEFAULTS``).
"""
from __future__ import annotations

import copy
from typing import Dict

WEIGHT_QUANTIZATION, ACTIVATION_QUANTIZATION = "weight_quantization", "activation_quantization"
SPARSE_PRUNING, ROW_PRUNING, HEAD_PRUNING, CHANNEL_PRUNING = "sparse_pruning", "row_pruning", "head_pruning", "channel_pruning"
LAYER_REDUCTION, COMPRESSION_TRAINING = "layer_reduction", "compression_training"
SHARED_PARAMETERS, DIFFERENT_GROUPS = "shared_parameters", "different_groups"

# technique -> {entry: default}; "_choices": {entry: allowed values}
_SHARED: Dict[str, Dict] = {
    WEIGHT_QUANTIZATION: {"enabled": False, "quantizer_kernel": False, "schedule_offset": 0, "quantize_groups": 1,
                          "quantize_verbose": False, "quantization_type": "symmetric", "quantize_weight_in_forward": False,
                          "rounding": "nearest",
                          "_choices": {"quantization_type": ("symmetric", "asymmetric"), "rounding": ("nearest", "stochastic")}},
    ACTIVATION_QUANTIZATION: {"enabled": False, "quantization_type": "symmetric", "range_calibration": "dynamic",
                              "schedule_offset": 1000,
                              "_choices": {"quantization_type": ("symmetric", "asymmetric"),
                                           "range_calibration": ("dynamic", "static")}},
    SPARSE_PRUNING: {"enabled": False, "method": "l1", "schedule_offset": 1000,
                     "_choices": {"method": ("l1", "topk", "snip_momentum")}},
    ROW_PRUNING: {"enabled": False, "method": "l1", "schedule_offset": 1000, "_choices": {"method": ("l1", "topk")}},
    HEAD_PRUNING: {"enabled": False, "method": "topk", "schedule_offset": 1000, "_choices": {"method": ("l1", "topk")}},
    CHANNEL_PRUNING: {"enabled": False, "method": "l1", "schedule_offset": 1000, "_choices": {"method": ("l1", "topk")}},
}
_REQUIRED_GROUP_PARAMS = {WEIGHT_QUANTIZATION: ("start_bits", "target_bits"), ACTIVATION_QUANTIZATION: ("bits", ),
                          SPARSE_PRUNING: ("dense_ratio", ), ROW_PRUNING: ("dense_ratio", ), HEAD_PRUNING: ("dense_ratio", ),
                          CHANNEL_PRUNING: ("dense_ratio", )}
_GROUP_DEFAULTS = {WEIGHT_QUANTIZATION: {"quantization_period": 1}}


def _shared(technique: str, section: Dict) -> Dict:
    table = _SHARED[technique]
    given = section.get(SHARED_PARAMETERS) or {}
    out = {k: given.get(k, d) for k, d in table.items() if k != "_choices"}
    for key, allowed in table["_choices"].items():
        assert out[key] in allowed, f"Invalid {technique} {key} '{out[key]}'. Supported: {list(allowed)}"
    if technique == WEIGHT_QUANTIZATION:
        mixed = given.get("fp16_mixed_quantize") or {}
        out["fp16_mixed_quantize"] = mixed.get("enabled", False)
        out["quantize_change_ratio"] = mixed.get("quantize_change_ratio", 0.001)
    if technique == SPARSE_PRUNING and out["method"] == "snip_momentum":
        out["block_pattern"] = given.get("block_pattern", "4x1")
        out["dense_ratio"] = given.get("dense_ratio", 0.1)
        assert 0 < out["dense_ratio"] < 1, "Invalid dense_ratio value. Must be less than 1"
        out["schedule_offset_stride"] = given.get("schedule_offset_stride", 1)
        out["excluded_modules"] = list(given.get("excluded_modules", []))
        out["schedule_offset_end"] = given.get("schedule_offset_end", out["schedule_offset"])
        assert out["schedule_offset"] <= out["schedule_offset_end"], "Invalid schedule_offset and schedule_offset_end values"
    if technique == HEAD_PRUNING and given:
        assert "num_heads" in given, "num_heads must be specified for head pruning"
        out["num_heads"] = given["num_heads"]
    return out


def _groups(technique: str, section: Dict) -> Dict:
    out = {}
    for name, group in (section.get(DIFFERENT_GROUPS) or {}).items():
        params = dict(group.get("params", {}))
        for req in _REQUIRED_GROUP_PARAMS[technique]:
            assert req in params, f"{req} must be specified for {technique} group {name}"
        for k, d in _GROUP_DEFAULTS.get(technique, {}).items():
            params.setdefault(k, d)
        out[name] = {"params": params, "modules": group.get("modules", ["*"]),
                     "related_modules": group.get("related_modules", None)}
    return out


def _technique(technique: str, param_dict: Dict) -> Dict:
    section = param_dict.get(technique) or {}
    shared = _shared(technique, section)
    if shared["enabled"]:
        assert DIFFERENT_GROUPS in section, f"{technique} is enabled, {DIFFERENT_GROUPS} must be specified"
    return {SHARED_PARAMETERS: shared, DIFFERENT_GROUPS: _groups(technique, section)}


def _wrappers(technique: str):

    def whole(param_dict: Dict) -> Dict:
        return _technique(technique, param_dict)

    def shared(param_dict: Dict) -> Dict:
        return _shared(technique, param_dict)

    def groups(param_dict: Dict) -> Dict:
        return _groups(technique, param_dict)

    for fn, suffix in ((whole, ""), (shared, "_shared_parameters"), (groups, "_different_groups")):
        fn.__name__ = f"get_{technique}{suffix}"
        fn.__doc__ = f"The ``{technique}`` block{' ' + suffix[1:].replace('_', ' ') if suffix else ''}, defaults filled in and checked."
    return whole, shared, groups


get_weight_quantization, get_weight_quantization_shared_parameters, get_weight_quantization_different_groups = _wrappers(WEIGHT_QUANTIZATION)
get_activation_quantization, get_activation_quantization_shared_parameters, get_activation_quantization_different_groups = \
    _wrappers(ACTIVATION_QUANTIZATION)
get_sparse_pruning, get_sparse_pruning_shared_parameters, get_sparse_pruning_different_groups = _wrappers(SPARSE_PRUNING)
get_row_pruning, get_row_pruning_shared_parameters, get_row_pruning_different_groups = _wrappers(ROW_PRUNING)
get_head_pruning, get_head_pruning_shared_parameters, get_head_pruning_different_groups = _wrappers(HEAD_PRUNING)
get_channel_pruning, get_channel_pruning_shared_parameters, get_channel_pruning_different_groups = _wrappers(CHANNEL_PRUNING)


def get_layer_reduction_enabled(param_dict: Dict) -> bool:
    return bool((param_dict.get(LAYER_REDUCTION) or {}).get("enabled", False))


def get_layer_reduction_params(param_dict: Dict):
    if LAYER_REDUCTION not in param_dict:
        return False
    params = copy.copy(param_dict[LAYER_REDUCTION])
    params.pop("enabled", None)
    return params


def get_layer_reduction(param_dict: Dict) -> Dict:
    out = {"enabled": False}
    if get_layer_reduction_enabled(param_dict):
        out["enabled"] = True
        out.update(get_layer_reduction_params(param_dict))
    return out


def get_quantize_enabled(param_dict: Dict) -> bool:
    """Is weight quantisation (MoQ) switched on in a whole config dict?"""
    section = (param_dict.get(COMPRESSION_TRAINING) or {}).get(WEIGHT_QUANTIZATION) or {}
    return bool((section.get(SHARED_PARAMETERS) or {}).get("enabled", False))


def get_compression_config(param_dict: Dict) -> Dict:
    """The whole ``compression_training`` block of a config dict, every technique with defaults and checks applied."""
    sub = param_dict.get(COMPRESSION_TRAINING) or {}
    out = {t: _technique(t, sub) for t in _SHARED}
    out[LAYER_REDUCTION] = get_layer_reduction(sub)
    return out
