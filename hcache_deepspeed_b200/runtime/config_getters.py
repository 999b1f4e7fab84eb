"""Readers of single entries of a raw config dict, under the reference's names (``deepspeed/runtime/config.py:96-560``).

``DeepSpeedConfig`` parses the whole dict into typed sections; integrations that only hold the raw JSON (HF Trainer's
``HfDeepSpeedConfig``, Megatron-DeepSpeed argument checks, the autotuner) call these instead. Scalars come from one table
(name -> path, default); the entries that combine several keys are written out below it.
"""
from __future__ import annotations

import copy
import json
from enum import Enum
from typing import Any, Dict, Optional

import torch


def get_scalar_param(param_dict: Dict, name: str, default: Any):
    return param_dict.get(name, default)


class DtypeEnum(Enum):
    """A dtype under every spelling configs use; ``DtypeEnum("bf16").value`` is the torch dtype."""
    fp16 = torch.float16, "torch.float16", "fp16", "float16", "half"
    fp32 = torch.float32, "torch.float32", "fp32", "float32", "float"
    fp64 = torch.float64, "torch.float64", "fp64", "float64", "double"
    int8 = torch.int8, "torch.int8", "int8"
    bf16 = torch.bfloat16, "torch.bfloat16", "bf16", "bfloat16", "bfloat"

    def __new__(cls, *values):
        obj = object.__new__(cls)
        obj._value_ = values[0]
        for other in values[1:]:
            cls._value2member_map_[other] = obj
        obj._all_values = values
        return obj

    def __repr__(self):
        return "<%s.%s: %s>" % (self.__class__.__name__, self._name_, ", ".join(repr(v) for v in self._all_values))


class ValidationMode:
    WARN, IGNORE, FAIL = "WARN", "IGNORE", "FAIL"


# getter name -> (path into the raw dict, default)
_SCALARS = {
    "get_gradient_accumulation_steps": (("gradient_accumulation_steps", ), None),
    "get_sparse_gradients_enabled": (("sparse_gradients", ), False),
    "get_prescale_gradients": (("prescale_gradients", ), False),
    "get_gradient_predivide_factor": (("gradient_predivide_factor", ), 1.0),
    "get_steps_per_print": (("steps_per_print", ), None),
    "get_disable_allgather": (("disable_allgather", ), False),
    "get_dump_state": (("dump_state", ), False),
    "get_gradient_clipping": (("gradient_clipping", ), 0.0),
    "get_graph_harvesting": (("graph_harvesting", ), False),
    "get_zero_allow_untested_optimizer": (("zero_allow_untested_optimizer", ), False),
    "get_zero_force_ds_cpu_optimizer": (("zero_force_ds_cpu_optimizer", ), True),
    "get_train_batch_size": (("train_batch_size", ), None),
    "get_train_micro_batch_size_per_gpu": (("train_micro_batch_size_per_gpu", ), None),
    "get_wall_clock_breakdown": (("wall_clock_breakdown", ), False),
    "get_memory_breakdown": (("memory_breakdown", ), False),
    "get_dataloader_drop_last": (("dataloader_drop_last", ), False),
    "get_expert_data_topo_config": (("use_data_before_expert_parallelism", ), False),
    "get_pld_enabled": (("progressive_layer_drop", "enabled"), False),
    "get_amp_enabled": (("amp", "enabled"), False),
    "get_fp16_enabled": (("fp16", "enabled"), False),
    "get_eigenvalue_enabled": (("eigenvalue", "enabled"), False),
    "get_eigenvalue_verbose": (("eigenvalue", "verbose"), False),
    "get_eigenvalue_max_iter": (("eigenvalue", "max_iter"), 100),
    "get_eigenvalue_tol": (("eigenvalue", "tol"), 1e-2),
    "get_eigenvalue_stability": (("eigenvalue", "stability"), 1e-6),
    "get_eigenvalue_gas_boundary_resolution": (("eigenvalue", "gas_boundary_resolution"), 1),
    "get_eigenvalue_layer_name": (("eigenvalue", "layer_name"), "bert.encoder.layer"),
    "get_eigenvalue_layer_num": (("eigenvalue", "layer_num"), 0),
    "get_optimizer_name": (("optimizer", "type"), None),
    "get_optimizer_legacy_fusion": (("optimizer", "legacy_fusion"), False),
    "get_scheduler_name": (("scheduler", "type"), None),
}


def _dig(d: Dict, path, default):
    cur = d
    for hop in path:
        if not isinstance(cur, dict) or hop not in cur:
            return default
        cur = cur[hop]
    return cur


def _make(name, path, default):

    def getter(param_dict: Dict):
        return _dig(param_dict, path, default)

    getter.__name__ = name
    getter.__doc__ = f"``{'.'.join(path)}`` of a raw config dict (default ``{default!r}``)."
    return getter


for _n, (_p, _d) in _SCALARS.items():
    globals()[_n] = _make(_n, _p, _d)
del _n, _p, _d


# ---- entries that combine keys -------------------------------------------------------------------------------------------------
def _section_without_enabled(param_dict: Dict, key: str):
    if key not in param_dict:
        return False
    sec = copy.copy(param_dict[key])
    sec.pop("enabled", None)
    return sec


def get_pld_params(param_dict: Dict):
    return _section_without_enabled(param_dict, "progressive_layer_drop")


def get_amp_params(param_dict: Dict):
    return _section_without_enabled(param_dict, "amp")


def _bf16_section(param_dict: Dict) -> Optional[Dict]:
    for key in ("bf16", "bfloat16"):
        if key in param_dict:
            return param_dict[key]
    return None


def get_bfloat16_enabled(param_dict: Dict) -> bool:
    sec = _bf16_section(param_dict)
    return bool(sec.get("enabled", False)) if sec is not None else False


def get_bfloat16_immediate_grad_update(param_dict: Dict) -> bool:
    sec = _bf16_section(param_dict)
    return bool(sec.get("immediate_grad_update", False)) if sec is not None else False


def get_fp16_master_weights_and_grads_enabled(param_dict: Dict) -> bool:
    return bool(param_dict["fp16"].get("fp16_master_weights_and_grads", False)) if get_fp16_enabled(param_dict) else False   # noqa: F821


def get_fp16_auto_cast(param_dict: Dict):
    return param_dict["fp16"].get("auto_cast", False) if get_fp16_enabled(param_dict) else None   # noqa: F821


def get_loss_scale(param_dict: Dict) -> float:
    """Static loss scale: the fp16 block's value (0 = dynamic), 1 under bf16 or fp32."""
    if get_fp16_enabled(param_dict):   # noqa: F821
        return param_dict["fp16"].get("loss_scale", 0)
    return 1.0 if get_bfloat16_enabled(param_dict) else 0


def get_initial_dynamic_scale(param_dict: Dict) -> float:
    if get_fp16_enabled(param_dict):   # noqa: F821
        power = param_dict["fp16"].get("initial_scale_power", 16)
    elif get_bfloat16_enabled(param_dict):
        power = 0
    else:
        power = 16
    return 2 ** power


def get_dynamic_loss_scale_args(param_dict: Dict):
    """The dynamic loss scaler's arguments, only when the fp16 block sets at least one of them explicitly."""
    if not get_fp16_enabled(param_dict):   # noqa: F821
        return None
    fp16 = param_dict["fp16"]
    keys = ("initial_scale_power", "loss_scale_window", "min_loss_scale", "hysteresis", "consecutive_hysteresis")
    if not any(k in fp16 for k in keys):
        return None
    return {"init_scale": 2 ** fp16.get("initial_scale_power", 16), "scale_window": fp16.get("loss_scale_window", 1000),
            "delayed_shift": fp16.get("hysteresis", 2), "consecutive_hysteresis": fp16.get("consecutive_hysteresis", False),
            "min_scale": fp16.get("min_loss_scale", 1)}


def get_communication_data_type(param_dict: Dict, comm_type: str = "communication_data_type", comm_data_type_default=None):
    val = param_dict.get(comm_type, comm_data_type_default)
    val = val.lower() if isinstance(val, str) else val
    table = {None: None, "fp32": torch.float32, "fp16": torch.float16, "bf16": torch.bfloat16, "bfp16": torch.bfloat16}
    if val not in table:
        raise ValueError(f"Invalid communication_data_type. Supported data types: ['fp16', 'bf16', 'fp32']. Got: {val}")
    return table[val]


def get_optimizer_params(param_dict: Dict):
    if get_optimizer_name(param_dict) is not None and "params" in param_dict["optimizer"]:   # noqa: F821
        return param_dict["optimizer"]["params"]
    return None


def get_optimizer_gradient_clipping(param_dict: Dict):
    params = get_optimizer_params(param_dict)
    return params["max_grad_norm"] if params is not None and "max_grad_norm" in params else None


def get_scheduler_params(param_dict: Dict):
    if get_scheduler_name(param_dict) is not None and "params" in param_dict["scheduler"]:   # noqa: F821
        return param_dict["scheduler"]["params"]
    return None


def get_pipeline_config(param_dict: Dict) -> Dict:
    cfg = {"stages": "auto", "partition": "best", "seed_layers": False, "activation_checkpoint_interval": 0,
           "pipe_partitioned": True, "grad_partitioned": True}
    cfg.update(param_dict.get("pipeline", {}))
    return cfg


def get_eigenvalue_config(param_dict: Dict):
    """(enabled, verbose, max_iter, tol, stability, gas_boundary_resolution, layer_name, layer_num); the block is only
    read together with quantize-training (MoQ), as in the reference."""
    if (param_dict.get("compression_training") or {}).get("weight_quantization", {}).get("shared_parameters", {}).get("enabled", False) \
            or "quantize_training" in param_dict:
        src = param_dict.get("quantize_training", param_dict)
        return (get_eigenvalue_enabled(src), get_eigenvalue_verbose(src), get_eigenvalue_max_iter(src), get_eigenvalue_tol(src),   # noqa: F821
                get_eigenvalue_stability(src), get_eigenvalue_gas_boundary_resolution(src), get_eigenvalue_layer_name(src),   # noqa: F821
                get_eigenvalue_layer_num(src))   # noqa: F821
    return (False, False, 100, 1e-2, 1e-6, 1, "bert.encoder.layer", 0)


def get_checkpoint_params(param_dict: Dict) -> Dict:
    return param_dict.get("checkpoint", {})


def get_data_types_params(param_dict: Dict) -> Dict:
    return param_dict.get("data_types", {})


def get_checkpoint_tag_validation_mode(checkpoint_params: Dict) -> str:
    mode = str(checkpoint_params.get("tag_validation", ValidationMode.WARN)).upper()
    if mode not in (ValidationMode.WARN, ValidationMode.IGNORE, ValidationMode.FAIL):
        raise ValueError(f"Checkpoint config contains invalid tag_validation value of {mode}, expecting one of "
                         f"{[ValidationMode.WARN, ValidationMode.IGNORE, ValidationMode.FAIL]}")
    return mode


def get_checkpoint_parallel_write_pipeline(checkpoint_params: Dict) -> bool:
    val = (checkpoint_params.get("parallel_write") or {}).get("pipeline_stage", False)
    if val not in (True, False):
        raise ValueError(f"checkpoint::parallel_write::pipeline_stage value of '{val}' is invalid, expecting: true or false")
    return val


# ---- sparse attention layouts --------------------------------------------------------------------------------------------------
def get_sparse_attention_mode(param_dict: Dict) -> str:
    return param_dict.get("mode", "fixed")


def get_sparse_attention_type(param_dict: Dict) -> str:
    return param_dict.get("attention", "bidirectional")


def get_sparse_dense_config(sparsity: Dict) -> Dict:
    return {"mode": "dense", "block": sparsity.get("block", 16)}


def _sparse_common(sparsity: Dict, mode: str) -> Dict:
    return {"mode": mode, "block": sparsity.get("block", 16),
            "different_layout_per_head": sparsity.get("different_layout_per_head", False)}


def get_sparse_fixed_config(sparsity: Dict) -> Dict:
    out = _sparse_common(sparsity, "fixed")
    out.update(num_local_blocks=sparsity.get("num_local_blocks", 4), num_global_blocks=sparsity.get("num_global_blocks", 1),
               attention=sparsity.get("attention", "bidirectional"),
               horizontal_global_attention=sparsity.get("horizontal_global_attention", False),
               num_different_global_patterns=sparsity.get("num_different_global_patterns", 1))
    return out


def get_sparse_variable_config(sparsity: Dict) -> Dict:
    out = _sparse_common(sparsity, "variable")
    out.update(num_random_blocks=sparsity.get("num_random_blocks", 0), local_window_blocks=sparsity.get("local_window_blocks", [4]),
               global_block_indices=sparsity.get("global_block_indices", [0]),
               global_block_end_indices=sparsity.get("global_block_end_indices", None),
               attention=sparsity.get("attention", "bidirectional"),
               horizontal_global_attention=sparsity.get("horizontal_global_attention", False))
    return out


def get_sparse_bigbird_config(sparsity: Dict) -> Dict:
    out = _sparse_common(sparsity, "bigbird")
    out.update(num_random_blocks=sparsity.get("num_random_blocks", 1),
               num_sliding_window_blocks=sparsity.get("num_sliding_window_blocks", 3),
               num_global_blocks=sparsity.get("num_global_blocks", 1))
    return out


def get_sparse_bslongformer_config(sparsity: Dict) -> Dict:
    out = _sparse_common(sparsity, "bslongformer")
    out.update(num_sliding_window_blocks=sparsity.get("num_sliding_window_blocks", 3),
               global_block_indices=sparsity.get("global_block_indices", [0]),
               global_block_end_indices=sparsity.get("global_block_end_indices", None))
    return out


def get_sparse_attention(param_dict: Dict):
    """The ``sparse_attention`` block with the defaults of its mode filled in (``None`` without the block)."""
    if "sparse_attention" not in param_dict:
        return None
    sparsity = param_dict["sparse_attention"]
    mode = get_sparse_attention_mode(sparsity)
    readers = {"dense": get_sparse_dense_config, "fixed": get_sparse_fixed_config, "variable": get_sparse_variable_config,
               "bigbird": get_sparse_bigbird_config, "bslongformer": get_sparse_bslongformer_config}
    if mode not in readers:
        raise NotImplementedError(f"Given sparsity mode, {mode}, has not been implemented yet!")
    return readers[mode](sparsity)


class DeepSpeedConfigWriter:
    """Assemble a config dict key by key and write it as JSON."""

    def __init__(self, data=None):
        self.data = data if data is not None else {}

    def add_config(self, key, value):
        self.data[key] = value

    def load_config(self, filename):
        with open(filename) as f:
            self.data = json.load(f)

    def write_config(self, filename):
        with open(filename, "w") as f:
            json.dump(self.data, f)


class HybridEngineConfig:
    """The ``hybrid_engine`` block (RLHF train + generate on one model) with its defaults."""
    _defaults = dict(enabled=False, max_out_tokens=512, inference_tp_size=1, release_inference_cache=False, pin_parameters=True,
                     tp_gather_partition_size=8)

    def __init__(self, **kwargs):
        unknown = set(kwargs) - set(self._defaults)
        if unknown:
            raise ValueError(f"unknown hybrid_engine keys: {sorted(unknown)}")
        for k, v in {**self._defaults, **kwargs}.items():
            setattr(self, k, v)

    def dict(self):
        return {k: getattr(self, k) for k in self._defaults}

    model_dump = dict


def get_hybrid_engine_config(param_dict: Dict) -> HybridEngineConfig:
    return HybridEngineConfig(**param_dict.get("hybrid_engine", {}))
