"""Accessors of the ``data_efficiency`` config tree with defaults filled in
(``deepspeed/runtime/data_pipeline/config.py:14-195``)."""
from __future__ import annotations

import copy

from .constants import *  # noqa: F401,F403
from . import constants as C


def _section(parent: dict, key: str, defaults: dict) -> dict:
    out = copy.deepcopy(defaults)
    out.update(copy.deepcopy(parent.get(key) or {}))
    return out


def get_data_efficiency_enabled(param_dict: dict) -> bool:
    return bool((param_dict.get(C.DATA_EFFICIENCY) or {}).get(C.DATA_EFFICIENCY_ENABLED, C.DATA_EFFICIENCY_ENABLED_DEFAULT))


def get_data_efficiency_seed(param_dict: dict) -> int:
    return int((param_dict.get(C.DATA_EFFICIENCY) or {}).get(C.DATA_EFFICIENCY_SEED, C.DATA_EFFICIENCY_SEED_DEFAULT))


def get_curriculum_learning(sampling: dict) -> dict:
    out = _section(sampling, C.CURRICULUM_LEARNING, {C.CURRICULUM_LEARNING_ENABLED: C.CURRICULUM_LEARNING_ENABLED_DEFAULT})
    if out[C.CURRICULUM_LEARNING_ENABLED]:
        assert C.CURRICULUM_LEARNING_METRICS in out, f"Curriculum learning is enabled, {C.CURRICULUM_LEARNING_METRICS} must be specified"
    return out


def get_dynamic_batching(sampling: dict) -> dict:
    out = _section(sampling, C.DYNAMIC_BATCHING, {C.DYNAMIC_BATCHING_ENABLED: C.DYNAMIC_BATCHING_ENABLED_DEFAULT})
    if out[C.DYNAMIC_BATCHING_ENABLED]:
        assert C.DYNAMIC_BATCHING_MAX_TOKENS in out, f"Dynamic batching is enabled, {C.DYNAMIC_BATCHING_MAX_TOKENS} must be specified"
    return out


def get_data_sampling(eff: dict) -> dict:
    out = _section(eff, C.DATA_SAMPLING, {C.DATA_SAMPLING_ENABLED: C.DATA_SAMPLING_ENABLED_DEFAULT,
                                          C.DATA_SAMPLING_NUM_EPOCHS: C.DATA_SAMPLING_NUM_EPOCHS_DEFAULT,
                                          C.DATA_SAMPLING_NUM_WORKERS: C.DATA_SAMPLING_NUM_WORKERS_DEFAULT,
                                          C.DATA_SAMPLING_PIN_MEMORY: C.DATA_SAMPLING_PIN_MEMORY_DEFAULT})
    out[C.CURRICULUM_LEARNING] = get_curriculum_learning(out)
    out[C.DYNAMIC_BATCHING] = get_dynamic_batching(out)
    return out


def get_random_ltd(routing: dict) -> dict:
    out = _section(routing, C.RANDOM_LTD, {C.RANDOM_LTD_ENABLED: C.RANDOM_LTD_ENABLED_DEFAULT})
    out.setdefault(C.RANDOM_LTD_LAYER_TOKEN_LR_SCHEDULE, {}).setdefault("enabled", False)
    return out


def get_data_routing(eff: dict) -> dict:
    out = _section(eff, C.DATA_ROUTING, {C.DATA_ROUTING_ENABLED: C.DATA_ROUTING_ENABLED_DEFAULT})
    out[C.RANDOM_LTD] = get_random_ltd(out)
    return out


def get_data_efficiency_config(param_dict: dict) -> dict:
    eff = param_dict.get(C.DATA_EFFICIENCY) or {}
    return {C.DATA_EFFICIENCY_ENABLED: get_data_efficiency_enabled(param_dict), C.DATA_EFFICIENCY_SEED: get_data_efficiency_seed(param_dict),
            C.DATA_SAMPLING: get_data_sampling(eff), C.DATA_ROUTING: get_data_routing(eff)}


def get_data_sampling_enabled(param_dict: dict) -> bool:
    return bool(get_data_efficiency_config(param_dict)[C.DATA_SAMPLING][C.DATA_SAMPLING_ENABLED])


def get_data_routing_enabled(param_dict: dict) -> bool:
    return bool(get_data_efficiency_config(param_dict)[C.DATA_ROUTING][C.DATA_ROUTING_ENABLED])


def get_random_ltd_enabled(param_dict: dict) -> bool:
    return bool(get_data_efficiency_config(param_dict)[C.DATA_ROUTING][C.RANDOM_LTD][C.RANDOM_LTD_ENABLED])


def get_curriculum_enabled_legacy(param_dict: dict) -> bool:
    return bool((param_dict.get(C.CURRICULUM_LEARNING) or {}).get(C.CURRICULUM_LEARNING_ENABLED, False))


def get_curriculum_params_legacy(param_dict: dict):
    sec = copy.deepcopy(param_dict.get(C.CURRICULUM_LEARNING) or {})
    if not sec:
        return False
    sec.pop(C.CURRICULUM_LEARNING_ENABLED, None)
    return sec


# ---- single-entry readers (ref: data_pipeline/config.py:59-193) ---------------------------------------------------------------
def _sub(param_dict, section, key, default):
    return (param_dict.get(section) or {}).get(key, default)


def get_data_sampling_num_epochs(param_dict):
    return _sub(param_dict, "data_sampling", "num_epochs", 1000)


def get_data_sampling_num_workers(param_dict):
    return _sub(param_dict, "data_sampling", "num_workers", 0)


def get_data_sampling_pin_memory(param_dict):
    return _sub(param_dict, "data_sampling", "pin_memory", False)


def get_curriculum_learning_enabled(param_dict):
    return bool(_sub(param_dict, "curriculum_learning", "enabled", False))


def _params_of(param_dict, section):
    import copy
    if section not in param_dict:
        return {}
    out = copy.copy(param_dict[section])
    out.pop("enabled", None)
    return out


def get_curriculum_learning_params(param_dict):
    return _params_of(param_dict, "curriculum_learning")


def get_random_ltd_params(param_dict):
    return _params_of(param_dict, "random_ltd")
