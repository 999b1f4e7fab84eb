"""``deepspeed.utils`` namespace: logging, process groups, NVTX ranges, the ``safe_get_* / safe_set_*`` accessors for
ZeRO-partitioned tensors, ZeRO-3 leaf-module flags, ``OnDevice``, ``RepeatingLoader``. Names resolve on first use (the
sub-modules import the communication layer, which imports this package)."""
from __future__ import annotations

import importlib

_EXPORTS = {
    "logger": ".logging", "log_dist": ".logging",
    "OnDevice": ".init_on_device",
    "instrument_w_nvtx": ".nvtx",
    "get_numactl_cmd": ".numa",
    "tensor_fragment": ".mixed_precision_linkage", "fragment_address": ".mixed_precision_linkage",
    "get_hp_fragment_mapping": ".mixed_precision_linkage", "link_hp_params": ".mixed_precision_linkage",
    "lazy_init_hp_params_optimizer_state": ".mixed_precision_linkage",
    "set_z3_leaf_modules": ".z3_leaf_module", "unset_z3_leaf_modules": ".z3_leaf_module", "get_z3_leaf_modules": ".z3_leaf_module",
    "z3_leaf_module": ".z3_leaf_module", "z3_leaf_parameter": ".z3_leaf_module", "set_z3_leaf_module": ".z3_leaf_module",
    "RepeatingLoader": "..runtime.dataloader",
    "get_caller_func": "..comm.comms_logging",
}
for _n in ("safe_get_full_fp32_param", "safe_get_full_grad", "safe_get_full_optimizer_state", "safe_set_full_fp32_param",
           "safe_set_full_optimizer_state", "safe_set_full_grad", "safe_get_local_fp32_param", "safe_get_local_grad",
           "safe_get_local_optimizer_state", "safe_set_local_fp32_param", "safe_set_local_grad", "safe_set_local_optimizer_state"):
    _EXPORTS[_n] = ".tensor_fragment"
_SUBMODULES = ("groups", "logging", "timer", "nvtx", "numa", "memory", "tensor_fragment", "z3_leaf_module", "init_on_device",
               "mixed_precision_linkage", "cuda_graph", "debug", "exceptions", "bwc", "types", "torch", "config", "zero_to_fp32")


def __getattr__(name: str):
    if name in _EXPORTS:
        value = getattr(importlib.import_module(_EXPORTS[name], __name__), name)
    elif name in _SUBMODULES:
        value = importlib.import_module("." + name, __name__)
    else:
        # ``from deepspeed.utils import *`` also re-exports the process-group helpers of ``groups``
        groups = importlib.import_module(".groups", __name__)
        if not name.startswith("__") and hasattr(groups, name):
            value = getattr(groups, name)
        else:
            raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
    globals()[name] = value
    return value


def __dir__():
    return sorted(set(globals()) | set(_EXPORTS) | set(_SUBMODULES))
