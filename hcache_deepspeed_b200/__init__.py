"""hcache_deepspeed_b200 — a B200-native (sm_100a) LLM inference / serving framework with the capabilities of
the HCache fork of DeepSpeed. Top-level API mirrors ``deepspeed/__init__.py`` (``init_inference`` :291,
``init_distributed`` :43, ``add_config_arguments`` :268, ``initialize`` :69). See DESIGN.md."""
from __future__ import annotations

import json
from typing import Optional, Union

from . import _import_paths

_import_paths.register()   # reference module paths that have no separate implementation here (one table, no shim files)

from . import comm  # noqa: F401,E402
from .comm.comm import init_distributed  # noqa: F401,E402
from .utils.logging import log_dist, logger  # noqa: F401
from .version import __version__  # noqa: F401

__git_hash__ = None
__git_branch__ = None


def _parse_version(version_str: str):
    import re
    m = re.match(r"(\d+)\.(\d+)\.(\d+)", version_str)
    return tuple(int(x) for x in m.groups()) if m else (0, 0, 0)


__version_major__, __version_minor__, __version_patch__ = _parse_version(__version__)


def init_inference(model, config=None, **kwargs):
    """Initialize the inference engine (v1 API).

    ``config`` may be a dict, a path to a JSON file, or None; keyword arguments are merged into it and a key given
    both ways with different values is an error (same rule as ``deepspeed/__init__.py:341-362``).
    Returns an :class:`hcache_deepspeed_b200.inference.engine.InferenceEngine`.
    """
    from .inference.config import DeepSpeedInferenceConfig
    from .inference.engine import InferenceEngine
    log_dist(f"hcache_deepspeed_b200 info: version={__version__}", ranks=[0])
    if config is None:
        config = {}
    if isinstance(config, str):
        with open(config, "r") as f:
            config = json.load(f)
    elif isinstance(config, DeepSpeedInferenceConfig):
        config = config.model_dump(exclude_unset=True)
    elif not isinstance(config, dict):
        raise ValueError(f"'config' argument expected string or dictionary, got {type(config)}")
    overlap = set(config.keys()).intersection(kwargs.keys())
    for key in overlap:
        if config[key] != kwargs[key]:
            raise ValueError(f"Conflicting argument '{key}' in 'config':{config[key]} and kwargs:{kwargs[key]}")
    config = dict(config)
    config.update(kwargs)
    return InferenceEngine(model, config=DeepSpeedInferenceConfig(**config))


def add_config_arguments(parser):
    """Add ``--deepspeed`` / ``--deepspeed_config`` (and the deprecated ``--deepscale*``) arguments."""
    group = parser.add_argument_group("DeepSpeed", "DeepSpeed configurations")
    group.add_argument("--deepspeed", default=False, action="store_true",
                       help="Enable DeepSpeed (helper flag for user code, no impact on DeepSpeed backend)")
    group.add_argument("--deepspeed_config", default=None, type=str, help="DeepSpeed json configuration file.")
    group.add_argument("--deepscale", default=False, action="store_true",
                       help="Deprecated enable DeepSpeed (helper flag for user code, no impact on DeepSpeed backend)")
    group.add_argument("--deepscale_config", default=None, type=str, help="Deprecated DeepSpeed json configuration file.")
    return parser


def default_inference_config() -> dict:
    from .inference.config import DeepSpeedInferenceConfig
    return DeepSpeedInferenceConfig().model_dump()


def initialize(*args, **kwargs):
    """Training entry point (``deepspeed.initialize``): see :mod:`hcache_deepspeed_b200.runtime`."""
    from .runtime.engine import initialize as _initialize
    return _initialize(*args, **kwargs)


# ---- the rest of the ``deepspeed`` top-level namespace, resolved on first use (keeps ``import`` light) --------------
_LAZY_MODULES = {
    "zero": ".runtime.zero", "moe": ".moe", "pipe": ".pipe", "checkpoint": ".checkpoint", "module_inject": ".module_inject",
    "ops": ".ops", "utils": ".utils", "runtime": ".runtime", "inference": ".inference", "sequence": ".sequence",
    "compression": ".compression", "autotuning": ".autotuning", "elasticity": ".elasticity", "launcher": ".launcher",
    "monitor": ".monitor", "profiling": ".profiling", "accelerator": ".accelerator", "linear": ".linear", "nvme": ".nvme",
    "model_implementations": ".model_implementations", "checkpointing": ".runtime.activation_checkpointing.checkpointing",
    "domino": ".runtime.domino", "comm": ".comm", "compile": ".compile",
}
_LAZY_ATTRS = {
    "DeepSpeedEngine": (".runtime.engine", "DeepSpeedEngine"), "DeepSpeedConfig": (".runtime.config", "DeepSpeedConfig"),
    "DeepSpeedHybridEngine": (".runtime.hybrid_engine", "DeepSpeedHybridEngine"),
    "PipelineEngine": (".runtime.pipe.engine", "PipelineEngine"), "PipelineModule": (".runtime.pipe.module", "PipelineModule"),
    "InferenceEngine": (".inference.engine", "InferenceEngine"),
    "DeepSpeedInferenceConfig": (".inference.config", "DeepSpeedInferenceConfig"),
    "DeepSpeedTransformerInference": (".model_implementations", "DeepSpeedTransformerInference"),
    "DeepSpeedTransformerLayer": (".ops.transformer.transformer", "DeepSpeedTransformerLayer"),
    "DeepSpeedTransformerConfig": (".ops.transformer.transformer", "DeepSpeedTransformerConfig"),
    "OnDevice": (".utils.init_on_device", "OnDevice"), "tp_model_init": (".runtime.tensor_parallel.tp_manager", "tp_model_init"),
    "get_accelerator": (".accelerator", "get_accelerator"),
    "DeepSpeedConfigError": (".runtime.config", "DeepSpeedConfigError"), "ADAM_OPTIMIZER": (".runtime.config", "ADAM_OPTIMIZER"),
    "LAMB_OPTIMIZER": (".runtime.config", "LAMB_OPTIMIZER"), "add_tuning_arguments": (".runtime.lr_schedules", "add_tuning_arguments"),
    "replace_transformer_layer": (".module_inject.replace_module", "replace_transformer_layer"),
    "revert_transformer_layer": (".module_inject.replace_module", "revert_transformer_layer"),
    "set_autotp_mode": (".runtime.tensor_parallel", "set_autotp_mode"),
    "git_hash": (".git_version_info", "git_hash"), "git_branch": (".git_version_info", "git_branch"),
    "TORCH_DISTRIBUTED_DEFAULT_PORT": (".constants", "TORCH_DISTRIBUTED_DEFAULT_PORT"),
}


def __getattr__(name: str):
    import importlib
    if name in _LAZY_MODULES:
        mod = importlib.import_module(_LAZY_MODULES[name], __name__)
        globals()[name] = mod
        return mod
    if name in _LAZY_ATTRS:
        mod_name, attr = _LAZY_ATTRS[name]
        value = getattr(importlib.import_module(mod_name, __name__), attr)
        globals()[name] = value
        return value
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")


def __dir__():
    return sorted(set(globals()) | set(_LAZY_MODULES) | set(_LAZY_ATTRS))


def is_compile_supported() -> bool:
    """``torch.compile`` is never the hot path here (kernels are hand-written); DeepCompile is declined (``compile/``)."""
    return False
