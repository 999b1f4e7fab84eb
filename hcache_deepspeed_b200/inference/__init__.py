"""``deepspeed.inference``: the v1 engine / config and the ragged (v2) engine's entry points (resolved on first use)."""
import importlib

_EXPORTS = {
    "InferenceEngine": ".engine", "DeepSpeedInferenceConfig": ".config",
    "RaggedInferenceEngineConfig": ".v2", "DeepSpeedTPConfig": ".v2.config_v2", "InferenceEngineV2": ".v2.engine_v2",
    "build_hf_engine": ".v2", "build_engine_from_ds_checkpoint": ".v2.engine_factory",
}


def __getattr__(name):
    if name in _EXPORTS:
        value = getattr(importlib.import_module(_EXPORTS[name], __name__), name)
        globals()[name] = value
        return value
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
