"""Engine construction helpers (ref: ``deepspeed/inference/v2/engine_factory.py:32-135``)."""
from __future__ import annotations

import os
import pickle
from typing import Optional, Union

import torch

from ...models.arch import ArchSpec, named_arch
from ...models.flat_model import make_model_config_filename
from ...utils.logging import inference_logger
from .checkpoint import HuggingFaceCheckpointEngine, InMemoryModelEngine
from .config_v2 import RaggedInferenceEngineConfig
from .engine_v2 import InferenceEngineV2
from .model_implementations import MODEL_TYPE_TO_POLICY, InferenceV2Policy


def _as_config(engine_config) -> RaggedInferenceEngineConfig:
    if engine_config is None:
        return RaggedInferenceEngineConfig()
    if isinstance(engine_config, dict):
        return RaggedInferenceEngineConfig(**engine_config)
    return engine_config


def build_engine_from_ds_checkpoint(path: str, engine_config=None, debug_level: int = 20) -> InferenceEngineV2:
    """Engine from a checkpoint written by ``InferenceEngineV2.serialize`` (TP degree must match)."""
    inference_logger(level=debug_level)
    if not os.path.exists(make_model_config_filename(path)):
        raise FileNotFoundError(f"{path} holds no serialized inference checkpoint (ds_model_config.pkl missing)")
    with open(make_model_config_filename(path), "rb") as f:
        saved = pickle.load(f)
    policy = InferenceV2Policy(model_config=None, inf_checkpoint_path=path, arch=saved["arch"])
    return InferenceEngineV2(policy, _as_config(engine_config))


def build_hf_engine(path: str, engine_config=None, debug_level: int = 20,
                    dtype: Optional[torch.dtype] = None) -> InferenceEngineV2:
    """Engine from a HuggingFace checkpoint directory (or a directory written by ``serialize``)."""
    if os.path.exists(os.path.join(path, "ds_model_config.pkl")):
        return build_engine_from_ds_checkpoint(path, engine_config, debug_level=debug_level)
    inference_logger(level=debug_level)
    ckpt = HuggingFaceCheckpointEngine(path)
    model_config = ckpt.model_config
    mt = model_config.model_type
    if mt not in MODEL_TYPE_TO_POLICY:
        raise ValueError(f"Unsupported model type {mt}")
    policy = MODEL_TYPE_TO_POLICY[mt](model_config, checkpoint_engine=ckpt, dtype=dtype)
    return InferenceEngineV2(policy, _as_config(engine_config))


def build_engine_from_module(model: torch.nn.Module, engine_config=None, dtype: Optional[torch.dtype] = None,
                             device: Optional[torch.device] = None) -> InferenceEngineV2:
    """Engine from a live HuggingFace model object (weights are re-laid-out for the kernels)."""
    ckpt = InMemoryModelEngine(model)
    mt = model.config.model_type
    if mt not in MODEL_TYPE_TO_POLICY:
        raise ValueError(f"Unsupported model type {mt}")
    policy = MODEL_TYPE_TO_POLICY[mt](model.config, checkpoint_engine=ckpt, dtype=dtype, device=device)
    return InferenceEngineV2(policy, _as_config(engine_config))


def build_engine_from_arch(arch: Union[str, ArchSpec], engine_config=None, seed: int = 0,
                           dtype: torch.dtype = torch.bfloat16, device: Optional[torch.device] = None) -> InferenceEngineV2:
    """Engine with random-init weights of a named architecture (benchmarks: there are no checkpoints offline)."""
    spec = named_arch(arch) if isinstance(arch, str) else arch
    policy = InferenceV2Policy(model_config=None, arch=spec, random_seed=seed, dtype=dtype, device=device)
    return InferenceEngineV2(policy, _as_config(engine_config))
