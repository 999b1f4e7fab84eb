"""Configuration of the ragged (v2) engine.

Field names and aliases follow ``deepspeed/inference/v2/config_v2.py:13-44`` (``tp`` / ``manager`` aliases
included) so existing config dicts load unchanged; ``hcache`` is new.
"""
from __future__ import annotations

from typing import Optional

from pydantic import Field

from ...config import DeepSpeedConfigModel
from .ragged.manager_configs import DSStateManagerConfig


class DeepSpeedTPConfig(DeepSpeedConfigModel):
    """Tensor-parallel settings."""
    tp_size: int = 1
    """Number of devices the model is split across."""
    fused_collectives: bool = True
    """Reduce row-parallel GEMM results over NVLink from inside the GEMM epilogue (peer / multicast stores)
    instead of running NCCL all-reduce as a separate kernel. Falls back to NCCL when symmetric memory is
    unavailable."""


class QuantizationConfig(DeepSpeedConfigModel):
    quantization_mode: Optional[str] = None
    """Weight quantization mode. The reference's only mode ('wf6af16') is Ampere-only
    (``modules/heuristics.py:93-104``); accepted values here: None, 'wf8af8' (E4M3 weights per output channel and
    activations per token on the fp8 tensor cores — ``modules/implementations.py::Fp8TensorCoreLinear``) or 'wf6af16'
    (E3M2 weights packed 6 bits each, fed to the same tensor-core kernels — ``Fp6TensorCoreLinear``)."""


class HCacheConfig(DeepSpeedConfigModel):
    """Hidden-state cache behaviour (no counterpart in the reference: capture is unconditional there)."""
    enabled: bool = True
    """Build the model with a latent buffer so ``put`` can return per-layer hidden states."""
    capture: str = "sync"
    """'sync': ``put`` returns host tensors that are ready to read (reference semantics);
    'async': returns :class:`LatentHandle` objects, call ``.wait()`` before reading;
    'off': ``put`` returns ``None`` latents unless overridden per call."""
    capture_decode_steps: bool = True
    """Also capture single-token steps (needed if the generated tokens are to be restorable)."""
    overlap_restore: bool = True
    """``restore_kv`` defers its per-layer projections into the next ``put`` so the new turn's prefill overlaps the
    host-to-device transfer of the saved hidden states."""
    shard_capture: bool = False
    """Tensor parallel: every rank exports only its 1/tp row slice of each sequence's hidden states (they are
    replicated on the device), to be passed back to ``restore_kv(..., sharded=True)``."""


class RaggedInferenceEngineConfig(DeepSpeedConfigModel):
    """Sets parameters for the ragged inference engine."""
    tensor_parallel: DeepSpeedTPConfig = Field(default_factory=DeepSpeedTPConfig, alias="tp")
    state_manager: DSStateManagerConfig = Field(default_factory=DSStateManagerConfig, alias="manager")
    quantization: QuantizationConfig = Field(default_factory=QuantizationConfig)
    use_cuda_graph_decode: bool = True
    """Replay a captured CUDA graph for pure decode batches (one token per sequence, no capture): ~230 kernel
    launches per step become one graph launch. Graphs are keyed by (sequence count, attention split count)."""
    hcache: HCacheConfig = Field(default_factory=HCacheConfig)
