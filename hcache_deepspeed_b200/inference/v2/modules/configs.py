"""Module configs (``deepspeed/inference/v2/modules/configs/*.py``)."""
from __future__ import annotations

from enum import Enum
from typing import Optional

import torch
from pydantic import ConfigDict

from .ds_module import DSModuleConfig


class PositionalEmbeddingType(Enum):
    none = "none"
    rotate_half = "rotate_half"
    alibi = "alibi"


class MaskingType(Enum):
    none = "none"
    causal = "causal"
    local = "local"
    asymmetric = "asymmetric"


class NormTypeEnum(Enum):
    LayerNorm = "layer_norm"
    RMSNorm = "rms_norm"


class _TorchCfg(DSModuleConfig):
    model_config = ConfigDict(arbitrary_types_allowed=True, extra="forbid", populate_by_name=True)


class RotateHalfConfig(_TorchCfg):
    use_trained_freqs: bool = False
    theta_base: float = 10_000.0
    rotate_dim: Optional[int] = None


class DSSelfAttentionConfig(_TorchCfg):
    n_heads_q: int
    n_heads_kv: int
    head_size: int
    max_sequences: int = 64
    scale_factor: float = 1.0
    input_dtype: torch.dtype = torch.float16
    output_dtype: torch.dtype = torch.float16
    masking_type: MaskingType = MaskingType.causal
    positional_embedding_type: PositionalEmbeddingType = PositionalEmbeddingType.none
    positional_embedding_config: Optional[RotateHalfConfig] = None


class DSEmbeddingsConfig(_TorchCfg):
    residual_dtype: torch.dtype = torch.float16
    embedding_dim: int
    positional_embedding: bool = False
    positional_offset: int = 0
    use_token_type: bool = False
    output_normalization: Optional[NormTypeEnum] = None


class DSLinearConfig(_TorchCfg):
    in_channels: int
    out_channels: int
    activation: str = "identity"   # identity / gelu / relu / silu / geglu / reglu / siglu (the last three are gated)
    input_dtype: torch.dtype = torch.float16
    output_dtype: torch.dtype = torch.float16


class DSMoEConfig(_TorchCfg):
    model_dim: int
    intermediate_features: int
    n_experts: int
    top_k: int = 1
    input_dtype: torch.dtype = torch.float16
    output_dtype: torch.dtype = torch.float16
    activation: str = "identity"
    normalize_scores: bool = False


class DSNormConfig(_TorchCfg):
    type: NormTypeEnum
    channels: int
    residual_dtype: torch.dtype = torch.float16
    input_dtype: torch.dtype = torch.float16
    output_dtype: torch.dtype = torch.float16
    eps: float = 1e-5


class DSUnembedConfig(_TorchCfg):
    dtype: torch.dtype = torch.float16
    norm_type: Optional[NormTypeEnum] = None
    vocab_size: int
    max_sequences: int = 64
    model_dim: int = 0


# the reference has this as a package of small files; `__path__` lets its sub-module paths (served from the alias table in
# _import_paths.py) import through this module
__path__: list = []
