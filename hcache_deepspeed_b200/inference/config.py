"""Configuration of ``init_inference`` (the v1 API).

Same field names, aliases and deprecated spellings as ``deepspeed/inference/config.py:118-323`` so a config
dict / JSON written for the reference loads unchanged. Fields that select reference-only back ends
(``use_triton``, ``triton_autotune``, ``transposed_mode``) are accepted and ignored with a warning:
this framework has exactly one device back end (sm_100a). ``hcache`` is new.
"""
from __future__ import annotations

from enum import Enum
from typing import Dict, Optional, Union

import torch
from pydantic import Field, field_validator

from ..config import DeepSpeedConfigModel, DtypeEnum
from ..utils.logging import logger


class MoETypeEnum(str, Enum):
    residual = "residual"
    standard = "standard"


class DeepSpeedTPConfig(DeepSpeedConfigModel):
    """Tensor-parallel settings."""
    enabled: bool = True
    tp_size: int = 1
    tp_grain_size: int = 64
    mpu: object = None
    tp_group: object = None


class DeepSpeedMoEConfig(DeepSpeedConfigModel):
    enabled: bool = True
    ep_size: int = 1
    moe_experts: list = Field([1], alias="num_experts")
    type: MoETypeEnum = MoETypeEnum.standard
    ep_mp_group: object = None
    ep_group: object = Field(None, alias="expert_group")


class QuantTypeEnum(str, Enum):
    asym = "asymmetric"
    sym = "symmetric"


class BaseQuantConfig(DeepSpeedConfigModel):
    enabled: bool = True
    num_bits: int = 8
    q_type: QuantTypeEnum = QuantTypeEnum.sym
    q_groups: int = 1


class WeightQuantConfig(BaseQuantConfig):
    enabled: bool = True
    quantized_initialization: Dict = {}
    post_init_quant: Dict = {}


class ActivationQuantConfig(BaseQuantConfig):
    enabled: bool = True


class QKVQuantConfig(DeepSpeedConfigModel):
    enabled: bool = True


class QuantizationConfig(DeepSpeedConfigModel):
    enabled: bool = True
    activation: ActivationQuantConfig = ActivationQuantConfig()
    weight: WeightQuantConfig = WeightQuantConfig()
    qkv: QKVQuantConfig = QKVQuantConfig()


class InferenceCheckpointConfig(DeepSpeedConfigModel):
    checkpoint_dir: Optional[str] = None
    save_mp_checkpoint_path: Optional[str] = None
    base_dir: Optional[str] = None


class DeepSpeedZeroConfigLite(DeepSpeedConfigModel):
    """ZeRO-Inference hook (``inference/config.py:159``): parameters may be kept on the host and streamed in."""
    stage: int = 0
    offload_param: Optional[Dict] = None


class V1HCacheConfig(DeepSpeedConfigModel):
    """Hidden-state capture for the v1 engine (new: the reference's v1 path has no HCache hooks)."""
    enabled: bool = False
    max_sessions: int = 64
    max_context: int = 4096
    max_batch_tokens: int = 4096
    kv_memory_bytes: int = 1 << 30


class DeepSpeedInferenceConfig(DeepSpeedConfigModel):
    replace_with_kernel_inject: bool = Field(False, alias="kernel_inject")
    """Run the transformer stack on this framework's sm_100a executor instead of the module's own forward."""
    dtype: torch.dtype = torch.float16
    tensor_parallel: DeepSpeedTPConfig = Field(default_factory=DeepSpeedTPConfig, alias="tp")
    enable_cuda_graph: bool = False
    use_triton: bool = False
    triton_autotune: bool = False
    zero: DeepSpeedZeroConfigLite = Field(default_factory=DeepSpeedZeroConfigLite)
    triangular_masking: bool = Field(True, alias="tm")
    moe: Union[bool, DeepSpeedMoEConfig] = Field(default_factory=DeepSpeedMoEConfig)
    keep_module_on_host: bool = False
    quant: QuantizationConfig = Field(default_factory=QuantizationConfig)
    checkpoint: Optional[Union[str, Dict]] = None
    base_dir: str = ""
    set_empty_params: bool = False
    save_mp_checkpoint_path: Optional[str] = None
    checkpoint_config: InferenceCheckpointConfig = Field(default_factory=InferenceCheckpointConfig, alias="ckpt_config")
    return_tuple: bool = True
    training_mp_size: int = 1
    replace_method: str = Field("auto", json_schema_extra={"deprecated": True,
                                                            "deprecated_msg": "This parameter is no longer needed"})
    injection_policy: Optional[Dict] = Field(None, alias="injection_dict")
    injection_policy_tuple: Optional[tuple] = None
    config: Optional[Dict] = Field(None, alias="args")
    max_out_tokens: int = Field(1024, alias="max_tokens")
    min_out_tokens: int = Field(1, alias="min_tokens")
    transposed_mode: bool = Field(False, alias="transposed_mode")
    hcache: V1HCacheConfig = Field(default_factory=V1HCacheConfig)
    # deprecated spellings, forwarded to their new home
    mp_size: int = Field(1, json_schema_extra={"deprecated": True, "new_param": "tensor_parallel.tp_size"})
    mpu: object = Field(None, json_schema_extra={"deprecated": True, "new_param": "tensor_parallel.mpu"})
    ep_size: int = Field(1, json_schema_extra={"deprecated": True, "new_param": "moe.ep_size"})
    ep_group: object = Field(None, alias="expert_group", json_schema_extra={"deprecated": True, "new_param": "moe.ep_group"})
    ep_mp_group: object = Field(None, alias="expert_mp_group",
                                json_schema_extra={"deprecated": True, "new_param": "moe.ep_mp_group"})
    moe_experts: list = Field([1], json_schema_extra={"deprecated": True, "new_param": "moe.moe_experts"})
    moe_type: MoETypeEnum = Field(MoETypeEnum.standard, json_schema_extra={"deprecated": True, "new_param": "moe.type"})

    @field_validator("dtype", mode="before")
    def validate_dtype(cls, field_value):
        if isinstance(field_value, str):
            return DtypeEnum.from_any(field_value).value
        if isinstance(field_value, DtypeEnum):
            return field_value.value
        if isinstance(field_value, torch.dtype):
            return field_value
        raise TypeError(f"Invalid type for dtype: {type(field_value)}")

    @field_validator("moe")
    def moe_backward_compat(cls, field_value):
        if isinstance(field_value, bool):
            return DeepSpeedMoEConfig(enabled=field_value)
        return field_value

    @field_validator("use_triton")
    def has_triton(cls, field_value):
        if field_value:
            logger.warning("use_triton is ignored: this framework has a single hand-written sm_100a back end")
        return field_value
