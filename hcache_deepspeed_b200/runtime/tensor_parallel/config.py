"""``tensor_parallel`` training config (``deepspeed/runtime/tensor_parallel/config.py:13-90``)."""
from __future__ import annotations

from enum import Enum
from typing import Optional

import torch
from pydantic import Field

from ...config.config_utils import DeepSpeedConfigModel


class AUTOTP_MODE(Enum):
    TRAINING = "TRAINING"
    INFERENCE = "INFERENCE"


class TPConfig(DeepSpeedConfigModel):
    tp_size: int = 1
    tp_grain_size: int = 1
    mpu: Optional[object] = None
    tp_group: Optional[object] = None


class TPTrainingConfig(DeepSpeedConfigModel):
    dtype: torch.dtype = torch.float16
    autotp_size: int = 0
    tp_overlap_comm: bool = False
    tensor_parallel: TPConfig = Field({}, alias="tp")
    injection_policy_tuple: Optional[tuple] = None
    keep_module_on_host: bool = False
    replace_with_kernel_inject: bool = Field(False, alias="kernel_inject")


def get_tensor_parallel_config(ds_config: dict) -> TPTrainingConfig:
    sec = dict(ds_config.get("tensor_parallel") or {})
    if "tp_grain_size" in sec:
        sec.setdefault("tp", {})["tp_grain_size"] = sec.pop("tp_grain_size")
    return TPTrainingConfig(**sec)
