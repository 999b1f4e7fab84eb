"""ZeRO config names (the pydantic model lives in :mod:`..config`)."""
from enum import IntEnum

from ..config import OffloadDeviceEnum, OffloadOptimizerConfig, OffloadParamConfig, ZeroConfig  # noqa: F401

DeepSpeedZeroConfig = ZeroConfig
DeepSpeedZeroOffloadParamConfig = OffloadParamConfig
DeepSpeedZeroOffloadOptimizerConfig = OffloadOptimizerConfig


class ZeroStageEnum(IntEnum):
    disabled = 0
    optimizer_states = 1
    gradients = 2
    weights = 3
    max_stage = 3


def read_zero_config_deprecated(param_dict):
    return {"stage": 1 if param_dict.get("zero_optimization") else 0}


def get_zero_config(param_dict):
    """The typed ``zero_optimization`` block of a raw config dict (a bare ``true`` means stage 1, as of old)."""
    from ..config import ZeroConfig
    block = param_dict.get("zero_optimization", {})
    if isinstance(block, bool):
        block = {"stage": 1 if block else 0}
    return ZeroConfig(**block)
