"""Shared enums / helpers of the v2 inference stack (``deepspeed/inference/v2/inference_utils.py:13-105``)."""
from __future__ import annotations

from enum import Enum, IntEnum

import torch

from ...config.config_utils import DtypeEnum  # noqa: F401
from .modules.configs import NormTypeEnum  # noqa: F401


class ActivationType(IntEnum):
    """Ids the kernels take (gated variants sit ``GATED_OFFSET`` above their base activation)."""
    GELU = 0
    RELU = 1
    SILU = 2
    GEGLU = 3
    ReGLU = 4
    SiGLU = 5
    IDENTITY = 6
    InvalidType = -1


_GATED = {ActivationType.GEGLU: ActivationType.GELU, ActivationType.ReGLU: ActivationType.RELU,
          ActivationType.SiGLU: ActivationType.SILU}
# ids of ``ops/reference.apply_act`` / the GEMM epilogues
_NATIVE_ACT = {ActivationType.GELU: 1, ActivationType.RELU: 2, ActivationType.SILU: 3, ActivationType.IDENTITY: 0}


def is_gated(act_fn: ActivationType) -> bool:
    if not isinstance(act_fn, ActivationType):
        act_fn = ActivationType(act_fn)
    return act_fn in _GATED


def native_act_id(act_fn: ActivationType) -> int:
    """Activation id of this framework's kernels for a (possibly gated) reference activation type."""
    act_fn = ActivationType(act_fn)
    return _NATIVE_ACT[_GATED.get(act_fn, act_fn)]


def elem_size(dtype: torch.dtype) -> int:
    try:
        return {torch.float16: 2, torch.bfloat16: 2, torch.float32: 4, torch.float64: 8, torch.int8: 1, torch.uint8: 1,
                torch.int16: 2, torch.int32: 4, torch.int64: 8, torch.bool: 1}[dtype]
    except KeyError:
        raise ValueError("Unknown dtype size for {}".format(dtype))


def ceil_div(a: int, b: int) -> int:
    return -(-a // b)
