"""``InferenceParameter``: a tensor that carries named auxiliary tensors (quantisation scales, ...) with it.
Reference: ``deepspeed/inference/v2/inference_parameter.py:24-89``."""
from __future__ import annotations

from typing import Dict

import torch

CORE_PARAM = "_ds_core_param_key"
STR_TO_DTYPE = {"torch.float32": torch.float32, "torch.float64": torch.float64, "torch.float16": torch.float16,
                "torch.bfloat16": torch.bfloat16, "torch.int64": torch.int64, "torch.int32": torch.int32,
                "torch.int16": torch.int16, "torch.int8": torch.int8, "torch.uint8": torch.uint8, "torch.bool": torch.bool}


class InferenceParameter(torch.Tensor):
    """Create with :meth:`initialize` / :meth:`initialize_raw`; aux tensors become attributes and follow ``.to()``."""

    @staticmethod
    def __new__(cls, tensor, *args, **kwargs):
        new = super().__new__(cls, tensor, *args, **kwargs)
        if hasattr(tensor, "_aux_attrs"):
            object.__setattr__(new, "_aux_attrs", tensor.aux_attrs)
        return new

    @property
    def aux_attrs(self) -> Dict[str, torch.Tensor]:
        return getattr(self, "_aux_attrs", {})

    def to(self, *args, **kwargs):
        new = InferenceParameter(super().to(*args, **kwargs))
        aux = {}
        for name, attr in self.aux_attrs.items():
            moved = attr.to(*args, **kwargs)
            aux[name] = moved
            setattr(new, name, moved)
        object.__setattr__(new, "_aux_attrs", aux)
        return new

    @classmethod
    def initialize(cls, core_param: torch.Tensor, **kwargs) -> "InferenceParameter":
        param = InferenceParameter(core_param)
        object.__setattr__(param, "_aux_attrs", dict(kwargs))
        for name, attr in kwargs.items():
            if hasattr(param, name):
                raise ValueError(f"Attribute {name} already exists on param.")
            if not isinstance(attr, torch.Tensor):
                raise ValueError(f"Attribute {name} must be a tensor.")
            setattr(param, name, attr)
        return param

    @classmethod
    def initialize_raw(cls, **kwargs) -> "InferenceParameter":
        """From a flat dict that stores the core tensor under :data:`CORE_PARAM` (deserialisation path)."""
        if CORE_PARAM not in kwargs:
            raise ValueError(f"Must provide core parameter, with key {CORE_PARAM}.")
        kwargs = dict(kwargs)
        return cls.initialize(kwargs.pop(CORE_PARAM), **kwargs)
