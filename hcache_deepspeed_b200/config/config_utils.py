"""pydantic base for every config block.

Same contract as the reference's ``DeepSpeedConfigModel`` (``deepspeed/runtime/config_utils.py:17-117``):
unknown keys are rejected, fields can be given by name or alias, the string ``"auto"`` is accepted
where the caller fills the value later, and a field tagged
``json_schema_extra={"deprecated": True, "new_param": "a.b"}`` forwards its value to the new
location with a warning.
"""
from __future__ import annotations

import logging
from enum import Enum
from functools import reduce
from typing import Any, Dict

import torch
from pydantic import BaseModel, ConfigDict, field_serializer

logger = logging.getLogger("hcache_deepspeed_b200")


class DeepSpeedConfigModel(BaseModel):
    model_config = ConfigDict(
        validate_default=True,
        validate_assignment=True,
        use_enum_values=True,
        populate_by_name=True,
        extra="forbid",
        arbitrary_types_allowed=True,
        protected_namespaces=(),
    )

    def __init__(self, strict: bool = False, **data):
        if not strict:  # "auto" placeholders are dropped so defaults apply
            data = {k: v for k, v in data.items() if (v != "auto" or k == "replace_method")}
        super().__init__(**data)
        self._deprecated_fields_check()

    def _process_deprecated_field(self, dep_field: str) -> None:
        fields_set = self.model_fields_set
        kwargs = type(self).model_fields[dep_field].json_schema_extra or {}
        new_param = kwargs.get("new_param", "")
        dep_msg = kwargs.get("deprecated_msg", "")
        new_param_fn = kwargs.get("new_param_fn", lambda x: x)
        if dep_field not in fields_set:
            return
        logger.warning("Config parameter %s is deprecated%s%s", dep_field,
                       f", use {new_param} instead" if new_param else "", f". {dep_msg}" if dep_msg else "")
        if new_param and kwargs.get("set_new_param", True):
            new_name = new_param.split(".")[0]
            if new_name in fields_set and "." not in new_param:
                raise ValueError(f"Cannot provide deprecated parameter '{dep_field}' and replacing parameter "
                                 f"'{new_param}' together")
            try:
                if "." in new_param:
                    obj = reduce(getattr, new_param.split(".")[:-1], self)
                    setattr(obj, new_param.split(".")[-1], new_param_fn(getattr(self, dep_field)))
                else:
                    setattr(self, new_param, new_param_fn(getattr(self, dep_field)))
            except Exception as e:  # pragma: no cover - configuration error path
                logger.error("Could not forward deprecated parameter %s to %s", dep_field, new_param)
                raise e

    def _deprecated_fields_check(self) -> None:
        for name, info in type(self).model_fields.items():
            if (info.json_schema_extra or {}).get("deprecated", False):
                self._process_deprecated_field(name)

    @field_serializer("dtype", check_fields=False)
    def _serialize_torch_dtype(self, dtype: Any) -> str:
        return str(dtype)


def get_config_default(config: type, field_name: str) -> Any:
    assert field_name in config.model_fields, f"'{field_name}' is not a field in {config}"
    assert not config.model_fields[field_name].is_required(), f"'{field_name}' is a required field"
    return config.model_fields[field_name].get_default()


class pp_int(int):
    """int that prints with thousands separators (for config dumps)."""

    def __new__(cls, val, custom_print_str=None):
        inst = super().__new__(cls, val)
        inst.custom_print_str = custom_print_str
        return inst

    def __repr__(self):
        return self.custom_print_str or f"{self.real:,}"


class DtypeEnum(Enum):
    """Accepts torch dtypes and their usual string spellings (ref: deepspeed/inference/config.py:16-42)."""
    fp16 = (torch.float16, "torch.float16", "fp16", "float16", "half")
    fp32 = (torch.float32, "torch.float32", "fp32", "float32", "float")
    bf16 = (torch.bfloat16, "torch.bfloat16", "bf16", "bfloat16", "bfloat")
    int8 = (torch.int8, "torch.int8", "int8")

    def __new__(cls, *values):
        obj = object.__new__(cls)
        obj._value_ = values[0]
        for other in values[1:]:
            cls._value2member_map_[other] = obj
        obj._all_values = values
        return obj

    def __repr__(self):
        return "<%s.%s: %s>" % (self.__class__.__name__, self._name_, ", ".join(repr(v) for v in self._all_values))

    @classmethod
    def from_any(cls, v) -> "DtypeEnum":
        if isinstance(v, cls):
            return v
        return cls(v)


def dict_raise_error_on_duplicate_keys(ordered_pairs) -> Dict:
    """json.load object_pairs_hook rejecting duplicate keys (ref: config_utils.py:205)."""
    d = dict(ordered_pairs)
    if len(d) != len(ordered_pairs):
        import collections
        counter = collections.Counter(k for k, _ in ordered_pairs)
        raise ValueError("Duplicate keys in config: {}".format([k for k, c in counter.items() if c > 1]))
    return d
