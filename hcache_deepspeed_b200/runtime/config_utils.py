"""Config helpers under their ``deepspeed/runtime/config_utils.py`` names: the pydantic base lives in
:mod:`hcache_deepspeed_b200.config.config_utils`; the dict accessors and the JSON encoder are here."""
from __future__ import annotations

import collections
import json
from typing import Any, Dict

from ..config.config_utils import (DeepSpeedConfigModel, DtypeEnum, dict_raise_error_on_duplicate_keys,  # noqa: F401
                                   get_config_default, pp_int)


class ScientificNotationEncoder(json.JSONEncoder):
    """Large numbers (bucket sizes, thresholds) print as ``5e+08`` instead of ``500000000`` in config dumps.
    Reference ``config_utils.py:146-173``."""

    def iterencode(self, o, _one_shot: bool = False, level: int = 0):
        pad = " " * (self.indent or 0) * level
        inner = " " * (self.indent or 0) * (level + 1)
        if isinstance(o, bool):
            return "true" if o else "false"
        if isinstance(o, (int, float)):
            return f"{o:e}" if abs(o) > 1e3 else f"{o}"
        if isinstance(o, collections.abc.Mapping):
            items = [f'\n{inner}"{k}": {self.iterencode(v, level=level + 1)}' for k, v in o.items()]
            return "{" + ", ".join(items) + f"\n{pad}" + "}"
        if isinstance(o, collections.abc.Sequence) and not isinstance(o, str):
            return "[" + ", ".join(self.iterencode(v, level=level + 1) for v in o) + "]"
        return "\n, ".join(super().iterencode(o, _one_shot))


class DeepSpeedConfigObject:
    """Plain config holder whose ``repr`` is its JSON."""

    def repr(self) -> Dict[str, Any]:
        return self.__dict__

    def __repr__(self) -> str:
        return json.dumps(self.__dict__, sort_keys=True, indent=4, cls=ScientificNotationEncoder)


def get_scalar_param(param_dict: dict, param_name: str, param_default_value):
    return param_dict.get(param_name, param_default_value)


def get_list_param(param_dict: dict, param_name: str, param_default_value):
    return param_dict.get(param_name, param_default_value)


def get_dict_param(param_dict: dict, param_name: str, param_default_value):
    return param_dict.get(param_name, param_default_value)
