"""ZeRO-3 "leaf" modules: a flagged module gathers the parameters of its whole subtree on entry instead of hooking
every child (needed when the order in which children run changes between steps, e.g. MoE experts).
Reference: ``deepspeed/utils/z3_leaf_module.py:10-93``; honoured by ``runtime/zero/stage3.py::_install``."""
from __future__ import annotations

from typing import List, Type, Union

import torch


def z3_leaf_module(model: torch.nn.Module) -> bool:
    return bool(getattr(model, "_z3_leaf", False))


def z3_leaf_parameter(param: torch.nn.Parameter) -> bool:
    return hasattr(param, "ds_z3_leaf_module")


def get_z3_leaf_modules(model: torch.nn.Module) -> List[torch.nn.Module]:
    return [m for m in model.modules() if z3_leaf_module(m)]


def set_z3_leaf_module(model: torch.nn.Module, flag: bool) -> None:
    model._z3_leaf = flag


def _flag(model: torch.nn.Module, classes: Union[List[Type], List[str]], flag: bool) -> List[torch.nn.Module]:
    assert all(isinstance(c, (type, str)) for c in classes), f"leaf_module_classes must be types or class names, got {classes}"
    types_ = tuple(c for c in classes if isinstance(c, type))
    names = {c for c in classes if isinstance(c, str)}
    hits = [m for m in model.modules() if type(m) in types_ or type(m).__name__ in names]
    if not hits:
        raise ValueError(f"No modules of type {classes} found in model {model}")
    for m in hits:
        m._z3_leaf = flag
    return hits


def set_z3_leaf_modules(model: torch.nn.Module, leaf_module_classes: Union[List[Type], List[str]]) -> List[torch.nn.Module]:
    return _flag(model, leaf_module_classes, True)


def unset_z3_leaf_modules(model: torch.nn.Module, leaf_module_classes: Union[List[Type], List[str]]) -> List[torch.nn.Module]:
    return _flag(model, leaf_module_classes, False)
