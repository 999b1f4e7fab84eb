"""Declarative checkpoint parameters: a :class:`ParameterBase` subclass names its source tensors as annotations,
the checkpoint loader assigns them in any order, and the moment the last one lands ``finalize()`` runs (fuse / shard /
transform through the model) and the result replaces the parameter object on its :class:`LayerContainer`.

    class UnfusedQKVParameter(ParameterBase):
        q_params: torch.Tensor
        k_params: torch.Tensor
        v_params: torch.Tensor
        def finalize(self):
            return self.inference_model.transform_qkv_param(torch.cat([self.q_params, self.k_params, self.v_params]))

Reference: ``deepspeed/inference/v2/model_implementations/parameter_base.py:54-255`` (metaclass-generated properties
with name-mangled storage). Here dependencies are gathered in ``__init_subclass__`` and stored in one dict per instance;
assignment is intercepted by ``__setattr__``.
"""
from __future__ import annotations

import weakref
from typing import Dict, List, Optional, get_type_hints

import torch


class ParametrizedList:
    """A dependency that is a LIST of tensors whose length comes from the model (e.g. one per expert): subclass via
    :func:`ParamList` naming the model attribute that holds the count. Counts as one dependency of its parameter."""
    count_attr: str = ""

    def __init__(self, param: "ParameterBase", n_params: int) -> None:
        self.n_params, self.set_params = n_params, 0
        self.param = weakref.ref(param)
        self._params: List[Optional[torch.Tensor]] = [None] * n_params

    def __getitem__(self, index):
        return self._params[index]

    def __setitem__(self, index, value) -> None:
        if self._params[index] is not None:
            raise ValueError("Cannot set a parameter twice.")
        self._params[index] = value
        self.set_params += 1
        if self.set_params == self.n_params:
            self.param().complete_component()

    def __iter__(self):
        return iter(self._params)

    def __len__(self) -> int:
        return self.n_params


def ParamList(attr: str):
    """``experts: ParamList("n_experts")`` — a list dependency sized by ``inference_model.n_experts``."""
    return type("ParametrizedListInstance", (ParametrizedList, ), {"count_attr": attr})


def _dependencies(cls) -> Dict[str, type]:
    deps: Dict[str, type] = {}
    for klass in reversed(cls.__mro__):
        for name, ann in (klass.__dict__.get("__annotations__") or {}).items():
            if isinstance(ann, str):
                try:
                    ann = get_type_hints(klass).get(name, ann)
                except Exception:
                    continue
            if isinstance(ann, type) and issubclass(ann, (torch.Tensor, ParametrizedList)):
                deps[name] = ann
    return deps


class ParameterBase:
    dependencies: Dict[str, type] = {}
    n_dependencies: int = 0

    def __init_subclass__(cls, **kw) -> None:
        super().__init_subclass__(**kw)
        cls.dependencies = _dependencies(cls)
        cls.n_dependencies = len(cls.dependencies)
        for name, ann in cls.dependencies.items():
            if issubclass(ann, ParametrizedList) and not ann.count_attr:
                raise TypeError(f"{cls.__name__}.{name}: use ParamList('<model attribute>') to size a list dependency")

    def __init__(self, model, parent_container) -> None:
        d = self.__dict__
        d["inference_model"] = model
        d["completed_components"] = 0
        d["parent_container"] = weakref.ref(parent_container)
        d["dest_param"] = None
        d["_values"] = {}
        for name, ann in self.dependencies.items():
            if issubclass(ann, ParametrizedList):
                if not hasattr(model, ann.count_attr):
                    raise AttributeError(f"{type(self).__name__}.{name} needs model attribute '{ann.count_attr}'")
                d["_values"][name] = ann(self, getattr(model, ann.count_attr))
            else:
                d["_values"][name] = None

    def __getattr__(self, name):
        values = self.__dict__.get("_values")
        if values is not None and name in values:
            return values[name]
        raise AttributeError(f"{type(self).__name__} has no attribute {name!r}")

    def __setattr__(self, name, value) -> None:
        ann = self.dependencies.get(name)
        if ann is None:
            return object.__setattr__(self, name, value)
        if issubclass(ann, ParametrizedList):
            raise ValueError("Cannot set a ParametrizedList directly.")
        if self._values[name] is not None:
            raise ValueError(f"Cannot set {type(self).__name__}.{name} twice.")
        self._values[name] = value
        self.complete_component()

    def finalize(self) -> torch.Tensor:
        """All dependencies present: return the transformed tensor (runs once, automatically)."""
        raise NotImplementedError

    def complete_component(self) -> None:
        self.completed_components += 1
        if self.completed_components == self.n_dependencies:
            self.parent_container().finalization_callback(self, self.finalize())
