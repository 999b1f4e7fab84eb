"""``LayerContainer``: the set of (declarative) parameters of one layer plus the routing table from checkpoint names
to their dependencies.

    class Llama2TransformerContainer(LayerContainer):
        qkv_w: UnfusedQKVParameter
        mlp_1_w: GatedMLPParameter
        PARAM_MAPPING = {"self_attn.q_proj.weight": "qkv_w.q_params", ..., "mlp.experts.*.w1.weight": "moe_1.gating_experts"}

``set_dependency("self_attn.q_proj.weight", tensor)`` feeds the tensor to ``qkv_w.q_params``; once a parameter has all
of its dependencies it finalizes and the attribute becomes the resulting tensor. Mapping tables are validated when the
class is defined (unknown targets, doubly-mapped or unmapped dependencies, list targets without exactly one ``*``).

Reference: ``deepspeed/inference/v2/model_implementations/layer_container_base.py:42-355``.
"""
from __future__ import annotations

import re
from typing import Dict, List, Tuple, Union, get_type_hints

import torch

from ..inference_parameter import InferenceParameter
from .parameter_base import ParameterBase, ParametrizedList

MAPPING_KEY = "PARAM_MAPPING"


def _parameter_fields(cls) -> Dict[str, type]:
    out: Dict[str, type] = {}
    for klass in reversed(cls.__mro__):
        anns = klass.__dict__.get("__annotations__") or {}
        hints = None
        for name, ann in anns.items():
            if isinstance(ann, str):
                if hints is None:
                    try:
                        hints = get_type_hints(klass)
                    except Exception:
                        hints = {}
                ann = hints.get(name, ann)
            if isinstance(ann, type) and issubclass(ann, ParameterBase):
                out[name] = ann
    return out


class LayerContainer:
    PARAM_MAPPING: Dict[str, Union[str, List[str]]] = {}
    _fields: Dict[str, type] = {}
    _plist_rules: List[Tuple[str, str, List[str]]] = []

    def __init_subclass__(cls, **kw) -> None:
        super().__init_subclass__(**kw)
        cls._fields = _parameter_fields(cls)
        mapping: Dict[str, Union[str, List[str]]] = {}
        for klass in reversed(cls.__mro__):
            mapping.update(klass.__dict__.get(MAPPING_KEY) or {})
        cls.PARAM_MAPPING = mapping
        cls._plist_rules = []
        if not mapping:
            return
        unmapped = {f"{p}.{d}" for p, ptype in cls._fields.items() for d in ptype.dependencies}
        for src, targets in mapping.items():
            targets = [targets] if isinstance(targets, str) else list(targets)
            kinds = []
            for t in targets:
                pname, _, dep = t.partition(".")
                if pname not in cls._fields:
                    raise ValueError(f'Target parameter "{pname}" not found in this layer. Valid targets are {list(cls._fields)}')
                deps = cls._fields[pname].dependencies
                if dep not in deps:
                    raise ValueError(f'Target dependency "{dep}" not found on parameter "{pname}". Valid targets are {list(deps)}')
                if t not in unmapped:
                    raise ValueError(f'Target dependency "{t}" was targeted with multiple mapping rules.')
                unmapped.remove(t)
                kinds.append(issubclass(deps[dep], ParametrizedList))
            if all(kinds):
                if src.count("*") != 1:
                    raise ValueError(f"ParametrizedList mapping needs exactly one wildcard (*) in the source name: {src}")
                prefix, suffix = src.split("*")
                cls._plist_rules.append((prefix, suffix, targets))
            elif any(kinds):
                raise ValueError("Cannot mix ParametrizedLists and Tensors in a single mapping rule.")
        if unmapped:
            raise ValueError(f"A parameter mapping was provided for {cls.__name__}, but the following dependencies were "
                             f"not mapped: {unmapped}")

    def __init__(self, model) -> None:
        self.inference_model = model
        self._finalized_params = 0
        for name, ptype in self._fields.items():
            setattr(self, name, ptype(model, self))

    # ---- bookkeeping ---------------------------------------------------------------------------------------------
    def finalization_callback(self, param: ParameterBase, finalized: torch.Tensor) -> None:
        """Called by a parameter when it finalizes: the attribute that held it now holds the tensor."""
        for name in self._fields:
            if getattr(self, name) is param:
                setattr(self, name, finalized)
                self._finalized_params += 1
                return
        raise ValueError(f"{type(param).__name__} is not a member of {type(self).__name__}")

    @property
    def n_params(self) -> int:
        return len(self._fields)

    @property
    def annotation_attrs(self) -> List[str]:
        return list(self._fields)

    @property
    def mapping_params(self) -> dict:
        return self.PARAM_MAPPING

    @property
    def plist_helpers(self) -> list:
        return self._plist_rules

    def _check(self, check_device: bool) -> bool:
        if self._finalized_params != self.n_params:
            return False
        for name in self._fields:
            t = getattr(self, name)
            if t is None:
                continue
            if not isinstance(t, InferenceParameter):
                raise ValueError(f"Layer should be finalized, but {name} ({type(t)}) is neither InferenceParameter nor None")
            if check_device and torch.cuda.is_available() and t.device != torch.device("cuda", torch.cuda.current_device()):
                raise RuntimeError(f"Layer should be finalized, but {name} is not on device {torch.cuda.current_device()}")
        return True

    @property
    def is_populated(self) -> bool:
        return self._check(check_device=False)

    @property
    def is_initialized(self) -> bool:
        return self._check(check_device=True)

    def direct_injection(self, name: str, tensor) -> None:
        """Install an already-final tensor (flat-checkpoint restore path)."""
        if name not in self._fields:
            raise ValueError(f"Cannot directly inject {name}, not a valid parameter.")
        setattr(self, name, tensor)
        self._finalized_params += 1

    # ---- routing ---------------------------------------------------------------------------------------------------
    def set_dependency(self, dep_name: str, dep_value: torch.Tensor) -> None:
        """``dep_name`` is the checkpoint name with the layer prefix removed."""
        targets = self.PARAM_MAPPING.get(dep_name)
        if targets is None:
            for prefix, suffix, plist_targets in self._plist_rules:
                if dep_name.startswith(prefix) and dep_name.endswith(suffix) and len(dep_name) > len(prefix) + len(suffix):
                    middle = dep_name[len(prefix):len(dep_name) - len(suffix)]
                    if middle.isdigit():
                        for t in plist_targets:
                            pname, _, dep = t.partition(".")
                            getattr(getattr(self, pname), dep)[int(middle)] = dep_value
                        return
            matches = [t for k, t in self.PARAM_MAPPING.items() if "*" in k and re.fullmatch(re.escape(k).replace(r"\*", ".*"), dep_name)]
            if len(matches) > 1:
                raise ValueError(f"Multiple targets matched for dependency {dep_name}: {matches}")
            if not matches:
                raise ValueError(f"Could not find a mapping for dependency \"{dep_name}\". Check that it is included in "
                                 f"the ``{MAPPING_KEY}``.")
            targets = matches[0]
        for t in ([targets] if isinstance(targets, str) else targets):
            pname, _, dep = t.partition(".")
            setattr(getattr(self, pname), dep, dep_value)
