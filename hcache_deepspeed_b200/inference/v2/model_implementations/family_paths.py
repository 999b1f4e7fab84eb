"""The reference keeps one package per model family (``deepspeed/inference/v2/model_implementations/<family>/`` with
``model.py``, ``policy.py``, ``container.py``). Here every family runs on ONE executor parameterised by its ``ArchSpec``,
so those packages would be empty shells; instead this single table is the compatibility map, and :func:`install` makes
the reference's import paths resolve to the real objects:

    from <pkg>.inference.v2.model_implementations.llama_v2.policy import Llama2Policy
    from <pkg>.inference.v2.model_implementations.qwen_v2.model import Qwen2InferenceModel
    from <pkg>.inference.v2.model_implementations.opt.container import MAPPER

``<family>.model.<X>InferenceModel`` is the executor class tagged with the family's ``model_type``;
``<family>.policy.<X>Policy`` is the shared policy class; ``<family>.container`` exposes the checkpoint-name mapper
(``MAPPER``) that plays the role of the reference's parameter containers."""
from __future__ import annotations

import sys
import types

#: reference package name -> (HF model_type, model class name, policy class name)
FAMILIES = {
    "falcon": ("falcon", "FalconInferenceModel", "FalconPolicy"),
    "llama_v2": ("llama", "Llama2InferenceModel", "Llama2Policy"),
    "mistral": ("mistral", "MistralInferenceModel", "MistralPolicy"),
    "mixtral": ("mixtral", "MixtralInferenceModel", "MixtralPolicy"),
    "opt": ("opt", "OPTInferenceModel", "OPTPolicy"),
    "phi": ("phi", "PhiInferenceModel", "PhiPolicy"),
    "phi3": ("phi3", "Phi3InferenceModel", "Phi3Policy"),
    "qwen": ("qwen", "QwenInferenceModel", "QwenPolicy"),
    "qwen_v2": ("qwen2", "Qwen2InferenceModel", "Qwen2Policy"),
    "qwen_v2_moe": ("qwen2_moe", "Qwen2MoeInferenceModel", "Qwen2MoePolicy"),
}


def install(parent) -> None:
    """Register ``<parent>.<family>`` (+ ``.model`` / ``.policy`` / ``.container``) modules for every family."""
    from ....models.families import get_mapper
    from ....models.transformer import RaggedTransformer
    for pkg_name, (model_type, model_cls_name, policy_cls_name) in FAMILIES.items():
        base = f"{parent.__name__}.{pkg_name}"
        if base in sys.modules:
            continue
        policy_cls = getattr(parent, policy_cls_name)
        model_cls = type(model_cls_name, (RaggedTransformer, ), {"model_type": model_type, "__module__": base + ".model"})
        pkg = types.ModuleType(base, f"import path of the reference's `{pkg_name}` family package (see family_paths.py)")
        pkg.__path__ = []   # a package: its three sub-modules are registered below
        subs = {
            "model": {model_cls_name: model_cls, "MODEL_TYPE": model_type},
            "policy": {policy_cls_name: policy_cls, "MODEL_TYPE": model_type},
            "container": {"MODEL_TYPE": model_type, "MAPPER": get_mapper(model_type)},
        }
        for sub, attrs in subs.items():
            mod = types.ModuleType(f"{base}.{sub}")
            mod.__dict__.update(attrs)
            sys.modules[mod.__name__] = mod
            setattr(pkg, sub, mod)
        setattr(pkg, model_cls_name, model_cls)
        setattr(pkg, policy_cls_name, policy_cls)
        sys.modules[base] = pkg
        setattr(parent, pkg_name, pkg)


class _FamilyLoader:
    """Hands the already-built alias module to the import system."""

    @staticmethod
    def create_module(spec):
        return sys.modules[spec.name]

    @staticmethod
    def exec_module(module):
        return None


class _FamilyFinder:
    """Resolves ``<parent>.<family>[.model|.policy|.container]`` on first import (the alias modules are built lazily:
    building them needs the executor, which itself imports this package)."""

    def __init__(self, parent_name: str):
        self.parent_name = parent_name

    def find_spec(self, fullname, path=None, target=None):
        if not fullname.startswith(self.parent_name + "."):
            return None
        rest = fullname[len(self.parent_name) + 1:].split(".")
        if rest[0] not in FAMILIES or len(rest) > 2 or (len(rest) == 2 and rest[1] not in ("model", "policy", "container")):
            return None
        install(sys.modules[self.parent_name])
        import importlib.util
        return importlib.util.spec_from_loader(fullname, _FamilyLoader(), is_package=len(rest) == 1)


def register(parent_name: str) -> None:
    if not any(isinstance(f, _FamilyFinder) and f.parent_name == parent_name for f in sys.meta_path):
        sys.meta_path.insert(0, _FamilyFinder(parent_name))
