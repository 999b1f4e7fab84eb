"""Injection policies: which model families can be lifted onto the sm_100a executor, and how their checkpoints map.

The reference keeps one container + policy class per family under ``module_inject/containers/`` (bert, bloom, clip,
distil_bert, gpt2, gptj, gptneo, gptneox, internlm, llama, llama2, megatron_gpt, megatron_gpt_moe, opt, unet, vae) that
extract q/k/v/o/mlp/norm tensors from one HF block. Here the per-family knowledge lives in
``models/families.py`` (name mappers) + ``models/arch.py`` (ArchSpec builders); a policy is just the pair. Families
without a mapper get per-layer injection of the v1 fused layer through ``module_inject/containers`` (gpt2, gptj, gptneo,
gptneox, bloom, opt, bert, distil_bert) or fall back to AutoTP.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, Dict, Optional


@dataclass
class InjectionPolicy:
    name: str
    model_types: tuple
    supported: bool
    note: str = ""


def _table() -> Dict[str, InjectionPolicy]:
    from ..inference.v2.model_implementations import MODEL_TYPE_TO_POLICY
    native = set(MODEL_TYPE_TO_POLICY)
    rows = [
        InjectionPolicy("llama", ("llama", ), True), InjectionPolicy("llama2", ("llama", ), True),
        InjectionPolicy("internlm", ("internlm", ), "internlm" in native), InjectionPolicy("opt", ("opt", ), True),
        InjectionPolicy("mistral", ("mistral", ), True), InjectionPolicy("mixtral", ("mixtral", ), True),
        InjectionPolicy("falcon", ("falcon", ), True), InjectionPolicy("phi", ("phi", ), True),
        InjectionPolicy("phi3", ("phi3", ), True), InjectionPolicy("qwen", ("qwen", ), True),
        InjectionPolicy("qwen2", ("qwen2", ), True), InjectionPolicy("qwen2_moe", ("qwen2_moe", ), True),
        InjectionPolicy("gpt2", ("gpt2", ), False, "v1 fused layer (module_inject/containers) or AutoTP"),
        InjectionPolicy("gptj", ("gptj", ), False, "v1 fused layer or AutoTP"), InjectionPolicy("gptneo", ("gpt_neo", ), False, "v1 fused layer or AutoTP"),
        InjectionPolicy("gptneox", ("gpt_neox", ), False, "v1 fused layer or AutoTP"), InjectionPolicy("bloom", ("bloom", ), False, "v1 fused layer or AutoTP"),
        InjectionPolicy("bert", ("bert", ), False, "v1 fused layer or AutoTP"), InjectionPolicy("distil_bert", ("distilbert", ), False, "v1 fused layer or AutoTP"),
        InjectionPolicy("megatron_gpt", ("megatron-gpt", ), False, "AutoTP"),
        InjectionPolicy("megatron_gpt_moe", ("megatron-gpt-moe", ), False, "AutoTP + moe.MoE"),
        InjectionPolicy("clip", ("clip", ), False, "generic_injection"), InjectionPolicy("unet", ("unet", ), False, "generic_injection"),
        InjectionPolicy("vae", ("vae", ), False, "generic_injection"),
    ]
    for r in rows:
        if r.supported:
            r.supported = any(t in native for t in r.model_types)
    return {r.name: r for r in rows}


def replace_policies():
    return list(_table().values())


def policy_for(model_type: Optional[str]) -> Optional[InjectionPolicy]:
    for p in _table().values():
        if model_type in p.model_types:
            return p
    return None
