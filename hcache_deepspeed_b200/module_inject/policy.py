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


# ---- the reference's policy base classes and per-tensor checkpoint copy helpers (module_inject/policy.py:26-224) -----------
import abc  # noqa: E402

import torch  # noqa: E402


class DSPolicy(abc.ABC):
    """What a kernel-injection policy has to expose about one transformer block."""
    _orig_layer_class = None

    def __init__(self):
        self.cuda_graph_supported = False

    @abc.abstractmethod
    def attention(self):
        """(qkv weight ``[3*hidden, hidden]``, qkv bias, dense weight ``[hidden, hidden]``, dense bias)."""
        raise NotImplementedError


class TransformerPolicy(DSPolicy):
    """Policy of a standard block: subclasses return the block's tensors through ``attention`` / ``mlp`` / ``layernorm``
    and its sizes through ``get_hidden_heads``. The flags describe the block to the fused layer."""
    hf_model_config = None

    def __init__(self, inference: bool = True, linear_layer: bool = True, scale_attention: bool = True, megatron_v2: bool = False,
                 use_mup: bool = False, mlp_act_func_type=None, pre_attn_norm: bool = True, use_load_prefix: bool = False,
                 split_qkv: bool = True, norm_type=None):
        super().__init__()
        self.inference, self.linear_layer, self.scale_attention = inference, linear_layer, scale_attention
        self.is_megatron_v2, self.use_mup, self.mlp_act_func_type = megatron_v2, use_mup, mlp_act_func_type
        self.pre_attn_norm, self.use_load_prefix, self.split_qkv, self.norm_type = pre_attn_norm, use_load_prefix, split_qkv, norm_type

    @abc.abstractmethod
    def get_hidden_heads(self):
        """(hidden size, attention heads, layer-norm epsilon, intermediate size)."""
        raise NotImplementedError

    @abc.abstractmethod
    def mlp(self):
        """(up weight, up bias, down weight, down bias)."""
        raise NotImplementedError

    @abc.abstractmethod
    def layernorm(self):
        """(post-attention norm gamma, beta, input norm gamma, beta)."""
        raise NotImplementedError


def transpose(data: torch.Tensor) -> torch.Tensor:
    """The last two dimensions swapped, written back into the SAME storage (no second full-size tensor stays alive)."""
    with torch.no_grad():
        data = data.contiguous()
        flipped = data.transpose(-1, -2).reshape(-1).clone()
        data.reshape(-1).copy_(flipped)
    return data.reshape(*data.shape[:-2], data.shape[-1], data.shape[-2])


def _regroup_qkv_by_head(x: torch.Tensor, heads: int, mp_replace) -> torch.Tensor:
    """Megatron-v2 stores fused QKV interleaved per head ``[h0:(q,k,v), h1:(q,k,v), ...]``; the fused layer wants
    ``[q | k | v]``. Works on the last dimension."""
    heads = heads // mp_replace.mp_size
    per_head = x.shape[-1] // heads
    q, k, v = torch.split(x.reshape(*x.shape[:-1], heads, per_head), per_head // 3, dim=-1)
    lead = x.shape[:-1]
    return torch.cat((q.reshape(*lead, -1), k.reshape(*lead, -1), v.reshape(*lead, -1)), dim=-1).reshape(x.shape)


def _quantize_or_transpose(weight_quantizer, t: torch.Tensor) -> torch.Tensor:
    if getattr(weight_quantizer, "q_int8", False):
        return weight_quantizer.quantize(t)
    t = transpose(t).contiguous()
    return weight_quantizer.quantize(t) if weight_quantizer is not None else t


def maybe_copy(module, sd, weight_quantizer, mp_replace, dst_name: str, src_name: str, qkv: bool = False,
               megatron_v2: bool = False, split_qkv: bool = False, heads: int = 1) -> None:
    """If ``src_name`` is in this checkpoint shard: copy this tensor-parallel rank's part into ``module.<dst_name>``
    (weights go in transposed - ``[in, out]`` - or int8-quantised; fused QKV is cut per part; Megatron-v2 QKV is
    regrouped from per-head interleaving). Shards that do not hold the tensor are skipped silently."""
    if src_name not in sd:
        return
    dst, tmp = getattr(module, dst_name), sd[src_name]
    int8 = bool(getattr(weight_quantizer, "q_int8", False))
    if dst.dim() == 1:
        dst = mp_replace.strided_copy(dst, tmp, num_splits=3) if split_qkv else mp_replace.copy(dst, tmp)
        if qkv and megatron_v2:
            dst = torch.nn.Parameter(_regroup_qkv_by_head(dst, heads, mp_replace).contiguous(), requires_grad=False)
    elif split_qkv:
        dst = mp_replace.strided_copy(dst, _quantize_or_transpose(weight_quantizer, tmp), num_splits=3, int8=int8)
    else:
        if qkv and megatron_v2:
            tmp = _regroup_qkv_by_head(transpose(tmp), heads, mp_replace).contiguous()
            tmp = transpose(tmp) if not int8 else tmp
        dst = mp_replace.copy(dst, _quantize_or_transpose(weight_quantizer, tmp), int8=int8)
    setattr(module, dst_name, dst)


def maybe_copy_qkv(module, sd, weight_quantizer, mp_replace, dst_name: str, src_names, split_qkv: bool = False) -> None:
    """Like ``maybe_copy`` for checkpoints that store Q, K and V separately: they are stacked first."""
    if src_names[0] not in sd:
        return
    fused = torch.cat([sd[n] for n in src_names[:3]], dim=0)
    dst = getattr(module, dst_name)
    int8 = bool(getattr(weight_quantizer, "q_int8", False))
    if dst.dim() == 1:
        dst = mp_replace.strided_copy(dst, fused.contiguous(), num_splits=3) if split_qkv else mp_replace.copy(dst, fused)
    elif split_qkv:
        dst = mp_replace.strided_copy(dst, _quantize_or_transpose(weight_quantizer, fused), num_splits=3, int8=int8)
    else:
        dst = mp_replace.copy(dst, _quantize_or_transpose(weight_quantizer, fused), int8=int8)
    setattr(module, dst_name, dst)


def maybe_copy_geglu(module, sd, weight_quantizer, mp_replace, dst_name: str, src_names) -> None:
    """GeGLU first MLP matrix stored as two tensors (projection, gate): stacked, then cut per half for this rank."""
    if src_names[0] not in sd:
        return
    fused = torch.cat((sd[src_names[0]], sd[src_names[1]]), dim=0)
    int8 = bool(getattr(weight_quantizer, "q_int8", False))
    dst = mp_replace.strided_copy(getattr(module, dst_name), _quantize_or_transpose(weight_quantizer, fused), num_splits=2, int8=int8)
    setattr(module, dst_name, dst)


def pack_lora_weights(p):
    return [p.lora_right_weight, p.lora_left_weight, p.lora_scaling]


def maybe_get_lora(p):
    """The LoRA triple hung on a parameter by DeepSpeed-Chat style adapters, or an empty list."""
    return pack_lora_weights(p) if hasattr(p, "lora_right_weight") else []
