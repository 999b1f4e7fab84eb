"""Kernel injection for the v1 fused layer: per-family policies that read one HuggingFace transformer block
(attention / MLP / norm tensors + wiring flags) and the wrapper that puts a
:class:`~hcache_deepspeed_b200.model_implementations.DeepSpeedTransformerInference` in its place.

Reference: ``deepspeed/module_inject/containers/base.py`` (``BaseTransformerContainer``: 11-step create / copy /
slice pipeline), ``module_inject/policy.py`` (``TransformerPolicy.attention / mlp / layernorm``) and
``replace_module.py:210-265`` (``replace_with_policy``). Here a policy is a function ``block -> LayerTensors`` and the
pipeline is: read tensors -> (tensor-parallel slice) -> build the fused layer -> wrap it in the block's call
signature. Families that the v2 ragged executor covers (llama, mistral, opt, falcon, phi, qwen, mixtral ...) are lifted
whole by ``init_inference`` instead; this path serves the encoder / GPT-style families it does not.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, Dict, Optional

import torch
from torch import nn

from ...model_implementations.transformers.ds_transformer import DeepSpeedTransformerInference
from ...ops.transformer.inference import DeepSpeedInferenceConfig


@dataclass
class LayerTensors:
    """Everything the fused layer needs from one block. Weights are ``[in, out]``; QKV columns are ``[q | k | v]``."""
    qkvw: torch.Tensor
    qkvb: Optional[torch.Tensor]
    ow: torch.Tensor
    ob: Optional[torch.Tensor]
    inter_w: torch.Tensor
    inter_b: Optional[torch.Tensor]
    out_w: torch.Tensor
    out_b: Optional[torch.Tensor]
    attn_nw: Optional[torch.Tensor]      # norm feeding the MLP (None: parallel block sharing the input norm)
    attn_nb: Optional[torch.Tensor]
    input_nw: torch.Tensor               # pre-LN: norm feeding attention; post-LN: the block's final norm
    input_nb: Optional[torch.Tensor]
    hidden_size: int
    heads: int
    num_kv: int = -1
    eps: float = 1e-5
    pre_layer_norm: bool = True
    mlp_after_attn: bool = True
    act: str = "gelu"
    norm_type: str = "layer_norm"
    triangular_masking: bool = True
    rotary_dim: int = -1
    rotate_half: bool = False
    rotate_every_two: bool = False
    rope_theta: float = 10000.0
    scale_attn_by_inverse_layer_idx: bool = False
    use_alibi: bool = False
    scale_attention: bool = True


POLICIES: Dict[str, Callable[[nn.Module, object], LayerTensors]] = {}
LAYER_CLASS_NAMES: Dict[str, str] = {}


def register_policy(block_class_name: str, family: str):
    """``@register_policy("GPT2Block", "gpt2")``: the policy applies to HF modules whose class has that name."""

    def deco(fn):
        POLICIES[block_class_name] = fn
        LAYER_CLASS_NAMES[block_class_name] = family
        return fn

    return deco


def t(linear_weight: torch.Tensor) -> torch.Tensor:
    """``nn.Linear`` stores ``[out, in]``; the fused layer multiplies ``x @ W`` with ``W: [in, out]``."""
    return linear_weight.detach().t().contiguous()


def _slice_tp(lt: LayerTensors, rank: int, size: int) -> LayerTensors:
    """Megatron split over ``size`` ranks: QKV / MLP-1 by columns (whole heads), attention-out / MLP-2 by rows; the
    row-parallel biases stay on rank 0 (they are added once, after the all-reduce)."""
    if size == 1:
        return lt
    d = lt.hidden_size // lt.heads
    kv = lt.num_kv if lt.num_kv > 0 else lt.heads
    assert lt.heads % size == 0 and kv % size == 0, "head counts must divide the tensor-parallel size"
    hq, hk = lt.heads // size, kv // size
    q0, k0, v0 = 0, lt.heads * d, (lt.heads + kv) * d
    cols = torch.cat([torch.arange(q0 + rank * hq * d, q0 + (rank + 1) * hq * d),
                      torch.arange(k0 + rank * hk * d, k0 + (rank + 1) * hk * d),
                      torch.arange(v0 + rank * hk * d, v0 + (rank + 1) * hk * d)])
    gated = lt.act.startswith("gated_")
    inter = lt.inter_w.shape[1] // (2 if gated else 1)
    assert inter % size == 0
    step = inter // size
    icols = torch.arange(rank * step, (rank + 1) * step)
    icols_w = torch.cat([icols, icols + inter]) if gated else icols
    zero = (lambda b: None if b is None else (b if rank == 0 else torch.zeros_like(b)))
    return LayerTensors(**{**lt.__dict__,
                           "qkvw": lt.qkvw[:, cols].contiguous(), "qkvb": None if lt.qkvb is None else lt.qkvb[cols].contiguous(),
                           "ow": lt.ow[rank * hq * d:(rank + 1) * hq * d].contiguous(), "ob": zero(lt.ob),
                           "inter_w": lt.inter_w[:, icols_w].contiguous(),
                           "inter_b": None if lt.inter_b is None else lt.inter_b[icols_w].contiguous(),
                           "out_w": lt.out_w[icols].contiguous(), "out_b": zero(lt.out_b)})


def build_fused_layer(lt: LayerTensors, n_layers: int, dtype: torch.dtype, mp_size: int = 1, mp_rank: int = 0, mp_group=None,
                      max_out_tokens: int = 1024, layer_index: int = 0) -> DeepSpeedTransformerInference:
    lt = _slice_tp(lt, mp_rank, mp_size)
    inter = lt.inter_w.shape[1] // (2 if lt.act.startswith("gated_") else 1) * mp_size
    cfg = DeepSpeedInferenceConfig(
        hidden_size=lt.hidden_size, intermediate_size=inter, heads=lt.heads, num_hidden_layers=n_layers,
        layer_norm_eps=lt.eps, mp_size=mp_size, dtype=dtype, pre_layer_norm=lt.pre_layer_norm, norm_type=lt.norm_type,
        triangular_masking=lt.triangular_masking, rotary_dim=lt.rotary_dim, rotate_half=lt.rotate_half,
        rotate_every_two=lt.rotate_every_two, mlp_after_attn=lt.mlp_after_attn, mlp_act_func_type=lt.act,
        num_kv=lt.num_kv, rope_theta=lt.rope_theta, max_out_tokens=max_out_tokens,
        scale_attn_by_inverse_layer_idx=lt.scale_attn_by_inverse_layer_idx, scale_attention=lt.scale_attention,
        return_tuple=False)
    cfg.layer_index = layer_index
    layer = DeepSpeedTransformerInference(cfg, mp_group=mp_group)
    dev = lt.qkvw.device

    def put(param: nn.Parameter, value: Optional[torch.Tensor]):
        if value is None:
            param.data = torch.zeros(param.shape, dtype=dtype, device=dev)
        else:
            assert tuple(param.shape) == tuple(value.shape), f"shape mismatch {tuple(param.shape)} vs {tuple(value.shape)}"
            param.data = value.detach().to(dtype).contiguous()

    a, m = layer.attention, layer.mlp
    put(a.attn_qkvw, lt.qkvw), put(a.attn_qkvb, lt.qkvb), put(a.attn_ow, lt.ow), put(a.attn_ob, lt.ob)
    put(m.inter_w, lt.inter_w), put(m.inter_b, lt.inter_b), put(m.output_w, lt.out_w), put(m.output_b, lt.out_b)
    m.has_own_norm = lt.attn_nw is not None
    if lt.attn_nw is not None:
        put(m.attn_nw, lt.attn_nw)
        if m.attn_nb is not None:
            put(m.attn_nb, lt.attn_nb)
    put(layer.norm_w, lt.input_nw)
    if layer.norm_b is not None:
        put(layer.norm_b, lt.input_nb)
    return layer


def _find_cache(args, kwargs):
    for v in list(kwargs.values()) + list(args):
        if hasattr(v, "get_seq_length") and hasattr(v, "update"):
            return v
    return None


def _find_mask(args, kwargs):
    m = kwargs.get("attention_mask")
    if m is None:
        for v in args:
            if torch.is_tensor(v) and v.dim() == 4:
                return v
    return m


class InjectedLayer(nn.Module):
    """Stands where the HF block stood: accepts the block's call (positional / keyword mix differs per family), runs the
    fused layer, returns what the model's loop expects (a tensor in transformers >= 4.56, a 1-tuple before).

    KV state lives in the fused layer's own contiguous cache. When the model passes a HF ``Cache`` the new tokens' K/V
    are also recorded there so that ``generate`` sees the right sequence length and feeds one token per step."""

    def __init__(self, fused: DeepSpeedTransformerInference, layer_idx: int, use_alibi: bool = False, returns_tuple: bool = False):
        super().__init__()
        self.fused, self.layer_idx, self.use_alibi, self.returns_tuple = fused, layer_idx, use_alibi, returns_tuple

    def forward(self, hidden_states, *args, **kwargs):
        cache = _find_cache(args, kwargs)
        mask = _find_mask(args, kwargs)
        if mask is not None and mask.dtype == torch.bool:
            mask = torch.zeros_like(mask, dtype=hidden_states.dtype).masked_fill(~mask, torch.finfo(hidden_states.dtype).min)
        past = cache.get_seq_length(self.layer_idx) if cache is not None else 0
        alibi = kwargs.get("alibi") if self.use_alibi else None
        out, (k, v) = self.fused(hidden_states, input_mask=mask, alibi=alibi, get_present=True, is_prompt=past == 0)
        if cache is not None:
            s = hidden_states.shape[1]
            cache.update(k[:, :, -s:], v[:, :, -s:], self.layer_idx)
        return (out, ) if self.returns_tuple else out


def _block_returns_tuple(block: nn.Module) -> bool:
    """Whether the HF block hands back ``(hidden_states, ...)`` or the bare tensor (it changed per family across
    transformers releases): read off the last ``return`` of its ``forward``."""
    import inspect
    import re
    try:
        src = inspect.getsource(type(block).forward)
    except (OSError, TypeError):
        return False
    returns = [ln.strip() for ln in src.splitlines() if ln.strip().startswith("return")]
    return bool(returns) and re.fullmatch(r"return\s+\w+", returns[-1]) is None


def inject_fused_layers(model: nn.Module, dtype: Optional[torch.dtype] = None, mp_size: int = 1, mp_rank: int = 0, mp_group=None,
                        max_out_tokens: int = 1024, config=None) -> int:
    """Replace every block of ``model`` that has a registered policy with an :class:`InjectedLayer`; returns how many
    were replaced (0: no policy knows this architecture)."""
    from . import families  # noqa: F401  (registers the policies)
    cfg = config if config is not None else getattr(model, "config", None)
    targets = [(name, mod) for name, mod in model.named_modules() if type(mod).__name__ in POLICIES and name]
    if not targets:
        return 0
    dtype = dtype or next(model.parameters()).dtype
    n_layers = len(targets)
    for idx, (name, block) in enumerate(targets):
        lt = POLICIES[type(block).__name__](block, cfg)
        fused = build_fused_layer(lt, n_layers, dtype, mp_size, mp_rank, mp_group, max_out_tokens, layer_index=idx)
        parent_name, _, leaf = name.rpartition(".")
        parent = model.get_submodule(parent_name) if parent_name else model
        wrapped = InjectedLayer(fused, idx, use_alibi=lt.use_alibi, returns_tuple=_block_returns_tuple(block))
        if isinstance(parent, (nn.ModuleList, nn.Sequential)):
            parent[int(leaf)] = wrapped
        else:
            setattr(parent, leaf, wrapped)
    return n_layers
