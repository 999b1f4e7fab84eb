"""Model-level injection entry points of the v1 API.

The reference swaps every transformer block for a fused CUDA module through an 11-step container pipeline
(``module_inject/replace_module.py:183-265``). Here a supported architecture is lifted WHOLE onto the ragged sm_100a
executor (``inference/engine.py::_apply_injection``), so ``replace_transformer_layer`` only has to decide between that
path and AutoTP; ``generic_injection`` covers diffusers pipelines by moving their sub-modules to half precision;
``GroupQuantizer`` implements the int8 group quantisation used by ``quant`` configs.
"""
from __future__ import annotations

from typing import Optional

import torch

from ..utils.logging import logger
from .auto_tp import AutoTP
from .policy import policy_for, replace_policies  # noqa: F401


class GroupQuantizer:
    """Symmetric int8 group-wise weight quantiser (``replace_module.py:44-86``)."""

    def __init__(self, q_int8: bool = True, group_size: int = 1, num_bits: int = 8, num_groups: int = 0):
        self.group_size, self.num_bits, self.q_int8, self.num_groups = group_size, num_bits, q_int8, num_groups

    def quantize(self, inputs: torch.Tensor, qkv: bool = True, count: int = 1, parallel_dim: int = 0):
        if not self.q_int8 or not qkv:
            p = torch.nn.Parameter(inputs, requires_grad=False)
            p.scale = None
            return p
        q_range = 2 ** self.num_bits
        num_groups = self.num_groups if self.num_groups > 0 else max(1, inputs.shape[0] // max(1, self.group_size))
        x = inputs.to(torch.float32).contiguous()
        g = x.reshape(num_groups, -1)
        absmax = g.abs().amax(dim=1, keepdim=True).clamp_min(1e-12)
        scale = q_range / (2 * absmax)
        q = (g * scale).round().clamp(-q_range // 2, q_range // 2 - 1).to(torch.int8).reshape(inputs.shape)
        out = torch.nn.Parameter(q, requires_grad=False)
        out.scale = (1.0 / scale).reshape(-1).contiguous()
        return out

    @staticmethod
    def dequantize(q: torch.Tensor, scale: torch.Tensor, dtype=torch.float16) -> torch.Tensor:
        g = q.to(torch.float32).reshape(scale.numel(), -1) * scale.reshape(-1, 1)
        return g.reshape(q.shape).to(dtype)


def generic_injection(module, dtype=None, enable_cuda_graph: bool = True):
    """Diffusers pipelines: cast the heavy sub-modules (unet / vae / text_encoder) to ``dtype`` and put them behind
    CUDA graphs (``DSUNet`` / ``DSVAE`` / ``DSClipEncoder``); their attention runs on torch SDPA. (The reference
    additionally swaps in Triton/CUDA attention blocks, ``replace_module.py:88-181``.)"""
    from ..model_implementations.diffusers import DSUNet, DSVAE
    from ..model_implementations.transformers.clip_encoder import DSClipEncoder
    wrappers = {"unet": DSUNet, "vae": DSVAE, "text_encoder": DSClipEncoder}
    for name in ("unet", "vae", "text_encoder", "transformer"):
        sub = getattr(module, name, None)
        if not isinstance(sub, torch.nn.Module):
            continue
        if dtype is not None:
            sub.to(dtype)
        if name in wrappers and not isinstance(sub, tuple(wrappers.values())):
            try:
                setattr(module, name, wrappers[name](sub, enable_cuda_graph=enable_cuda_graph))
            except Exception:  # pipelines that validate component types keep the bare module
                pass
    return module


def replace_transformer_layer(orig_layer_impl, model: torch.nn.Module, checkpoint_dict=None, config=None,
                              model_config=None) -> torch.nn.Module:
    """Apply tensor slicing to ``model`` according to ``config`` (a ``DeepSpeedInferenceConfig``): AutoTP for models
    without a policy or when kernel injection is off; models with a policy are handled by the inference engine."""
    tp = config.tensor_parallel.tp_size if config is not None else 1
    if tp > 1 and not getattr(config, "replace_with_kernel_inject", False):
        from .. import comm as dist
        group = config.tensor_parallel.tp_group
        rank = dist.get_rank(group=group) if dist.is_initialized() else 0
        AutoTP.apply(model, tp, rank, group)
    return model


def revert_transformer_layer(orig_layer_impl, model: torch.nn.Module, config=None, preln: bool = False) -> torch.nn.Module:
    """Undo AutoTP on a single-rank model (gathers nothing: only valid when tp == 1) — mirrors the reference's
    ``revert_transformer_layer`` (:550) used by tests to restore the HF module."""
    from .layers import LinearAllreduce, LinearLayer
    for name, child in list(model.named_children()):
        if isinstance(child, (LinearLayer, LinearAllreduce)):
            lin = torch.nn.Linear(child.weight.shape[1], child.weight.shape[0], bias=child.bias is not None)
            lin.weight.data = child.weight.data
            if child.bias is not None:
                lin.bias.data = child.bias.data
            setattr(model, name, lin)
        else:
            revert_transformer_layer(orig_layer_impl, child, config, preln)
    return model


# ---- generic traversal helpers of the reference (replace_module.py:29, 619, 662) ------------------------------------------
def replace_module(model: torch.nn.Module, orig_class, replace_fn, _replace_policy=None, checkpoint=None) -> torch.nn.Module:
    """Replace every sub-module of type ``orig_class`` by ``replace_fn(child, _replace_policy, layer_id, prefix=...,
    state_dict=...)``. With ``orig_class=None`` the registered injection policies decide (their block classes are matched
    by class name). ``checkpoint`` (``.pt`` / ``.bin`` / ``.safetensors``) is loaded once and handed to ``replace_fn``."""
    import inspect
    sd = None
    if checkpoint is not None:
        if str(checkpoint).endswith(".safetensors"):
            from safetensors.torch import load_file
            sd = load_file(checkpoint)
        else:
            sd = torch.load(checkpoint, map_location="cpu", weights_only=False)
    from .replace_policy import policy_class_names
    accepts = set(inspect.signature(replace_fn).parameters)
    counter = {"n": 0}

    def wanted(child) -> bool:
        if orig_class is not None:
            return isinstance(child, orig_class) if isinstance(orig_class, type) else type(child).__name__ == str(orig_class)
        return type(child).__name__ in policy_class_names

    def walk(parent, prefix):
        for name, child in list(parent.named_children()):
            full = f"{prefix}.{name}" if prefix else name
            if wanted(child):
                kw = {k: v for k, v in (("prefix", full), ("state_dict", sd)) if k in accepts}
                setattr(parent, name, replace_fn(child, _replace_policy, counter["n"], **kw))
                counter["n"] += 1
            else:
                walk(child, full)

    walk(model, "")
    return model


def get_transformer_name(replaced_module: torch.nn.Module) -> str:
    """Dotted path of the ModuleList that holds the transformer blocks (``"transformer.h"``, ``"model.layers"``, ...)."""
    for n, c in replaced_module.named_children():
        for name, child in c.named_children():
            if isinstance(child, torch.nn.ModuleList):
                return f"{n}.{name}"
    return ""


def skip_level_0_prefix(model, state_dict) -> bool:
    """BLOOM / OPT checkpoints name their tensors without the wrapper's first attribute (``transformer.`` / ``model.``):
    True when loading must skip that level - unless the keys do start with ``model.``."""
    import re
    text = str(model)
    key = re.search(r": (.*?)Model", text) or re.search(r": (.*?)Stack", text) or re.match(r"(.*?)Model", text)
    if state_dict is not None and any(re.match(r"^model[.]", k) for k in state_dict.keys()):
        return False
    return key is not None and key.group(1).lower() in ("bloom", "opt")
