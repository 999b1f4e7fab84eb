"""Helpers to separate expert parameters from dense ones (they use different data-parallel groups for gradient
reduction and norms). Reference: ``deepspeed/moe/utils.py:15-180``."""
from __future__ import annotations

from typing import Dict, List, Tuple

import torch


def has_moe_layers(m: torch.nn.Module) -> Tuple[bool, int]:
    from .layer import MoE
    n = 0
    found = False
    for mod in m.modules():
        if isinstance(mod, MoE):
            found = True
            n = mod.num_experts
            break
    return found, n


def is_moe_param(param: torch.Tensor) -> bool:
    return hasattr(param, "allreduce") and not param.allreduce


def split_params_into_shared_and_expert_params(params):
    shared, expert = [], []
    for p in params:
        (expert if is_moe_param(p) else shared).append(p)
    return shared, expert


def split_params_grads_into_shared_and_expert_params(group):
    shared, expert = [], []
    for p in group:
        if p.grad is not None:
            (expert if is_moe_param(p) else shared).append(p.grad.to(p.dtype))
    return shared, expert


def split_params_into_different_moe_groups_for_optimizer(param_groups, max_group_size=178956971) -> List[Dict]:
    """Return optimizer param groups where expert parameters of each expert group get their own group
    (``moe=True``, ``name=<group>``), so ZeRO partitions them over the expert-data-parallel group."""
    if isinstance(param_groups, tuple):
        param_groups = list(param_groups)
    elif isinstance(param_groups, dict):
        param_groups = [param_groups]
    elif not isinstance(param_groups, list):
        raise ValueError(f"Unknown param group type of {type(param_groups)}")
    out = []
    moe_groups: Dict[str, Dict] = {}
    for g in param_groups:
        dense = []
        for p in g["params"]:
            if is_moe_param(p):
                name = p.group_name
                if name not in moe_groups:
                    ng = {k: v for k, v in g.items() if k != "params"}
                    ng.update(name=name, moe=True, params=[])
                    moe_groups[name] = ng
                moe_groups[name]["params"].append(p)
            else:
                dense.append(p)
        ng = dict(g)
        ng["params"] = dense
        out.append(ng)
    for name, g in moe_groups.items():
        cur, size = [], 0
        for p in g["params"]:
            if size + p.numel() > max_group_size and cur:
                out.append({**g, "params": cur})
                cur, size = [], 0
            cur.append(p)
            size += p.numel()
        if cur:
            out.append({**g, "params": cur})
    return out


def is_moe_param_group(param_group) -> bool:
    return param_group.get("moe", False)


def configure_moe_param_groups(model_parameters: List):
    assert isinstance(model_parameters, list), "model_parameters must be a list"
    if not model_parameters:
        return model_parameters
    if isinstance(model_parameters[0], torch.nn.Parameter):
        return split_params_into_different_moe_groups_for_optimizer({"params": model_parameters})
    if isinstance(model_parameters[0], dict):
        return split_params_into_different_moe_groups_for_optimizer(model_parameters)
    return model_parameters
