"""Random layer-token-drop: middle layers process a random subset of tokens whose size grows on a schedule.
``RandomLTDScheduler`` (kept-token count per step), ``RandomLayerTokenDrop`` (wrapper that gathers the kept tokens,
runs the layer, scatters the result back) and the token-sort / gather / scatter ops.
Reference: ``deepspeed/runtime/data_pipeline/data_routing/{scheduler,basic_layer,helper}.py``, ``csrc/random_ltd/*.cu``."""
from __future__ import annotations

import math
from typing import Dict

import torch


def token_sort_(reserved_length: int, seq_length: int, batch: int, layers: int, device="cpu", generator=None) -> torch.Tensor:
    """[layers, batch, reserved_length] sorted random token indices (one independent subset per layer and sample)."""
    noise = torch.rand(layers, batch, seq_length, device=device, generator=generator)
    idx = noise.argsort(dim=-1)[..., :reserved_length]
    return idx.sort(dim=-1).values


def gather_tokens(x: torch.Tensor, idx: torch.Tensor, batch_first: bool = True) -> torch.Tensor:
    """x: [b, s, h] (or [s, b, h]); idx: [b, r] -> [b, r, h]."""
    if not batch_first:
        x = x.transpose(0, 1)
    out = torch.gather(x, 1, idx.unsqueeze(-1).expand(-1, -1, x.shape[-1]))
    return out if batch_first else out.transpose(0, 1)


def scatter_tokens(full: torch.Tensor, part: torch.Tensor, idx: torch.Tensor, batch_first: bool = True) -> torch.Tensor:
    if not batch_first:
        full, part = full.transpose(0, 1), part.transpose(0, 1)
    out = full.scatter(1, idx.unsqueeze(-1).expand(-1, -1, full.shape[-1]), part)
    return out if batch_first else out.transpose(0, 1)


class RandomLTDScheduler:

    def __init__(self, config: Dict):
        self.model_layer_num = config.get("total_layer_num", config.get("random_ltd_layer_num", 0))
        self.random_ltd_layer_num = config.get("random_ltd_layer_num", 0)
        self.config_schedule = config.get("random_ltd_schedule", {})
        self.global_batch_size = config.get("global_batch_size", 1)
        sc = self.config_schedule.get("schedule_config", {})
        self.state = {"min_value": self.config_schedule.get("min_value", 128), "max_value": self.config_schedule.get("max_value", 2048),
                      "current_value": self.config_schedule.get("min_value", 128), "schedule_type": self.config_schedule.get("schedule_type", "fixed_linear"),
                      "require_steps": sc.get("require_steps", 1000), "seq_per_step": sc.get("seq_per_step", 16),
                      "consumed_layer_tokens": 0, "curr_step": -1}

    def get_total_layer_tokens(self, train_iters: int) -> int:
        total = 0
        for step in range(train_iters):
            self.update_seq(step)
            total += self.state["consumed_layer_tokens_this_step"]
        return total

    def get_current_seq(self) -> int:
        return self.state["current_value"]

    def set_current_seq(self, v: int):
        self.state["current_value"] = v

    def get_value(self, global_steps: int) -> int:
        s = self.state
        v = math.floor(global_steps / s["require_steps"] * (s["max_value"] - s["min_value"]) + s["min_value"])
        v -= v % s["seq_per_step"]
        return min(max(v, s["min_value"]), s["max_value"])

    def update_seq(self, global_steps: int) -> int:
        s = self.state
        if s["current_value"] < s["max_value"]:
            s["current_value"] = self.get_value(global_steps)
        if global_steps != s["curr_step"]:
            this = self.global_batch_size * (s["current_value"] * self.random_ltd_layer_num +
                                             s["max_value"] * (self.model_layer_num - self.random_ltd_layer_num))
            s["consumed_layer_tokens"] += this
            s["consumed_layer_tokens_this_step"] = this
            s["curr_step"] = global_steps
        return s["current_value"]

    def state_dict(self):
        return dict(self.state)

    def load_state_dict(self, sd):
        self.state.update(sd)


class RandomLayerTokenDrop(torch.nn.Module):
    """Wrap a transformer layer: during training it only sees ``scheduler.get_current_seq()`` random tokens."""

    def __init__(self, layer: torch.nn.Module, layer_id: int = 0, scheduler: RandomLTDScheduler = None, batch_first: bool = True):
        super().__init__()
        self.random_ltd_layer, self.layer_id, self.scheduler, self.batch_first = layer, layer_id, scheduler, batch_first

    def forward(self, hidden_states, *args, **kwargs):
        if not self.training or self.scheduler is None:
            return self.random_ltd_layer(hidden_states, *args, **kwargs)
        b, s = (hidden_states.shape[0], hidden_states.shape[1]) if self.batch_first else (hidden_states.shape[1], hidden_states.shape[0])
        keep = min(self.scheduler.get_current_seq(), s)
        if keep >= s:
            return self.random_ltd_layer(hidden_states, *args, **kwargs)
        idx = token_sort_(keep, s, b, 1, device=hidden_states.device)[0]
        part = gather_tokens(hidden_states, idx, self.batch_first)
        out = self.random_ltd_layer(part, *args, **kwargs)
        rest = out[1:] if isinstance(out, tuple) else None
        out0 = out[0] if isinstance(out, tuple) else out
        full = scatter_tokens(hidden_states, out0, idx, self.batch_first)
        return (full, ) + tuple(rest) if rest is not None else full


def convert_to_random_ltd(model, convert_type, scheduler: RandomLTDScheduler = None, batch_first: bool = True):
    """Wrap every sub-module of type ``convert_type`` in :class:`RandomLayerTokenDrop`
    (``data_routing/helper.py:11-25``). ``model`` may be an engine (``.module`` is then converted)."""
    root = model.module if hasattr(model, "module") and isinstance(model.module, torch.nn.Module) else model
    targets = [(n, m) for n, m in root.named_modules() if isinstance(m, convert_type) and n]
    for layer_id, (name, mod) in enumerate(targets):
        parent_name, _, leaf = name.rpartition(".")
        parent = root.get_submodule(parent_name) if parent_name else root
        setattr(parent, leaf, RandomLayerTokenDrop(mod, layer_id, scheduler, batch_first))
    if hasattr(model, "random_ltd_initialize"):
        model.random_ltd_initialize()
    return model


def remove_random_ltd_state_dict(state_dict):
    """Checkpoint keys without the wrapper level, so the file loads into the unconverted model."""
    return type(state_dict)((k.replace(".random_ltd_layer", ""), v) for k, v in state_dict.items())


def save_without_random_ltd(model):
    root = model.module if hasattr(model, "module") and isinstance(model.module, torch.nn.Module) else model
    return remove_random_ltd_state_dict(root.state_dict())
