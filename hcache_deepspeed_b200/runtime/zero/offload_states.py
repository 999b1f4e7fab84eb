"""Moving optimizer / parameter states off the accelerator and back between phases (e.g. while another model uses the
GPU in RLHF): ``deepspeed/runtime/zero/offload_states.py:15-86`` and ``offload_config.py`` (``OffloadStateTypeEnum``)."""
from __future__ import annotations

from enum import Enum
from typing import Set

import torch


class OffloadStateTypeEnum(str, Enum):
    optim_states = "optim_states"
    hp_params = "hp_params"
    lp_params = "lp_params"
    lp_grads = "lp_grads"
    contiguous_grad_buffer = "contiguous_grad_buffer"


def _make_offload_state_key(key: str) -> str:
    return f"{key}_offload_buffer"


def offload_adam_states(optimizer, device, pin_memory: bool = False, non_blocking: bool = False) -> None:
    """``exp_avg`` / ``exp_avg_sq`` of every parameter -> ``device`` (buffers are created once and reused)."""
    for state in optimizer.state.values():
        for key in ("exp_avg", "exp_avg_sq"):
            if key not in state:
                continue
            buf_key = _make_offload_state_key(key)
            if buf_key not in state:
                buf = torch.empty_like(state[key], device=device)
                state[buf_key] = buf.pin_memory() if pin_memory and torch.cuda.is_available() and buf.device.type == "cpu" else buf
            state[buf_key].copy_(state[key], non_blocking=non_blocking)
            state[key].data = state[buf_key]


def reload_adam_states(optimizer, device, non_blocking: bool = False) -> None:
    for state in optimizer.state.values():
        for key in ("exp_avg", "exp_avg_sq"):
            if key in state and _make_offload_state_key(key) in state:
                state[key].data = state[_make_offload_state_key(key)].to(device, non_blocking=non_blocking)


def get_state_devices(model, state: OffloadStateTypeEnum) -> Set[torch.device]:
    """Devices on which the given state of an engine currently lives."""
    opt = getattr(model, "optimizer", model)
    base = getattr(opt, "optimizer", opt)
    if state == OffloadStateTypeEnum.optim_states:
        return {t.device for st in base.state.values() for k, t in st.items()
                if torch.is_tensor(t) and k in ("exp_avg", "exp_avg_sq")}
    if state == OffloadStateTypeEnum.lp_params:
        return {p.device for p in model.module.parameters()} if hasattr(model, "module") else set()
    if state == OffloadStateTypeEnum.hp_params:
        groups = getattr(opt, "fp32_partitions", None) or getattr(opt, "single_partition_of_fp32_groups", None) or []
        return {t.device for t in groups}
    if state == OffloadStateTypeEnum.lp_grads:
        return {p.grad.device for p in model.module.parameters() if p.grad is not None} if hasattr(model, "module") else set()
    return set()
