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


# ---- whole-engine state movement (engine.offload_states / reload_states) -------------------------------------------------
def _engine_state_tensors(engine, kind: OffloadStateTypeEnum):
    """The tensors that make up one kind of state of ``engine``, whichever ZeRO stage (or none) it runs."""
    from .partition_parameters import is_zero_param
    opt = engine.optimizer
    base = getattr(opt, "optimizer", opt)
    if kind == OffloadStateTypeEnum.optim_states:
        return [t for st in getattr(base, "state", {}).values() for k, t in st.items()
                if torch.is_tensor(t) and t.dim() > 0 and not k.endswith("_offload_buffer")]
    if kind == OffloadStateTypeEnum.hp_params:
        return [t for t in (getattr(opt, "fp32_partitioned_groups_flat", None) or getattr(opt, "single_partition_of_fp32_groups", None)
                            or getattr(opt, "fp32_groups_flat", None) or []) if t.numel()]
    if kind == OffloadStateTypeEnum.lp_params:
        flats, _ = _flat_lp_layout(opt)
        if flats is not None:      # parameters are views of flat buffers: the buffers move, the views are re-made
            return [f for f in flats if f.numel()]
        out = []
        for p in engine.module.parameters():
            if is_zero_param(p):
                if getattr(p, "ds_tensor", None) is not None and not getattr(p, "ds_nvme", False):
                    out.append(p.ds_tensor)
            else:
                out.append(p)
        return out
    if kind == OffloadStateTypeEnum.lp_grads:
        parts = getattr(opt, "grad_partitions", None)
        if isinstance(parts, dict) and parts:
            return [g for g in parts.values() if torch.is_tensor(g)]
        return [p.grad for p in engine.module.parameters() if p.grad is not None]
    if kind == OffloadStateTypeEnum.contiguous_grad_buffer:
        buf = getattr(opt, "ipg_buffer", None) or getattr(opt, "_grad_flat", None)
        return [b for b in (buf if isinstance(buf, (list, tuple)) else [buf]) if torch.is_tensor(b)]
    return []


def _flat_lp_layout(opt):
    """(flat low-precision buffers, parameter lists) when the optimizer keeps its parameters as views of flat buffers
    (ZeRO-1/2, bf16 and fp16 wrappers), else (None, None)."""
    for flat_name, group_name in (("bit16_groups_flat", "bit16_groups"), ("fp16_groups_flat", "fp16_groups")):
        flats, groups = getattr(opt, flat_name, None), getattr(opt, group_name, None)
        if flats is not None and groups is not None:
            return flats, groups
    return None, None


def _rebind_lp_views(opt) -> None:
    """Point every parameter at its range of the (moved) flat buffer again."""
    flats, groups = _flat_lp_layout(opt)
    if flats is None:
        return
    for flat, params in zip(flats, groups):
        off = 0
        for p in params:
            if flat.numel():
                p.data = flat.data.narrow(0, off, p.numel()).view(p.shape)
            off += p.numel()


def offload_engine_states(engine, include=None, device=torch.device("cpu"), pin_memory: bool = True, non_blocking: bool = False) -> None:
    """Move the chosen kinds of state to ``device``, remembering where each tensor lived. Host buffers are kept across
    calls (pinned when asked and possible) so repeated offload / reload cycles - one per RLHF phase - do not re-allocate."""
    kinds = [OffloadStateTypeEnum(k) for k in include] if include is not None else list(OffloadStateTypeEnum)
    record = engine.__dict__.setdefault("_offloaded_states", {})      # id(tensor) -> (tensor, home device, kind)
    buffers = engine.__dict__.setdefault("_offload_host_buffers", {})  # id(tensor) -> host buffer
    can_pin = pin_memory and torch.cuda.is_available() and device.type == "cpu"
    for kind in kinds:
        for t in _engine_state_tensors(engine, kind):
            if id(t) in record:
                continue
            buf = buffers.get(id(t))
            if buf is None or buf.shape != t.shape or buf.dtype != t.dtype:
                buf = torch.empty(t.shape, dtype=t.dtype, device=device)
                if can_pin:
                    buf = buf.pin_memory()
                buffers[id(t)] = buf
            buf.copy_(t.data, non_blocking=non_blocking)
            record[id(t)] = (t, t.device, kind)
            t.data = buf
    if OffloadStateTypeEnum.lp_params in kinds:
        _rebind_lp_views(engine.optimizer)
    if torch.cuda.is_available() and not non_blocking:
        torch.cuda.synchronize()
        torch.cuda.empty_cache()


def reload_engine_states(engine, non_blocking: bool = False) -> None:
    record = engine.__dict__.get("_offloaded_states") or {}
    for t, home, _ in record.values():
        t.data = t.data.to(home, non_blocking=non_blocking) if t.data.device != home else t.data.clone()
    touched_lp = any(kind == OffloadStateTypeEnum.lp_params for _, _, kind in record.values())
    record.clear()
    if touched_lp:
        _rebind_lp_views(engine.optimizer)
    if torch.cuda.is_available() and not non_blocking:
        torch.cuda.synchronize()


def offloaded_state_kinds(engine) -> Set[OffloadStateTypeEnum]:
    return {kind for _, _, kind in (engine.__dict__.get("_offloaded_states") or {}).values()}
