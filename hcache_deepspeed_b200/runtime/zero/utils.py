"""ZeRO helpers (``deepspeed/runtime/zero/utils.py:23-180``): optimizer whitelist, cross-rank consistency asserts (the
``safe_mode`` guard that catches rank divergence BEFORE a collective deadlocks), tensor-tree mapping, flat-buffer maps."""
from __future__ import annotations

from typing import Callable, List, Tuple

import torch

from ... import comm as dist
from .partition_parameters import is_zero_param  # noqa: F401


class ZeRORuntimeException(Exception):
    pass


def is_zero_supported_optimizer(optimizer) -> bool:
    names = {"Adam", "AdamW", "SGD", "Adagrad", "FusedAdam", "FusedLamb", "FusedLion", "DeepSpeedCPUAdam", "DeepSpeedCPUAdagrad",
             "DeepSpeedCPULion", "Lion", "OnebitAdam", "OnebitLamb", "ZeroOneAdam"}
    return type(optimizer).__name__ in names


def get_lst_from_rank0(lst: List[int]) -> List[int]:
    """Rank 0's list, on every rank (lengths must already agree)."""
    t = torch.tensor(lst if dist.get_rank() == 0 else [-1] * len(lst), dtype=torch.int64)
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.broadcast(t, src=0)
    return t.tolist()


def assert_ints_same_as_other_ranks(ints: List[int]) -> None:
    """Raise on every rank that disagrees with rank 0 (e.g. a different parameter fetch order), instead of letting the
    next collective hang."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return
    rank0 = get_lst_from_rank0(ints)
    if list(ints) != rank0:
        raise RuntimeError(f"disagreement between rank0 and rank{dist.get_rank()}: rank0: {rank0}, rank{dist.get_rank()}: {ints}")


def is_builtin_type(obj) -> bool:
    return obj.__class__.__module__ in ("__builtin__", "builtins")


def isinstance_namedtuple(obj: object) -> bool:
    return isinstance(obj, tuple) and hasattr(obj, "_asdict") and hasattr(obj, "_fields")


def apply_to_tensors_only(function: Callable, value, warning_msg_fn=None):
    """Map ``function`` over every tensor in a nest of lists / tuples / named tuples / dicts; other leaves pass through."""
    if isinstance(value, (tuple, list)):
        mapped = [apply_to_tensors_only(function, v, warning_msg_fn) for v in value]
        return value.__class__(*mapped) if isinstance_namedtuple(value) else value.__class__(mapped)
    if isinstance(value, dict):
        return {k: apply_to_tensors_only(function, v, warning_msg_fn) for k, v in value.items()}
    if isinstance(value, torch.Tensor):
        return function(value)
    if warning_msg_fn is not None and not is_builtin_type(value):
        from ...utils.logging import logger
        logger.warning(warning_msg_fn(value))
    return value


def get_mapping_to_flat_buffer(tensors: List[torch.Tensor]) -> List[Tuple[torch.Tensor, int, int]]:
    """``(tensor, offset, numel)`` of every tensor inside the flat buffer that would hold them back to back."""
    out, offset = [], 0
    for t in tensors:
        out.append((t, offset, t.numel()))
        offset += t.numel()
    return out
