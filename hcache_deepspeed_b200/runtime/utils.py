"""Runtime helpers: gradient norms / clipping, overflow checks, partitioning, flatten helpers, memory report.
Behavioural reference: ``deepspeed/runtime/utils.py`` (clip_grad_norm_ :327, get_global_norm_of_tensors :862,
partition_uniform :577, partition_balanced :643, see_memory_usage :771, CheckOverflow :182)."""
from __future__ import annotations

import math
from bisect import bisect_left
from typing import Iterable, List, Optional, Sequence

import torch

from .. import comm as dist
from ..utils.logging import logger


def noop_decorator(func):
    return func


def flatten_dense_tensors(tensors: Sequence[torch.Tensor]) -> torch.Tensor:
    return torch.cat([t.contiguous().view(-1) for t in tensors]) if len(tensors) else torch.empty(0)


def unflatten_dense_tensors(flat: torch.Tensor, tensors: Sequence[torch.Tensor]) -> List[torch.Tensor]:
    outs, off = [], 0
    for t in tensors:
        n = t.numel()
        outs.append(flat.narrow(0, off, n).view_as(t))
        off += n
    return outs


def ensure_directory_exists(filename: str):
    import os
    os.makedirs(os.path.dirname(filename) or ".", exist_ok=True)


def set_random_seed(seed: int):
    import random
    import numpy as np
    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)


def is_model_parallel_parameter(p) -> bool:
    return bool(getattr(p, "model_parallel", False) or getattr(p, "tensor_model_parallel", False))


def get_global_norm(norm_list: Iterable[float]) -> float:
    return math.sqrt(sum(n ** 2 for n in norm_list))


def _all_reduce_scalar(value: float, op, group, device) -> float:
    if not dist.is_initialized() or (group is not None and dist.get_world_size(group) == 1):
        return value
    t = torch.tensor([float(value)], dtype=torch.float32, device=device)
    dist.all_reduce(t, op=op, group=group)
    return float(t[0])


def get_grad_norm(parameters, norm_type: float = 2, mpu=None) -> float:
    """Global gradient norm over ``parameters`` (model-parallel replicas counted once, summed over the MP group)."""
    if isinstance(parameters, torch.Tensor):
        parameters = [parameters]
    params = [p for p in parameters if p.grad is not None]
    if not params:
        return 0.0
    dev = params[0].grad.device
    mp_rank = mpu.get_model_parallel_rank() if mpu is not None else 0
    mp_group = mpu.get_model_parallel_group() if mpu is not None else None
    if norm_type == math.inf:
        total = max(float(p.grad.detach().abs().max()) for p in params)
        if mpu is not None:
            total = _all_reduce_scalar(total, dist.ReduceOp.MAX, mp_group, dev)
    else:
        total = 0.0
        for p in params:
            if mpu is not None and mp_rank > 0 and not is_model_parallel_parameter(p):
                continue
            total += float(p.grad.detach().float().norm(norm_type)) ** norm_type
        if mpu is not None:
            total = _all_reduce_scalar(total, dist.ReduceOp.SUM, mp_group, dev)
        total = total ** (1.0 / norm_type)
    if total in (float("inf"), -float("inf")) or total != total:
        total = -1.0
    return total


def get_global_norm_of_tensors(input_tensors, norm_type: float = 2, mpu=None, use_graph: bool = False,
                               moe_ep_group=None) -> float:
    tensors = [t for t in input_tensors if t is not None]
    if not tensors:
        return 0.0
    dev = tensors[0].device
    if norm_type == math.inf:
        total = max(float(t.detach().abs().max()) for t in tensors)
        for grp in ((mpu.get_model_parallel_group() if mpu is not None else None), moe_ep_group):
            if grp is not None:
                total = _all_reduce_scalar(total, dist.ReduceOp.MAX, grp, dev)
    else:
        total = sum(float(t.detach().float().norm(norm_type)) ** norm_type for t in tensors)
        for grp in ((mpu.get_model_parallel_group() if mpu is not None else None), moe_ep_group):
            if grp is not None:
                total = _all_reduce_scalar(total, dist.ReduceOp.SUM, grp, dev)
        total = total ** (1.0 / norm_type)
    if total in (float("inf"), -float("inf")) or total != total:
        total = -1.0
    return total


def clip_grad_norm_(parameters, max_norm: float, norm_type: float = 2, mpu=None) -> float:
    """Clip in place to ``max_norm``; returns the norm before clipping."""
    if isinstance(parameters, torch.Tensor):
        parameters = [parameters]
    params = [p for p in parameters if p.grad is not None]
    total = get_grad_norm(params, norm_type, mpu)
    if total > 0:
        coef = max_norm / (total + 1e-6)
        if coef < 1:
            for p in params:
                p.grad.detach().mul_(coef)
    return total


def clip_tensors_by_global_norm(input_tensors, max_norm: float = 1.0, global_norm: Optional[float] = None, mpu=None,
                                eps: float = 1e-6, use_graph: bool = False) -> float:
    if global_norm is None:
        global_norm = get_global_norm_of_tensors(input_tensors, mpu=mpu)
    coef = max_norm / (global_norm + eps)
    if coef < 1:
        for t in input_tensors:
            t.detach().mul_(coef)
    return global_norm


class CheckOverflow:
    """inf/nan detection over a parameter list, agreed on by every data/model-parallel rank (MAX all-reduce)."""

    def __init__(self, param_groups=None, mpu=None, zero_reduce_scatter: bool = False, deepspeed=None):
        self.mpu, self.deepspeed = mpu, deepspeed
        self.params = [p for g in (param_groups or []) for p in g]
        self.zero_reduce_scatter = zero_reduce_scatter

    @staticmethod
    def _has_inf_or_nan(x: torch.Tensor) -> bool:
        s = float(x.float().sum())
        return s in (float("inf"), -float("inf")) or s != s

    def has_overflow_serial(self, params) -> bool:
        return any(p.grad is not None and self._has_inf_or_nan(p.grad.data) for p in params)

    def has_overflow(self, params, has_moe_params: bool = False) -> bool:
        overflow = self.has_overflow_serial(params)
        if not dist.is_initialized():
            return overflow
        dev = next((p.device for p in params), torch.device("cpu"))
        flag = torch.tensor([1.0 if overflow else 0.0], device=dev)
        if self.zero_reduce_scatter:
            dist.all_reduce(flag, op=dist.ReduceOp.MAX)
        elif self.mpu is not None:
            dist.all_reduce(flag, op=dist.ReduceOp.MAX, group=self.mpu.get_model_parallel_group())
        return bool(flag[0] > 0)

    def check(self, param_groups=None) -> bool:
        params = [p for g in param_groups for p in g] if param_groups is not None else self.params
        return self.has_overflow(params)


def partition_uniform(num_items: int, num_parts: int) -> List[int]:
    """Boundaries ``parts[p] .. parts[p+1]`` of an even split; the remainder goes to the first parts."""
    parts = [0] * (num_parts + 1)
    if num_items <= num_parts:
        for p in range(num_parts + 1):
            parts[p] = min(p, num_items)
        return parts
    chunk, rem = divmod(num_items, num_parts)
    for p in range(num_parts):
        parts[p + 1] = parts[p] + chunk + (1 if p < rem else 0)
    return parts


def prefix_sum_inc(weights: Sequence[float]) -> List[float]:
    out, s = [], 0
    for w in weights:
        s += w
        out.append(s)
    return out


def _lprobe(weights_psum: List[float], num_parts: int, bottleneck: float):
    """Greedy probe: can the items be split into ``num_parts`` contiguous parts each weighing <= bottleneck?"""
    n = len(weights_psum)
    parts, total, bsum = [0] * (num_parts + 1), weights_psum[-1], bottleneck
    for p in range(1, num_parts + 1):
        parts[p] = n
    step = 0
    for p in range(1, num_parts):
        while step < n and weights_psum[step] <= bsum:  # bisect would also do; n is small
            step += 1
        parts[p] = step
        if step == n:
            return parts, True
        bsum = weights_psum[step - 1] + bottleneck if step > 0 else bottleneck
    return parts, (total - (weights_psum[parts[num_parts - 1] - 1] if parts[num_parts - 1] > 0 else 0)) <= bottleneck


def partition_balanced(weights: Sequence[float], num_parts: int) -> List[int]:
    """Contiguous partition minimising the heaviest part (binary search on the bottleneck + greedy probe)."""
    n = len(weights)
    if n <= num_parts:
        return partition_uniform(n, num_parts)
    ps = prefix_sum_inc(weights)
    lo, hi = max(weights), ps[-1]
    best = None
    for _ in range(64):
        mid = (lo + hi) / 2
        parts, ok = _lprobe(ps, num_parts, mid)
        if ok:
            best, hi = parts, mid
        else:
            lo = mid
        if hi - lo <= max(1e-9, 1e-6 * hi):
            break
    if best is None:
        best, _ = _lprobe(ps, num_parts, hi)
    return best


def get_only_unique_item(items):
    s = set(items)
    if len(s) != 1:
        raise RuntimeError(f"expected there to be only one unique element in {items}")
    return next(iter(s))


def align_dense_tensors(tensor_list, alignment: int):
    n = sum(t.numel() for t in tensor_list)
    rem = n % alignment
    if rem:
        pad = torch.zeros(alignment - rem, device=tensor_list[0].device, dtype=tensor_list[0].dtype)
        return list(tensor_list) + [pad]
    return list(tensor_list)


def see_memory_usage(message: str, force: bool = False):
    if not force:
        return
    if dist.is_initialized() and dist.get_rank() != 0:
        return
    import psutil
    vm = psutil.virtual_memory()
    if torch.cuda.is_available():
        gb = 1024 ** 3
        logger.info(f"{message} | MA {torch.cuda.memory_allocated() / gb:.2f} GB Max_MA "
                    f"{torch.cuda.max_memory_allocated() / gb:.2f} GB CA {torch.cuda.memory_reserved() / gb:.2f} GB")
        torch.cuda.reset_peak_memory_stats()
    logger.info(f"{message} | CPU Virtual Memory: used = {vm.used / 1024 ** 3:.2f} GB, percent = {vm.percent}%")


def call_to_str(base: str, *args, **kwargs) -> str:
    parts = [repr(a) for a in args] + [f"{k}={v!r}" for k, v in kwargs.items()]
    return f"{base}({', '.join(parts)})"


def get_numel(tensors) -> int:
    return sum(t.numel() for t in tensors)


def required_torch_version(min_version=None, max_version=None) -> bool:
    from packaging import version as pv
    v = pv.parse(torch.__version__.split("+")[0])
    if min_version is not None and v < pv.parse(str(min_version)):
        return False
    if max_version is not None and v > pv.parse(str(max_version)):
        return False
    return True


# ---- the rest of the reference's helper surface lives in utils_extra.py ----------------------------------------------------
from .utils_extra import (DummyOptim, PartitionedTensor, TLinear, all_gather_dp_groups, all_gather_into_tensor_dp_groups,  # noqa: E402,F401
                          compare_tensors_in_structures, copy_to_device, empty_cache, get_flattened_grad_norm, get_grad_zeros,
                          get_inactive_params, get_ma_status, get_norm_with_moe_layers, get_norm_with_moe_layers_fast,
                          get_weight_norm, graph_process, mask_nan_or_inf_with_val_inplace, memory_status, move_to_device,
                          noop_context)
from .zero.offload_states import offload_adam_states, reload_adam_states  # noqa: E402,F401
