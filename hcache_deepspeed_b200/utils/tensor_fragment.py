"""Access the full-precision view of a parameter whose master copy / gradient / optimizer state is partitioned by
ZeRO: ``safe_get_full_fp32_param``, ``safe_get_full_grad``, ``safe_get_full_optimizer_state`` and setters.
Reference: ``deepspeed/utils/tensor_fragment.py`` (safe_get_full_fp32_param :118, safe_set_full_fp32_param :138,
safe_get_full_optimizer_state :158, safe_get_full_grad :203) — there via per-parameter fragment mappings; here by slicing
the flat partitions the optimizers keep."""
from __future__ import annotations

from typing import Optional

import torch

from .. import comm as dist


def _find(engine_or_opt, param):
    opt = getattr(engine_or_opt, "optimizer", engine_or_opt)
    from ..runtime.zero.stage_1_and_2 import DeepSpeedZeroOptimizer
    if isinstance(opt, DeepSpeedZeroOptimizer):
        for gi, params in enumerate(opt.bit16_groups):
            off = 0
            for p in params:
                if p is param:
                    return opt, gi, off
                off += p.numel()
    return opt, None, None


def _gather_flat(opt, gi: int, local: torch.Tensor) -> torch.Tensor:
    if opt.partition_count == 1:
        return local
    dev = opt.bit16_groups_flat[gi].device
    full = torch.empty(local.numel() * opt.partition_count, dtype=local.dtype, device=dev)
    dist.all_gather_into_tensor(full, local.to(dev).contiguous(), group=opt.dp_process_group)
    return full


def safe_get_full_fp32_param(param, optimizer=None) -> Optional[torch.Tensor]:
    """Collective: every rank gets the full fp32 master value of ``param``."""
    opt = optimizer if optimizer is not None else getattr(param, "_ds_optimizer", None)
    if opt is None:
        return param.detach().float()
    opt, gi, off = _find(opt, param)
    if gi is None:
        return param.detach().float()
    full = _gather_flat(opt, gi, opt.single_partition_of_fp32_groups[gi].detach())
    return full[off:off + param.numel()].view(param.shape).clone()


def safe_set_full_fp32_param(param, value: torch.Tensor, optimizer=None):
    opt = optimizer if optimizer is not None else getattr(param, "_ds_optimizer", None)
    opt, gi, off = _find(opt, param) if opt is not None else (None, None, None)
    if gi is None:
        param.data.copy_(value.to(param.dtype))
        return
    psize = opt.bit16_groups_flat[gi].numel() // opt.partition_count
    lo = opt.rank * psize
    s, e = max(off, lo), min(off + param.numel(), lo + psize)
    if s < e:
        opt.single_partition_of_fp32_groups[gi].data[s - lo:e - lo].copy_(value.reshape(-1)[s - off:e - off])
    param.data.copy_(value.to(param.dtype))


def safe_get_full_optimizer_state(param, optim_state_key: str, optimizer=None) -> Optional[torch.Tensor]:
    opt = optimizer if optimizer is not None else getattr(param, "_ds_optimizer", None)
    opt, gi, off = _find(opt, param) if opt is not None else (None, None, None)
    if gi is None:
        return None
    master = opt.single_partition_of_fp32_groups[gi]
    st = opt.optimizer.state.get(master, {})
    if optim_state_key not in st:
        return None
    full = _gather_flat(opt, gi, st[optim_state_key].detach())
    return full[off:off + param.numel()].view(param.shape).clone()


def safe_get_full_grad(param, optimizer=None) -> Optional[torch.Tensor]:
    """Before ``step()``: the averaged gradient of ``param`` (gathered from the partitions under ZeRO-2)."""
    opt = optimizer if optimizer is not None else getattr(param, "_ds_optimizer", None)
    opt, gi, off = _find(opt, param) if opt is not None else (None, None, None)
    if gi is None or opt.grad_partitions[gi] is None:
        return None if param.grad is None else param.grad.detach().float()
    full = _gather_flat(opt, gi, opt.grad_partitions[gi].detach())
    return full[off:off + param.numel()].view(param.shape).float().clone() / opt.loss_scale


safe_get_local_fp32_param = safe_get_full_fp32_param
safe_get_local_grad = safe_get_full_grad
safe_get_local_optimizer_state = safe_get_full_optimizer_state
