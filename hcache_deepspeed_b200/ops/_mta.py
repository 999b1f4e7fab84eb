"""Shared plumbing of the fused (multi-tensor) and host (ZeRO-Offload) optimizers."""
from __future__ import annotations

import ctypes
from typing import List, Optional

import torch

from . import native

_DT = {torch.bfloat16: 0, torch.float16: 1, torch.float32: 2}


def dtype_code(t: torch.Tensor) -> int:
    try:
        return _DT[t.dtype]
    except KeyError:
        raise TypeError(f"unsupported dtype {t.dtype}") from None


def ptr_array(tensors: List[torch.Tensor]):
    arr = (ctypes.c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = t.data_ptr()
    return arr


def numel_array(tensors: List[torch.Tensor]):
    arr = (ctypes.c_int64 * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = t.numel()
    return arr


def stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def multi_tensor_adam(params, grads, exp_avg, exp_avg_sq, lr, beta1, beta2, eps, weight_decay, step, adamw_mode,
                      bias_correction, grad_scale: float = 1.0) -> None:
    """One fused step over same-dtype groups (CUDA tensors, contiguous)."""
    if not params:
        return
    lib = native.cuda_lib()
    rc = lib.hds_multi_tensor_adam(len(params), ptr_array(params), ptr_array(grads), ptr_array(exp_avg),
                                   ptr_array(exp_avg_sq), numel_array(params), dtype_code(params[0]), dtype_code(grads[0]),
                                   lr, beta1, beta2, eps, weight_decay, step, int(adamw_mode), int(bias_correction),
                                   grad_scale, stream_ptr())
    native.check(rc, "hds_multi_tensor_adam")


def multi_tensor_lion(params, grads, exp_avg, lr, beta1, beta2, weight_decay, grad_scale: float = 1.0) -> None:
    if not params:
        return
    lib = native.cuda_lib()
    rc = lib.hds_multi_tensor_lion(len(params), ptr_array(params), ptr_array(grads), ptr_array(exp_avg),
                                   numel_array(params), dtype_code(params[0]), dtype_code(grads[0]), lr, beta1, beta2,
                                   weight_decay, grad_scale, stream_ptr())
    native.check(rc, "hds_multi_tensor_lion")


def group_by_dtype(params, grads, *states):
    """Split parallel lists into buckets of identical (param dtype, grad dtype)."""
    buckets = {}
    for i, (p, g) in enumerate(zip(params, grads)):
        buckets.setdefault((p.dtype, g.dtype), []).append(i)
    for idx in buckets.values():
        yield ([params[i] for i in idx], [grads[i] for i in idx], *[[s[i] for i in idx] for s in states])


def host_copy_out(copy_out: Optional[torch.Tensor]):
    if copy_out is None:
        return None, 2
    assert copy_out.device.type == "cpu" and copy_out.is_contiguous()
    return copy_out.data_ptr(), dtype_code(copy_out)
