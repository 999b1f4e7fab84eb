"""Cached views into persistent buffers (``deepspeed/inference/v2/allocator.py:14-28``): steady-state forwards
slice pre-sized buffers instead of allocating."""
from functools import reduce
from typing import Dict, Iterable, Tuple

import torch

_views: Dict[Tuple[int, Tuple[int, ...]], torch.Tensor] = {}


def empty_from(tensor: torch.Tensor, shape: Iterable[int]) -> torch.Tensor:
    shape = tuple(int(s) for s in shape)
    key = (tensor.data_ptr(), shape)
    v = _views.get(key)
    if v is None or v.dtype != tensor.dtype:
        n = reduce(lambda a, b: a * b, shape, 1)
        if n > tensor.numel():
            raise ValueError(f"view of shape {shape} does not fit a buffer of {tensor.numel()} elements")
        v = tensor.flatten()[:n].view(shape)
        _views[key] = v
    return v


on_device = empty_from
