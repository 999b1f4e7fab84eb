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


class Allocator:
    """Views of a given shape over the front of a scratch tensor, cached per (tensor, shape) so repeated requests inside
    a forward return the same view object (ref: inference/v2/allocator.py:14)."""
    cache: dict = {}

    @staticmethod
    def empty_from(tensor: torch.Tensor, shape: Iterable[int]) -> torch.Tensor:
        shape = tuple(shape)
        per_tensor = Allocator.cache.setdefault(id(tensor), {})
        view = per_tensor.get(shape)
        if view is None or view.data_ptr() != tensor.data_ptr():
            view = per_tensor[shape] = empty_from(tensor, shape)
        return view


def on_device(method):
    """Decorator: a tensor returned by the wrapped method is moved to the current CUDA device (when there is one)."""

    def wrapped(self, *args, **kwargs):
        t = method(self, *args, **kwargs)
        if isinstance(t, torch.Tensor) and torch.cuda.is_available():
            return t.to(torch.device("cuda", torch.cuda.current_device()))
        return t

    return wrapped
