"""CUDA-graph capture of a fixed-shape callable: warm up on a side stream, capture once into static buffers, then
every call is input copies + one graph launch. Used by the diffusers wrappers and the CLIP encoder.
Reference: ``deepspeed/model_implementations/features/cuda_graph.py:9-24`` (abstract mix-in) and the copies of
``_create_cuda_graph`` / ``_graph_replay`` in ``diffusers/unet.py:31-59``, ``diffusers/vae.py``, ``transformers/clip_encoder.py``.
One implementation (:class:`GraphedCallable`) replaces those copies; static inputs are CLONES, so callers may reuse
or free their tensors (the reference aliases the first call's arguments)."""
from __future__ import annotations

from abc import ABC, abstractmethod
from typing import Any, Callable, Dict, Tuple

import torch


class CUDAGraph(ABC):

    def __init__(self, enable_cuda_graph: bool = False):
        super().__init__()
        self.enable_cuda_graph = enable_cuda_graph

    @abstractmethod
    def _create_cuda_graph(self, *args, **kwargs):
        raise NotImplementedError

    @abstractmethod
    def _graph_replay(self, *args, **kwargs):
        raise NotImplementedError


def _sig(args: Tuple[Any, ...], kwargs: Dict[str, Any]):
    def one(v):
        return (tuple(v.shape), v.dtype, v.device) if torch.is_tensor(v) else ("py", repr(v))
    return tuple(one(a) for a in args), tuple((k, one(v)) for k, v in sorted(kwargs.items()))


class GraphedCallable:
    """``GraphedCallable(fn)(*args, **kwargs)``: one captured graph per distinct (shape, dtype, python-arg) signature."""

    def __init__(self, fn: Callable, enabled: bool = True, warmup: int = 3):
        self.fn, self.enabled, self.warmup = fn, enabled, warmup
        self._graphs: Dict[Any, Tuple[torch.cuda.CUDAGraph, tuple, dict, Any]] = {}

    @property
    def cuda_graph_created(self) -> bool:
        return bool(self._graphs)

    def _capture(self, key, args, kwargs):
        s_args = tuple(a.clone() if torch.is_tensor(a) else a for a in args)
        s_kwargs = {k: (v.clone() if torch.is_tensor(v) else v) for k, v in kwargs.items()}
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(self.warmup):
                self.fn(*s_args, **s_kwargs)
        torch.cuda.current_stream().wait_stream(side)
        from ...utils.cuda_graph import capture
        graph, out = capture(lambda: self.fn(*s_args, **s_kwargs))
        self._graphs[key] = (graph, s_args, s_kwargs, out)

    def __call__(self, *args, **kwargs):
        tensors = [a for a in list(args) + list(kwargs.values()) if torch.is_tensor(a)]
        if not self.enabled or not tensors or not all(t.is_cuda for t in tensors):
            return self.fn(*args, **kwargs)
        key = _sig(args, kwargs)
        if key not in self._graphs:
            self._capture(key, args, kwargs)
        graph, s_args, s_kwargs, out = self._graphs[key]
        for s, a in zip(s_args, args):
            if torch.is_tensor(a):
                s.copy_(a)
        for k, v in kwargs.items():
            if torch.is_tensor(v):
                s_kwargs[k].copy_(v)
        graph.replay()
        return out
