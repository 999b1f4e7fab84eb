"""``DSVAE``: a diffusers autoencoder with ``encode`` / ``decode`` / ``forward`` each replayed from its own CUDA graph
(``deepspeed/model_implementations/diffusers/vae.py:11-151``)."""
from __future__ import annotations

import torch

from ..features.cuda_graph import GraphedCallable


class DSVAE(torch.nn.Module):

    def __init__(self, vae, enable_cuda_graph: bool = True):
        super().__init__()
        self.vae = vae
        self.config = getattr(vae, "config", None)
        self.device = getattr(vae, "device", None)
        self.dtype = getattr(vae, "dtype", None)
        self.vae.requires_grad_(False)
        self.enable_cuda_graph = enable_cuda_graph
        self._decode = GraphedCallable(lambda *a, **k: self.vae.decode(*a, **k), enabled=enable_cuda_graph)
        self._encode = GraphedCallable(lambda *a, **k: self.vae.encode(*a, **k), enabled=enable_cuda_graph)
        self._fwd = GraphedCallable(lambda *a, **k: self.vae(*a, **k), enabled=enable_cuda_graph)

    def decode(self, *inputs, **kwargs):
        return self._decode(*inputs, **kwargs)

    def encode(self, *inputs, **kwargs):
        return self._encode(*inputs, **kwargs)

    def forward(self, *inputs, **kwargs):
        return self._fwd(*inputs, **kwargs)
