"""``DSUNet``: a diffusers UNet frozen, switched to channels-last and replayed from a CUDA graph
(``deepspeed/model_implementations/diffusers/unet.py:12-85``)."""
from __future__ import annotations

import torch

from ..features.cuda_graph import GraphedCallable


class DSUNet(torch.nn.Module):

    def __init__(self, unet, enable_cuda_graph: bool = True):
        super().__init__()
        self.unet = unet
        # attributes the Stable Diffusion pipeline reads from its unet
        self.in_channels = getattr(unet, "in_channels", getattr(getattr(unet, "config", None), "in_channels", None))
        self.device = getattr(unet, "device", None)
        self.dtype = getattr(unet, "dtype", None)
        self.config = getattr(unet, "config", None)
        self.fwd_count = 0
        self.unet.requires_grad_(False)
        self.unet.to(memory_format=torch.channels_last)
        self.enable_cuda_graph = enable_cuda_graph
        self._graphed = GraphedCallable(self._forward, enabled=enable_cuda_graph)

    @property
    def cuda_graph_created(self) -> bool:
        return self._graphed.cuda_graph_created

    def _forward(self, sample, timestamp, encoder_hidden_states, return_dict=True, cross_attention_kwargs=None,
                 timestep_cond=None, added_cond_kwargs=None):
        if cross_attention_kwargs:
            return self.unet(sample, timestamp, encoder_hidden_states, return_dict, cross_attention_kwargs=cross_attention_kwargs)
        return self.unet(sample, timestamp, encoder_hidden_states, return_dict)

    def forward(self, *inputs, **kwargs):
        self.fwd_count += 1
        return self._graphed(*inputs, **kwargs)
