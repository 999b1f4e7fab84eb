"""``DSClipEncoder``: the text encoder of a diffusion pipeline behind CUDA graphs — one per call slot, because the
pipeline encodes the prompt and the negative prompt alternately
(``deepspeed/model_implementations/transformers/clip_encoder.py:11-76``)."""
from __future__ import annotations

import torch

from ..features.cuda_graph import GraphedCallable


class DSClipEncoder(torch.nn.Module):

    def __init__(self, enc, enable_cuda_graph: bool = False):
        super().__init__()
        if hasattr(enc, "text_model") and hasattr(enc.text_model, "_build_causal_attention_mask"):
            enc.text_model._build_causal_attention_mask = self._build_causal_attention_mask
        self.enc = enc
        self.device = getattr(enc, "device", None)
        self.dtype = getattr(enc, "dtype", None)
        self.config = getattr(enc, "config", None)
        self.enable_cuda_graph = enable_cuda_graph
        self._graphed = [GraphedCallable(lambda *a, **k: self.enc(*a, **k), enabled=enable_cuda_graph) for _ in range(2)]
        self.iter = 0

    @staticmethod
    def _build_causal_attention_mask(bsz, seq_len, dtype):
        mask = torch.full((bsz, seq_len, seq_len), torch.finfo(dtype).min, dtype=dtype,
                          device="cuda" if torch.cuda.is_available() else "cpu")
        return mask.triu_(1).unsqueeze(1)

    @property
    def cuda_graph_created(self):
        return [g.cuda_graph_created for g in self._graphed]

    def forward(self, *inputs, **kwargs):
        out = self._graphed[self.iter](*inputs, **kwargs)
        self.iter = (self.iter + 1) % 2
        return out
