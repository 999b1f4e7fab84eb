"""Token sampling over last-token logits (the reference leaves this to HF ``generate`` / DeepSpeed-MII)."""
from __future__ import annotations

import torch


def sample_next(logits: torch.Tensor, do_sample: bool = False, temperature: float = 1.0, top_k: int = 0,
                top_p: float = 1.0) -> torch.Tensor:
    """``logits [n, vocab]`` -> int32 ``[n]``. Greedy decoding runs the library's arg-max kernel on a GPU."""
    if not do_sample or temperature == 0.0:
        if logits.device.type == "cuda" and logits.dtype in (torch.bfloat16, torch.float16, torch.float32):
            from ..ops import kernels
            return kernels.argmax(logits.contiguous())
        return logits.float().argmax(-1).to(torch.int32)
    x = logits.float() / max(temperature, 1e-6)
    if top_k and top_k > 0:
        kth = torch.topk(x, min(top_k, x.shape[-1]), dim=-1).values[..., -1:]
        x = x.masked_fill(x < kth, float("-inf"))
    if top_p < 1.0:
        sorted_x, idx = torch.sort(x, descending=True, dim=-1)
        probs = torch.softmax(sorted_x, dim=-1)
        cum = probs.cumsum(-1)
        drop = cum - probs > top_p
        sorted_x = sorted_x.masked_fill(drop, float("-inf"))
        x = torch.full_like(x, float("-inf")).scatter(-1, idx, sorted_x)
    return torch.multinomial(torch.softmax(x, dim=-1), 1).squeeze(-1).to(torch.int32)
