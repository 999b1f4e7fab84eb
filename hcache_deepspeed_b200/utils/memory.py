"""Host/device memory helpers."""
from __future__ import annotations

import torch


def pinned_empty(shape, dtype) -> torch.Tensor:
    """Page-locked host tensor when a CUDA device exists, plain host tensor otherwise (CPU-only CI)."""
    if torch.cuda.is_available():
        return torch.empty(shape, dtype=dtype, pin_memory=True)
    return torch.empty(shape, dtype=dtype)


def pinned_zeros(shape, dtype) -> torch.Tensor:
    t = pinned_empty(shape, dtype)
    t.zero_()
    return t
