"""``CupyBackend``: bool tensor <-> packed sign bytes. The reference round-trips through CuPy over DLPack
(``runtime/compression/cupy.py:11-27``); torch has the bit ops needed, so the "cupy tensors" here are torch tensors
and the packing runs on whatever device they live on (big-endian bit order, as ``cupy.packbits``)."""
from __future__ import annotations

from typing import List

import torch

_WEIGHTS = (128, 64, 32, 16, 8, 4, 2, 1)


class CupyBackend:

    def torch2cupy(self, tensor: torch.Tensor) -> torch.Tensor:
        return tensor

    def cupy2torch(self, cupy_tensor: torch.Tensor) -> torch.Tensor:
        return cupy_tensor

    @staticmethod
    def packbits(bool_tensor: torch.Tensor) -> torch.Tensor:
        flat = bool_tensor.reshape(-1).to(torch.uint8)
        pad = (-flat.numel()) % 8
        if pad:
            flat = torch.cat([flat, flat.new_zeros(pad)])
        w = torch.tensor(_WEIGHTS, dtype=torch.uint8, device=flat.device)
        return (flat.view(-1, 8) * w).sum(dim=1, dtype=torch.int32).to(torch.uint8)

    @staticmethod
    def unpackbits(packed: torch.Tensor) -> torch.Tensor:
        shifts = torch.arange(7, -1, -1, device=packed.device, dtype=torch.uint8)
        return ((packed.reshape(-1, 1) >> shifts) & 1).reshape(-1)

    def compress_by_chunk(self, cupy_bool_tensor: torch.Tensor, num_chunks: int) -> List[torch.Tensor]:
        packed = self.packbits(cupy_bool_tensor)
        assert packed.numel() % num_chunks == 0, "packed sign bytes must split evenly over the chunks"
        return list(packed.chunk(num_chunks))
