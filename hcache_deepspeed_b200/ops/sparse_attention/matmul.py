"""Block-sparse matrix products over a layout (reference ``ops/sparse_attention/matmul.py``: Triton SDD / DSD / DDS
kernels). Only the layout's live blocks are multiplied: they are gathered into one batch and go through a single
batched GEMM (cuBLAS on the GPU). Sparse operands use the same packed format as the reference:
``[batch, nnz_blocks, block, block]``, blocks ordered like ``layout.nonzero()`` (head, block-row, block-col)."""
from __future__ import annotations

import torch


class MatMul:
    """``mode``: which of (C, A, B) is sparse — ``sdd``: C = A @ B sampled at the layout; ``dsd``: A sparse;
    ``dds``: B sparse. Dense operands are ``[batch, heads, rows, cols]``."""

    def __init__(self, layout: torch.Tensor, block: int, mode: str, trans_a: bool = False, trans_b: bool = False, bench: bool = False):
        if mode not in ("sdd", "dsd", "dds"):
            raise NotImplementedError("Supported modes are: sdd, dsd, dds")
        self.layout, self.block, self.mode, self.trans_a, self.trans_b = layout, block, mode, trans_a, trans_b
        nz = layout.nonzero(as_tuple=False)
        self.h, self.i, self.j = nz[:, 0], nz[:, 1], nz[:, 2]
        self.nnz = nz.shape[0]
        self.spdims = layout.shape

    def _blocks(self, x: torch.Tensor, heads: torch.Tensor, idx: torch.Tensor, dim: int) -> torch.Tensor:
        """Gather blocks of ``block`` consecutive rows (dim=2) or columns (dim=3): -> [batch, nnz, ...]."""
        b = self.block
        xb = x.unflatten(dim, (x.shape[dim] // b, b))             # rows: [B,H,nb,b,C]   cols: [B,H,R,nb,b]
        if dim == 2:
            return xb[:, heads.to(x.device), idx.to(x.device)]              # [B,nnz,b,C]
        return xb[:, heads.to(x.device), :, idx.to(x.device)].movedim(0, 1)  # [B,nnz,R,b]

    def __call__(self, a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
        blk = self.block
        if self.mode == "sdd":
            a = a.transpose(-1, -2) if self.trans_a else a
            b = b.transpose(-1, -2) if self.trans_b else b
            return torch.matmul(self._blocks(a, self.h, self.i, 2), self._blocks(b, self.h, self.j, 3))
        if self.mode == "dsd":
            a = a.transpose(-1, -2) if self.trans_a else a       # transposing every block ...
            rows, cols = (self.j, self.i) if self.trans_a else (self.i, self.j)   # ... and the block grid
            b = b.transpose(-1, -2) if self.trans_b else b
            part = torch.matmul(a, self._blocks(b, self.h, cols, 2))                   # [B,nnz,blk,N]
            out = torch.zeros(b.shape[0], self.spdims[0] * self.spdims[1 if not self.trans_a else 2], blk, b.shape[-1],
                              dtype=part.dtype, device=part.device)
            nb = self.spdims[1 if not self.trans_a else 2]
            out.index_add_(1, (self.h * nb + rows).to(part.device), part)
            return out.view(b.shape[0], self.spdims[0], nb * blk, b.shape[-1])
        # dds
        a = a.transpose(-1, -2) if self.trans_a else a
        b = b.transpose(-1, -2) if self.trans_b else b
        rows, cols = (self.j, self.i) if self.trans_b else (self.i, self.j)
        part = torch.matmul(self._blocks(a, self.h, rows, 3), b)                        # [B,nnz,M,blk]
        nb = self.spdims[2 if not self.trans_b else 1]
        out = torch.zeros(a.shape[0], self.spdims[0] * nb, a.shape[2], blk, dtype=part.dtype, device=part.device)
        out.index_add_(1, (self.h * nb + cols).to(part.device), part)
        return out.view(a.shape[0], self.spdims[0], nb, a.shape[2], blk).permute(0, 1, 3, 2, 4).reshape(
            a.shape[0], self.spdims[0], a.shape[2], nb * blk)
