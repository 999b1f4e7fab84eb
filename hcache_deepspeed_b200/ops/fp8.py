"""FP8 (E4M3) weights for the ragged executor (``quantization_mode = "wf8af8"``): a weight is stored as its fp8 codes
plus one fp32 scale per output channel; the op back ends quantise activations per token and run the
``tcgen05.mma.kind::f8f6f4`` GEMM (``ops/csrc/gemm_fp8.cu``, ``gemm_2sm.cu``). Takes the place of the reference's
weight-only FP6 linear (``modules/implementations/linear/quantized_linear.py:24-205``)."""
from __future__ import annotations

from typing import Tuple

import torch

E4M3_MAX = 448.0


def quantize_rows_reference(x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """torch restatement of ``hds_quantize_fp8_rows``: (uint8 codes, fp32 per-row scales)."""
    amax = x.float().abs().amax(dim=1)
    scale = torch.where(amax > 0, amax / E4M3_MAX, torch.ones_like(amax))
    q = (x.float() / scale[:, None]).to(torch.float8_e4m3fn).view(torch.uint8)
    return q, scale


E3M2_MAX = 28.0
#: the 32 non-negative E3M2 values in code order (code = exponent << 2 | mantissa; exponent bias 3, subnormals at e = 0)
_E3M2_POS = torch.tensor([(m / 4.0) * 2.0 ** -2 if e == 0 else (1.0 + m / 4.0) * 2.0 ** (e - 3) for e in range(8) for m in range(4)])


def quantize_rows_e3m2(x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Per-row E3M2 (FP6) quantisation: ``(codes uint8 [rows, K] in 0..63, fp32 scales [rows])``, round to nearest."""
    xf = x.float()
    amax = xf.abs().amax(dim=1)
    scale = torch.where(amax > 0, amax / E3M2_MAX, torch.ones_like(amax))
    y = xf / scale[:, None]
    pos = _E3M2_POS.to(x.device)
    mid = (pos[1:] + pos[:-1]) / 2
    idx = torch.bucketize(y.abs().contiguous(), mid)            # nearest representable magnitude
    code = idx.to(torch.uint8) | ((y < 0).to(torch.uint8) << 5)
    return code, scale


def pack_6bit(code: torch.Tensor) -> torch.Tensor:
    """``[rows, K]`` 6-bit codes -> ``[rows, 3K/4]`` bytes: value i of a row occupies bits [6i, 6i+6) of the row's
    little-endian bit string (16 values = 12 bytes, the group the TMA type 16U6_ALIGN16B unpacks)."""
    rows, K = code.shape
    assert K % 4 == 0
    c = code.view(rows, K // 4, 4).to(torch.int32)
    b0 = (c[..., 0] | (c[..., 1] << 6)) & 0xFF
    b1 = ((c[..., 1] >> 2) | (c[..., 2] << 4)) & 0xFF
    b2 = ((c[..., 2] >> 4) | (c[..., 3] << 2)) & 0xFF
    return torch.stack([b0, b1, b2], dim=-1).to(torch.uint8).reshape(rows, K * 3 // 4).contiguous()


def unpack_6bit(packed: torch.Tensor) -> torch.Tensor:
    rows, nb = packed.shape
    b = packed.view(rows, nb // 3, 3).to(torch.int32)
    c0 = b[..., 0] & 0x3F
    c1 = ((b[..., 0] >> 6) | (b[..., 1] << 2)) & 0x3F
    c2 = ((b[..., 1] >> 4) | (b[..., 2] << 4)) & 0x3F
    c3 = (b[..., 2] >> 2) & 0x3F
    return torch.stack([c0, c1, c2, c3], dim=-1).to(torch.uint8).reshape(rows, nb // 3 * 4)


def e3m2_to_float(code: torch.Tensor) -> torch.Tensor:
    pos = _E3M2_POS.to(code.device)
    mag = pos[(code & 0x1F).long()]
    return torch.where((code & 0x20) != 0, -mag, mag)


class Fp8Weight:
    """``[N, K]`` weight in a tensor-core 8-bit-class float format plus per-output-channel scales (``scale``: fp32 ``[N]``):
    ``fmt = "e4m3"`` — one byte per weight (``q``: uint8 ``[N, K]``); ``fmt = "e3m2"`` — FP6, packed 6 bits per weight
    (``q``: uint8 ``[N, 3K/4]``; the TMA unpacks it on the way into shared memory). Row slices (``w[r0:r1]``) are views,
    like the bf16 tensors they replace (the K/V rows of a fused QKV weight)."""

    __slots__ = ("q", "scale", "fmt")

    def __init__(self, q: torch.Tensor, scale: torch.Tensor, fmt: str = "e4m3"):
        assert q.dtype == torch.uint8 and q.dim() == 2 and scale.dtype == torch.float32 and scale.shape == (q.shape[0], )
        assert fmt in ("e4m3", "e3m2")
        self.q, self.scale, self.fmt = q, scale, fmt

    @classmethod
    def from_tensor(cls, w: torch.Tensor, fmt: str = "e4m3") -> "Fp8Weight":
        if fmt == "e3m2":
            if w.shape[1] % 128:
                raise ValueError(f"FP6 weights need an input dimension that is a multiple of 128, got {w.shape[1]}")
            code, s = quantize_rows_e3m2(w)
            return cls(pack_6bit(code), s, "e3m2")
        if w.is_cuda:
            from . import kernels
            q, s = kernels.quantize_fp8_rows(w.contiguous())
        else:
            q, s = quantize_rows_reference(w)
        return cls(q, s)

    @property
    def fmt_code(self) -> int:
        return 1 if self.fmt == "e3m2" else 0

    @property
    def shape(self):
        if self.fmt == "e3m2":
            return torch.Size((self.q.shape[0], self.q.shape[1] * 4 // 3))
        return self.q.shape

    @property
    def device(self):
        return self.q.device

    def __getitem__(self, rows) -> "Fp8Weight":
        assert isinstance(rows, slice), "Fp8Weight supports row slices only"
        return Fp8Weight(self.q[rows], self.scale[rows], self.fmt)

    def dequantize(self, dtype=torch.float32) -> torch.Tensor:
        if self.fmt == "e3m2":
            return (e3m2_to_float(unpack_6bit(self.q)) * self.scale[:, None]).to(dtype)
        return (self.q.view(torch.float8_e4m3fn).float() * self.scale[:, None]).to(dtype)

    def nbytes(self) -> int:
        return self.q.numel() + 4 * self.scale.numel()
